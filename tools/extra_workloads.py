"""BASELINE.json configs[3] and [4] on one GPU (they are parity / coverage cases, not the bench line):
  C4  BERT-large QA fine-tune, seq 512, batch 16   -> optimizer steps per second
  C5  BERT-base span decoding over Natural-Questions-shaped chunks, seq <= 512 -> chunks per second
Writes profiles/r01_extra_workloads.json."""
import importlib, json, os, sys, time, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
from __graft_entry__ import load_package
load_package()
dev = torch.device("cuda", 0)
init = importlib.import_module("bqa_b200.modules.init")
tm = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
ds = importlib.import_module("bqa_b200.modules.model.dataset")
out = {}


def timed(fn, n, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


# ---- C4
B, S = 16, 512
params, mp = bench.make_params(B, S)
mp.model = "bert-large-uncased"
torch.manual_seed(1234)
model, tok = init.init_model(mp, device=dev)
opt = init.init_optimizer(params, model)
loss = init.init_loss(params, {"label_weights": None})
tr = tm.Trainer(model=model, loss=loss, collate_fun=None, optimizer=opt, device=dev, local_rank=-1, n_epochs=1,
                train_batch_size=B, batch_split=1, n_jobs=0, warmup_coef=0.0, max_grad_norm=params.max_grad_norm)
tr.set_train()
hb = bench.host_batches(2, B, S, 1234, tok, compact=True)
db = [tr._to_device(b) for b in hb]


def step(i):
    x, y = db[i % 2]
    tr.train_micro_batch(x, y, last_of_step=True)
    tr.optimizer_step()


ms = timed(step, 10)
flop = 1005.0e9 * B  # 1005.0 GFLOP per sample (BERT-large, S = 512, training)
out["C4_bert_large_train_seq512_batch16"] = {"ms_per_step": round(ms, 3), "samples_per_s": round(B / ms * 1e3, 1),
                                             "model_tflops": round(flop / ms / 1e9, 1)}
print(out)
del tr, opt, model, db
torch.cuda.empty_cache()

# ---- C5: chunks of documents, ragged lengths, key mask, decode on the device
val = importlib.import_module("bqa_b200.modules.validate")
pred_mod = importlib.import_module("bqa_b200.modules.model.inference.predictor")
params, mp = bench.make_params(16, 512)
torch.manual_seed(1234)
model, tok = init.init_model(mp, device=dev)
model.eval()
data = val.SyntheticChunkDataset(64, max_seq_len=512, doc_stride=128)
chunks = [c for doc in data.docs for c in doc]
collate = init.init_collate_fun(tok, return_items=True)
pred = pred_mod.Predictor(model, dev, batch_size=64, n_jobs=0, collate_fun=collate)
batches = [collate(chunks[i:i + 64]) for i in range(0, len(chunks) - 63, 64)][:8]


def infer(i):
    inputs, labels, items = batches[i % len(batches)]
    with torch.no_grad():
        pred.predict_batch(inputs)   # forward + device-side decode + D2H of the result vectors


try:
    ms = timed(infer, 16)
    out["C5_bert_base_decode_seq512_batch64"] = {"ms_per_batch": round(ms, 3), "chunks_per_s": round(64 / ms * 1e3, 1),
                                                 "n_chunks": len(chunks)}
except Exception as e:  # report, do not hide
    out["C5_bert_base_decode_seq512_batch64"] = {"error": repr(e)}
print(out)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "extra_workloads.json"), "w"), indent=1)
