"""Resident-input step time of the default workload (BERT-base, S = 384, B = 32, dropout 0.1) for A/B runs on one box:
    python tools/step_time.py                       # in-tree libbqa.so
    BQA_LIB_PATH=<other build> python tools/step_time.py
Prints ms/step of three legs of 30 steps (clocks settle over the first seconds)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
import bench
from __graft_entry__ import load_package
pkg = load_package()
from _ab import maybe_swap_library
maybe_swap_library(pkg)
dev = torch.device("cuda", 0)
init = importlib.import_module("bqa_b200.modules.init")
tm = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
B, S = 32, 384
params, mp = bench.make_params(B, S)
if 'STEP_HIDDEN_DROPOUT' in os.environ:
    mp.hidden_dropout_prob = float(os.environ['STEP_HIDDEN_DROPOUT'])
if 'STEP_ATTN_DROPOUT' in os.environ:
    mp.attention_probs_dropout_prob = float(os.environ['STEP_ATTN_DROPOUT'])
torch.manual_seed(1234)
model, tok = init.init_model(mp, device=dev)
opt = init.init_optimizer(params, model)
tr = tm.Trainer(model=model, loss=init.init_loss(params, {"label_weights": None}), collate_fun=None, optimizer=opt,
                device=dev, local_rank=-1, n_epochs=1, train_batch_size=B, batch_split=1, n_jobs=0, warmup_coef=0.0,
                max_grad_norm=1.0)
tr.set_train()
hb = bench.host_batches(4, B, S, 1234, tok, compact=True)
db = [tr._to_device(b) for b in hb]


def step(i):
    x, y = db[i % 4]
    tr.train_micro_batch(x, y, last_of_step=True)
    tr.optimizer_step()


for i in range(5):
    step(i)
legs = []
for leg in range(3):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(30):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    legs.append(e0.elapsed_time(e1) / 30)
print("library:", pkg._lib.LIB_PATH, "ms/step per leg:", " ".join(f"{x:.3f}" for x in legs))
