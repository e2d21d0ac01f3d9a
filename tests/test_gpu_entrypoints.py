"""The reference's entry points executed on the GPU: `modules/train.py` and `modules/validate.py` as
scripts with the shipped config files (reference modules/train.py:125-148, validate.py:29-54),
`Trainer.train()` / `Trainer.test()` with a DataLoader and the three test callbacks
(trainer.py:266-353, callback.py:30-108), and the NCCL gradient exchange against a single-rank
step (trainer.py:136-142) when the box has two GPUs."""
import importlib
import json
import os
import subprocess
import sys
import types

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ml-recipe-distributed-pytorch_b200")
DEV = "cuda"
TINY = dict(num_layers=2, hidden=128, heads=2, intermediate=512, vocab=30522, max_pos=512, type_vocab=2)


def _run(cmd, timeout=900):
    env = dict(os.environ, PYTHONUNBUFFERED="1")
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    return r.returncode, r.stdout + "\n" + r.stderr


def test_train_script_runs_test_bert_cfg(tmp_path):
    """BASELINE configs[0] on the GPU: config/test_bert.cfg (debug mode, DummyDataset, BERT-base) with the
    sequence / batch overrides of SURVEY.md §8.  Two epochs of one optimizer step each, each followed by
    save_last / save_each (no-ops in debug mode) and Trainer.test with MAP / accuracy / save-best callbacks."""
    code, out = _run([sys.executable, os.path.join(PKG, "modules", "train.py"), "-c", "config/test_bert.cfg",
                      "--max_seq_len", "384", "--train_batch_size", "8", "--batch_split", "1", "--n_jobs", "2",
                      "--dump_dir", str(tmp_path), "--seed", "1"])
    assert code == 0, out[-4000:]
    logs = list((tmp_path / "test").glob("*.log"))
    text = out + "".join(p.read_text() for p in logs)
    assert "Test metrics after epoch 2" in text, text[-4000:]
    assert "map" in text and "s_acc" in text
    assert (tmp_path / "test" / "trainer.cfg").exists() and (tmp_path / "test" / "model.cfg").exists()


def test_validate_script_runs_validate_cfg(tmp_path):
    """config/validate.cfg (S = 512, doc stride 128) over the synthetic chunk dataset, random-init weights
    (`--checkpoint None`: the reference's checkpoint file is not shipped)."""
    code, out = _run([sys.executable, os.path.join(PKG, "modules", "validate.py"), "-c", "config/validate.cfg",
                      "--checkpoint", "None", "--limit", "6", "--n_jobs", "0"])
    assert code == 0, out[-4000:]


def test_trainer_train_and_test_with_dataloader_and_callbacks(pkg, tmp_path):
    """Trainer.train() over a DataLoader (collate_fun, pinned batches, _to_device), grad accumulation, warm-up
    schedule, after-epoch functions, then Trainer.test() with the three callbacks; checkpoints are written
    (not debug mode) and the device-side decode used by AccuracyCallback agrees with torch.argmax."""
    I = importlib.import_module("bqa_b200.modules.init")
    T = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
    C = importlib.import_module("bqa_b200.modules.model.trainer.callback")
    D = importlib.import_module("bqa_b200.modules.model.dataset")
    M = importlib.import_module("bqa_b200.modules.model.model")
    mp = types.SimpleNamespace(model="bert-base-uncased", model_shape=TINY, hidden_dropout_prob=0.1,
                               attention_probs_dropout_prob=0.1, layer_norm_eps=1e-12)
    torch.manual_seed(3)
    model = M.BertForQuestionAnswering(mp, 5).to(DEV)
    tok = I.DummyTokenizer()
    hp = types.SimpleNamespace(lr=1e-3, weight_decay=1e-2, optimizer="adam", finetune=False, apex_level=None,
                               loss="smooth", smooth_alpha=0.01, focal_alpha=1, focal_gamma=2, w_start=1, w_end=1,
                               w_start_reg=1, w_end_reg=1, w_cls=1, best_metric="map", best_order=">",
                               dump_dir=tmp_path, experiment_name="exp")
    train_ds = D.DummyDataset(tok, max_seq_len=128, max_question_len=16, dataset_len=32)
    test_ds = D.DummyDataset(tok, max_seq_len=128, max_question_len=16, dataset_len=12)
    tr = T.Trainer(model=model, loss=I.init_loss(hp, {"label_weights": None}), collate_fun=I.init_collate_fun(tok),
                   optimizer=I.init_optimizer(hp, model), train_dataset=train_ds, test_dataset=test_ds,
                   writer_dir=None, device=torch.device(DEV), n_epochs=2, train_batch_size=8, test_batch_size=4,
                   batch_split=2, n_jobs=0, warmup_coef=0.5, max_grad_norm=1.0)
    w0 = model.engine.params.clone()
    seen = []
    cbs = [C.MAPCallback(list(D.RawPreprocessor.labels2id.keys())), C.AccuracyCallback(), C.SaveBestCallback(hp)]

    def save_last(epoch_i):
        tr.save_state_dict(tmp_path / "exp" / "last.ch")

    def test_fun(epoch_i):
        tr.test(epoch_i, callbacks=cbs)
        seen.append(dict(tr.last_test_metrics))

    tr.train(after_epoch_funcs=[save_last, test_fun])
    torch.cuda.synchronize()
    # 32 items / (8 / 2 per micro-batch) = 8 micro-batches = 4 optimizer steps per epoch
    assert tr.global_step == 8
    assert all(tr.optimizer.state[p]["step"] == 8 for p in model.parameters())
    assert not torch.equal(model.engine.params, w0)
    assert torch.isfinite(model.engine.params).all()
    for k in ("start_class", "end_class", "start_reg", "end_reg", "cls", "loss", "lr"):
        assert k in tr.last_train_meters
    assert len(seen) == 2
    for m in seen:
        for k in ("loss", "cls", "s_acc", "e_acc", "c_acc", "map"):
            assert k in m, (k, m)
        assert 0.0 <= m["s_acc"] <= 1.0 and 0.0 <= m["c_acc"] <= 1.0
    # the dummy labels are constant (start 0, end S-1, class 0): two epochs at lr 1e-3 must reduce the loss
    assert seen[1]["loss"] < seen[0]["loss"] + 1e-3
    assert (tmp_path / "exp" / "last.ch").exists() and (tmp_path / "exp" / "best.ch").exists()
    ck = torch.load(tmp_path / "exp" / "last.ch", map_location="cpu")
    assert ck["global_step"] == 8 and ck["scheduler"] is not None

    # device decode of AccuracyCallback == torch.argmax on the same predictions (first index on ties)
    model.eval()
    inputs, labels = next(iter(tr.test_dataloader))
    inputs = tr._to_device(inputs)
    with torch.no_grad():
        preds = model(**inputs)
    dec = C.device_decode(preds)
    assert torch.equal(dec["start_ids"], preds["start_class"].argmax(-1))
    assert torch.equal(dec["end_ids"], preds["end_class"].argmax(-1))
    assert torch.equal(dec["cls_ids"], preds["cls"].argmax(-1))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_two_rank_nccl_step_equals_single_rank_step(tmp_path):
    out = tmp_path / "ddp.json"
    code, log = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                      "--master-addr", "127.0.0.1", "--master-port", "29611",
                      os.path.join(ROOT, "tests", "dist_parity_worker.py"), "--shape", "base", "--out", str(out)])
    assert code == 0, log[-4000:]
    rep = json.loads(out.read_text())
    assert rep["ok"] and rep["world"] == 2
