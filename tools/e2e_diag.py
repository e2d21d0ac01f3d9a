"""Where the e2e loop loses time against the resident loop: host time per step, variants."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from __graft_entry__ import load_package
load_package()
dev = torch.device("cuda", 0)
init = importlib.import_module("bqa_b200.modules.init")
tm = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
params, mp = bench.make_params(32, 384)
torch.manual_seed(1234)
model, tok = init.init_model(mp, device=dev)
opt = init.init_optimizer(params, model)
loss = init.init_loss(params, {"label_weights": None})
tr = tm.Trainer(model=model, loss=loss, collate_fun=None, optimizer=opt, device=dev, local_rank=-1, n_epochs=1,
                train_batch_size=32, batch_split=1, n_jobs=0, warmup_coef=0.0, max_grad_norm=params.max_grad_norm)
tr.set_train()
hb = bench.host_batches(4, 32, 384, 1234, tok, compact=True)
db = [tr._to_device(b) for b in hb]
rb = tm._LossReadback(dev)


def run(name, fn, n=30):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    host = 0.0
    e0.record()
    for i in range(n):
        t = time.perf_counter()
        fn(i)
        host += time.perf_counter() - t
    e1.record()
    torch.cuda.synchronize()
    print(f"{name:34s} gpu {e0.elapsed_time(e1) / n:7.3f} ms/step   host {host / n * 1e3:7.3f} ms/step")


def resident(i):
    x, y = db[i % 4]
    tr.train_micro_batch(x, y, last_of_step=True)
    tr.optimizer_step()


def resident_readback(i):
    x, y = db[i % 4]
    rb.push(tr.train_micro_batch(x, y, last_of_step=True))
    rb.pop_previous()
    tr.optimizer_step()


def h2d_only(i):
    x, y = tr._to_device(hb[i % 4])
    tr.train_micro_batch(x, y, last_of_step=True)
    tr.optimizer_step()


def full(i):
    x, y = tr._to_device(hb[i % 4])
    rb.push(tr.train_micro_batch(x, y, last_of_step=True))
    rb.pop_previous()
    tr.optimizer_step()


run("resident", resident)
run("resident + deferred readback", resident_readback)
run("h2d every step", h2d_only)
run("h2d + deferred readback", full)
run("resident (again)", resident)

# ---- host time of the individual library calls
import ctypes
L = importlib.import_module("bqa_b200._lib")
orig = {}
acc = {}
real = L.lib


def wrap(name):
    fn = getattr(real, name)

    def w(*a):
        t = time.perf_counter()
        r = fn(*a)
        acc[name] = acc.get(name, 0.0) + time.perf_counter() - t
        acc[name + "#"] = acc.get(name + "#", 0) + 1
        return r
    w.restype, w.argtypes = fn.restype, fn.argtypes
    return w


class Proxy:
    def __init__(self, lib):
        self._lib = lib
        self._cache = {}

    def __getattr__(self, n):
        if n not in self._cache:
            self._cache[n] = wrap(n)
        return self._cache[n]


L.lib = Proxy(real)
for m in list(sys.modules.values()):
    if m is not None and getattr(m, "L", None) is L:
        pass  # modules reference L.lib dynamically
torch.cuda.synchronize()
n = 20
t0 = time.perf_counter()
for i in range(n):
    resident(i)
tot = time.perf_counter() - t0
torch.cuda.synchronize()
print(f"host total {tot / n * 1e3:.3f} ms/step")
for k in sorted(acc):
    if not k.endswith("#"):
        print(f"  {k:28s} {acc[k] / n * 1e3:7.3f} ms/step  calls/step {acc[k + '#'] / n:.1f}")

# ---- pure host cost: one step launched into an empty queue
for k in list(acc):
    del acc[k]
ts = []
for i in range(10):
    torch.cuda.synchronize()
    t = time.perf_counter()
    resident(i)
    ts.append(time.perf_counter() - t)
torch.cuda.synchronize()
print(f"host cost of one step into an empty queue: median {sorted(ts)[5] * 1e3:.3f} ms")
for k in sorted(acc):
    if not k.endswith("#"):
        print(f"  {k:28s} {acc[k] / 10 * 1e3:7.3f} ms/step")
