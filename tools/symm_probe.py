"""Probe: is multicast (NVLS) symmetric memory available, and how fast is an in-switch all-reduce next to NCCL?"""
import os, time
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as sm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
gname = dist.group.WORLD.group_name
try:
    sm.enable_symm_mem_for_group(gname)
except Exception as e:
    print("enable:", repr(e))
n = 110_000_000
t = sm.empty(n, dtype=torch.float32, device=dev)
hdl = sm.rendezvous(t, gname)
if rank == 0:
    print("multicast support:", sm._SymmetricMemory.has_multicast_support(DeviceType := torch._C._autograd.DeviceType.CUDA, lr) if hasattr(sm._SymmetricMemory, "has_multicast_support") else None)
    print("multicast_ptr:", hex(hdl.multicast_ptr), "world", hdl.world_size, "signal pad", hdl.signal_pad_size)
t.fill_(rank + 1.0)
torch.cuda.synchronize(); dist.barrier()


def bench(name, fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    if rank == 0:
        print(f"{name}: {e0.elapsed_time(e1) / iters * 1e3:.1f} us")


for mb in (28, 96):
    k = mb * (1 << 20) // 4
    sl = t[:k]
    ref = torch.empty(k, device=dev)
    bench(f"nccl all_reduce {mb} MB", lambda: dist.all_reduce(ref))
    try:
        bench(f"symm_mem multimem_all_reduce_ {mb} MB", lambda: torch.ops.symm_mem.multimem_all_reduce_(sl, "sum", gname))
    except Exception as e:
        if rank == 0:
            print("multimem_all_reduce_ failed:", repr(e)[:300])
    try:
        bench(f"symm_mem two_shot_all_reduce_ {mb} MB", lambda: torch.ops.symm_mem.two_shot_all_reduce_(sl, "sum", gname))
    except Exception as e:
        if rank == 0:
            print("two_shot failed:", repr(e)[:300])
dist.destroy_process_group()
