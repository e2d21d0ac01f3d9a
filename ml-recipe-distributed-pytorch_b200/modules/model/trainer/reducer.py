"""Data-parallel gradient exchange: bucketed all-reduce (mean) of the flat gradient buffer,
issued bucket by bucket on a side stream while the backward of earlier layers is still running.

Replaces torch.nn.parallel.DistributedDataParallel as wrapped by the reference
(modules/model/trainer/trainer.py:136-142): same arithmetic (average of the ranks' gradients,
once per optimizer step), without the autograd-hook reducer, the per-iteration unused-parameter
graph walk (`find_unused_parameters=True`) or the parameter/buffer broadcasts.  Buckets are
contiguous slices of the flat buffer in the order the backward finalises them
(heads -> layer L-1 .. 0 -> embeddings, `bqa_phase_grad_begin`).  The collective itself is
NCCL's all-reduce over NVLink/NVSwitch through torch.distributed; device-agnostic so that the
bucket logic is testable with gloo on CPU tensors.

Exchange precision.  The reference's DDP exchanges whatever dtype the gradients have: fp32 under apex O0 / O1,
16-bit under O2 / O3 (trainer.py:128-142).  `exchange_dtype=torch.bfloat16` packs a bucket to bf16 (bqa_f32_to_bf16),
all-reduces half the bytes and unpacks into the fp32 gradient buffer: 219 MB instead of 438 MB per step for
BERT-base, the compute dtype of this implementation.  fp32 stays available (`exchange_dtype=torch.float32`).
"""
import torch
import torch.distributed as dist

# Bucket size.  Every all-reduce issued beside the backward costs ~0.1-0.2 ms of step time whatever its size: its CTAs take
# SMs as soon as a GEMM of the backward lets go of them and then spin until the peers' CTAs have been scheduled too
# (every rank is in the middle of a different GEMM).  Over NVLink 5 / NVSwitch the bytes themselves are cheap (438 MB in
# ~0.7 ms at 8 GPUs), so few large buckets beat many small ones.  Measured on B200s, BERT-base, ms per step at 25 / 60 /
# 120 / 160 / 220 MB: 8 GPUs 10.50 / 10.41 / 10.24 / 10.20 / 10.01, 2 GPUs 10.05 / - / - / 10.03 / 9.93
# (profiles/r02_bucket_sweep.json).  220 MB = two buckets for BERT-base (heads + layers 11..4 | layers 3..0 + embeddings).
DEFAULT_BUCKET_MB = 220.0


def bucket_plan(phase_begins, total, min_bucket_elems):
    """phase_begins[k] = first flat element that is final after backward phase k (decreasing, last
    one 0).  Returns [(last_phase, begin, end)] covering [0, total) from the top down, merging
    consecutive phases until a bucket holds at least `min_bucket_elems` elements."""
    plan, end = [], total
    for k, begin in enumerate(phase_begins):
        last = k == len(phase_begins) - 1
        if end - begin >= min_bucket_elems or last:
            if end > begin:
                plan.append((k, begin, end))
            end = begin
    assert not plan or plan[-1][1] == 0, "phases must cover the buffer down to element 0"
    return plan


class GradReducer:
    def __init__(self, flat_grads, phase_begins, *, group=None, min_bucket_mb=DEFAULT_BUCKET_MB, world=None,
                 exchange_dtype=torch.float32):
        self.flat = flat_grads
        self.group = group
        # world=1: bucket schedule and side stream only, no exchange (a single-rank Trainer, also inside a process
        # that has a process group)
        if world is None:
            world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.world = world
        self.plan = bucket_plan(phase_begins, flat_grads.numel(), int(min_bucket_mb * (1 << 20) / 4))
        self.cuda = flat_grads.is_cuda
        self.stream = torch.cuda.Stream(device=flat_grads.device) if self.cuda else None
        self._works = []
        self.exchange_dtype = exchange_dtype
        self._pack = None  # 16-bit staging of the largest bucket (allocated on first use)

    def phases(self):
        """[(first_phase, last_phase, bucket_index)] to run the backward bucket by bucket."""
        out, first = [], 0
        for i, (last, _, _) in enumerate(self.plan):
            out.append((first, last, i))
            first = last + 1
        return out

    def reduce_bucket(self, i, then=None):
        """All-reduce (mean) bucket i; on CUDA it is enqueued on the side stream behind everything
        issued so far on the current stream.  `then(begin, end)` (optional) is called with the side
        stream current right after the bucket's exchange has been enqueued: work on the final gradient
        of the bucket that can overlap the rest of the backward (the optimizer's sum of squares).
        With one rank there is no exchange, only `then`."""
        _, b, e = self.plan[i]
        if self.world == 1 and then is None:
            return
        t = self.flat[b:e]
        if self.cuda:
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(self.flat.device))
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ready)
                if self.world > 1:
                    if self.exchange_dtype == torch.bfloat16:
                        self._all_reduce_bf16(t)
                    else:
                        dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.group)
                if then is not None:
                    then(b, e)
            self._pending = True
        else:
            if self.world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
                t.div_(self.world)
            if then is not None:
                then(b, e)

    def _all_reduce_bf16(self, t):
        """fp32 bucket -> bf16 (library kernel) -> all-reduce (mean) of half the bytes -> back into the fp32 bucket.
        Runs on the side stream; consecutive buckets share one staging buffer (stream-ordered)."""
        if not t.is_cuda:
            raise RuntimeError('the bf16 gradient exchange packs with a CUDA kernel (bqa_f32_to_bf16): there is no CPU path')
        from .... import _lib as L
        if self._pack is None:
            n_max = max(e - b for _, b, e in self.plan)
            self._pack = torch.empty(n_max, dtype=torch.bfloat16, device=t.device)
        pk = self._pack[:t.numel()]
        L.check(L.lib.bqa_f32_to_bf16(t.data_ptr(), pk.data_ptr(), t.numel(), L.current_stream_ptr()), "bqa_f32_to_bf16")
        dist.all_reduce(pk, op=dist.ReduceOp.AVG, group=self.group)
        t.copy_(pk)

    def wait(self):
        """Make the current stream wait for every bucket enqueued so far."""
        if self.cuda and getattr(self, "_pending", False):
            torch.cuda.current_stream(self.flat.device).wait_stream(self.stream)
            self._pending = False
