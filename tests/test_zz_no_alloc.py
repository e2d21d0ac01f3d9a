"""bqa_grad_sumsq_noalloc: the sum of squares for callers that cannot hand over (or create) a scratch
buffer — e.g. a server that gave all device memory to its own pool at start-up, or a plugin with no
access to its host's allocator.  The partial sums go to a fixed array inside the library instead.

Checked through the C ABI against bqa_grad_sumsq (caller-provided scratch) on two calls in a row: a
large buffer first, then a small one that uses fewer partials, so that leftovers of the first call in
the shared array would show up if they were ever read.  Collected last (file name) so that a run that
stops at the first failure has already executed every other test."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(L, g):
    n = g.numel()
    st = L.current_stream_ptr()
    ref = torch.zeros(1, dtype=torch.float64, device=g.device)
    scratch = torch.empty(L.lib.bqa_sumsq_scratch_floats(), dtype=torch.float32, device=g.device)
    L.check(L.lib.bqa_grad_sumsq(g.data_ptr(), n, ref.data_ptr(), scratch.data_ptr(), st), "bqa_grad_sumsq")
    got = torch.zeros(1, dtype=torch.float64, device=g.device)
    L.check(L.lib.bqa_grad_sumsq_noalloc(g.data_ptr(), n, got.data_ptr(), st), "bqa_grad_sumsq_noalloc")
    torch.cuda.synchronize()
    return got.item(), ref.item()


def test_noalloc_sumsq_matches_scratch_version_twice(pkg):
    L = pkg._lib
    torch.manual_seed(0)
    big = torch.randn(109_489_664, device="cuda")          # the flat gradient buffer of BERT-base
    small = torch.randn(4_099, device="cuda") * 3.0        # fewer blocks than the first call, ragged tail
    for g in (big, small):
        got, ref = _both(L, g)
        exact = float((g.double() ** 2).sum().item())
        assert abs(got - ref) <= 1e-6 * ref                 # same kernels, same partition: agree to rounding
        assert abs(got - exact) <= 1e-5 * exact
    # accumulation semantics (+=) are the same as the scratch version's
    acc = torch.full((1,), 5.0, dtype=torch.float64, device="cuda")
    L.check(L.lib.bqa_grad_sumsq_noalloc(small.data_ptr(), small.numel(), acc.data_ptr(), L.current_stream_ptr()),
            "bqa_grad_sumsq_noalloc")
    torch.cuda.synchronize()
    assert abs(acc.item() - 5.0 - float((small.double() ** 2).sum())) <= 1e-5 * acc.item()


def test_noalloc_sumsq_rejects_oversize(pkg):
    L = pkg._lib
    g = torch.zeros(16, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    rc = L.lib.bqa_grad_sumsq_noalloc(g.data_ptr(), (1 << 31) + 4, out.data_ptr(), L.current_stream_ptr())
    assert rc != 0 and b"limit" in L.lib.bqa_last_error()
