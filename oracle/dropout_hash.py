"""Restatement (torch integer ops, any device) of the library's stateless dropout decision.

TEST INFRASTRUCTURE ONLY.  The reference uses torch's Philox stream for nn.Dropout
(site-packages/transformers/models/bert/modeling_bert.py:112,138,297,355 and reference
modules/model/model/model.py:33); that stream cannot be reproduced outside torch, so the
library defines its own counter-based rule (csrc/bqa_common.cuh: mix32/drop_keep,
csrc/bqa_host.cuh: make_drop_key) and this file restates it so that tests can compare
dropout-on results exactly instead of statistically.
"""
import torch

_M32 = 0xFFFFFFFF
_M64 = 0xFFFFFFFFFFFFFFFF


def splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & _M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M64
    return x ^ (x >> 31)


def drop_key(seed, site):
    h = splitmix64((seed ^ ((0xD1B54A32D192ED03 * (site + 1)) & _M64)) & _M64)
    return h & _M32, (h >> 32) & _M32


def _mix32(x):
    # x: int64 tensor holding uint32 values
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _M32
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _M32
    x = x ^ (x >> 16)
    return x


def keep_mask(seed, site, p, numel, device="cpu", start=0):
    """Boolean keep decisions for element indices [start, start+numel): one 32-bit hash per PAIR
    of consecutive indices, low 16 bits for the even element, high 16 bits for the odd one."""
    if p <= 0:
        return torch.ones(numel, dtype=torch.bool, device=device)
    k0, k1 = drop_key(seed, site)
    thresh = min(int(round(p * 65536.0)), 65535)
    idx = torch.arange(start, start + numel, dtype=torch.int64, device=device) & _M32
    h = _mix32(((idx >> 1) * 0x9E3779B1 + k0) & _M32) ^ k1
    r = torch.where((idx & 1) == 1, h >> 16, h & 0xFFFF)
    return r >= thresh
