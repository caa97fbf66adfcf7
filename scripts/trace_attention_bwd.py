#!/usr/bin/env python
"""Developer probe: per-iteration clock64 timeline of the attention backward kernels (needs a library built
with -DDLTI_ATTN_TRACE, selected with DLTI_GPU_LIB). Prints, for CTA (0,0,0), the cycle stamps of the
softmax warp, the MMA warp and the TMA producer relative to the first event."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops import _lib, attention as A  # noqa: E402

B, S, H, KV, hd = 2, 2048, 32, 32, 128
qkv = torch.randn(B * S, (H + 2 * KV) * hd, device="cuda", dtype=torch.bfloat16)
cu = torch.arange(0, (B + 1) * S, S, dtype=torch.int32, device="cuda")
o, lse = A.attention_fwd(qkv, cu, S, H, KV, hd)
do = torch.randn_like(o)
for _ in range(2):
    A.attention_bwd(do, qkv, o, lse, cu, S, H, KV, hd)
torch.cuda.synchronize()
buf = (C.c_longlong * (2 * 3 * 64 * 4))()
f = _lib.gpu_lib().dlti_attn_trace_read
f.argtypes = [C.POINTER(C.c_longlong)]
assert f(buf) == 0
for k, name in ((0, "dQ kernel"), (1, "dK/dV kernel")):
    base = k * 3 * 64 * 4
    ev = lambda role, it, e: buf[base + (role * 64 + it) * 4 + e]
    t0 = min(x for x in (ev(r, 0, e) for r in range(3) for e in range(4)) if x > 0)
    print(f"== {name}: cycles since first event; softmax[enter, scores, in-regs, published] | mma[Y, S-issued, P-present, acc-issued] | producer[slot free]")
    for it in range(34):
        if ev(0, it, 0) == 0:
            break
        sm = [ev(0, it, e) - t0 for e in range(4)]
        mm = [ev(1, it, e) - t0 if ev(1, it, e) else -1 for e in range(4)]
        print(f"it {it:2d} softmax {sm} | mma {mm} | prod {ev(2, it, 0) - t0 if ev(2, it, 0) else -1}")
