#!/usr/bin/env python
"""Attention kernel micro-benchmark at the training shape (B x S x H x 128, causal): forward and the
backward pair, CUDA-event timed, reported as TFLOP/s (causal FLOPs: fwd 2*2*S^2*hd*H*B/2, bwd x2.5
algorithmic / x3.5 executed by the two-kernel deterministic backward)."""
import json, os, sys, math
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops import attention as A
B, S, H, KV, hd = int(os.environ.get("B", 2)), int(os.environ.get("S", 2048)), 32, 32, 128
qkv = torch.randn(B * S, (H + 2 * KV) * hd, device="cuda", dtype=torch.bfloat16)
cu = torch.arange(0, (B + 1) * S, S, dtype=torch.int32, device="cuda")
o, lse = A.attention_fwd(qkv, cu, S, H, KV, hd); do = torch.randn_like(o)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize(); return s.elapsed_time(e) / n * 1e3
fw = t(lambda: A.attention_fwd(qkv, cu, S, H, KV, hd))
bw = t(lambda: A.attention_bwd(do, qkv, o, lse, cu, S, H, KV, hd))
fl = 4 * S * S * hd * H * B / 2
print(json.dumps({"B": B, "S": S, "fwd_us": fw, "bwd_us": bw, "fwd_tflops": fl / fw / 1e6, "bwd_tflops_algorithmic": 2.5 * fl / bw / 1e6}))
