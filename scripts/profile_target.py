#!/usr/bin/env python
"""Small deterministic workload for `ncu`: a few launches of the hot kernels at Llama-2-7B shapes
(GEMM NT/NN with LoRA K-extension, attention fwd/bwd, paged decode). Exits 0 quickly."""
import os, sys, math
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops.gemm import gemm
from dlti_b200.ops import attention as A, serving as S
torch.manual_seed(0)
M, d = 4096, 4096
x = torch.randn(M, d, device="cuda", dtype=torch.bfloat16)
w = torch.randn(12288, d, device="cuda", dtype=torch.bfloat16) * 0.02
t = torch.randn(M, 64, device="cuda", dtype=torch.bfloat16); bl = torch.randn(12288, 64, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    y = gemm(x, w, a2=t, b2=bl)                       # fused base + LoRA forward
    dx = gemm(y, w, b_mn=True)                        # input gradient through the same weight
H, KV, hd, S_ = 32, 32, 128, 2048
qkv = torch.randn(2 * S_, (H + 2 * KV) * hd, device="cuda", dtype=torch.bfloat16)
cu = torch.tensor([0, S_, 2 * S_], dtype=torch.int32, device="cuda")
for _ in range(2):
    o, lse = A.attention_fwd(qkv, cu, S_, H, KV, hd)
    dq = A.attention_bwd(torch.randn_like(o), qkv, o, lse, cu, S_, H, KV, hd)
torch.cuda.synchronize()
print("ok")
