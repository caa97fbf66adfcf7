#!/usr/bin/env python
"""`ncu` target for the K1 kernel on ONE GPU (ncu cannot wrap a multi-rank run: it replays kernels, and a kernel that
waits for another rank never replays). The gather job's only "owner" is this GPU, so the gather warp pulls the
weight's 64-row units from a local shard into the staging buffer with the same ld.global.nc / counter-release code
while the same kernel's TMA producer waits on the unit counters and its tcgen05 pipeline consumes the tiles —
every part of the fused kernel runs; only the NVLink hop is missing. Also launches the same GEMM on a complete
weight for comparison. Prints CUDA-event times (not under ncu: run plain first)."""
import ctypes as C, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops import _lib
from dlti_b200.ops.gemm import gemm

V, I, L, U = C.c_void_p, C.c_int, C.c_longlong, C.c_uint
M, R, K = 4096, 12288, 4096
x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
shard = torch.randn(R, K, device="cuda", dtype=torch.bfloat16) * 0.02
staging = torch.zeros_like(shard)
counters = torch.zeros(R // 64, dtype=torch.int32, device="cuda")
err = torch.zeros(1, dtype=torch.int32, device="cuda")
peers = (C.c_void_p * 16)(shard.data_ptr(), *([None] * 15))
attach = _lib.fn("dlti_gemm2_attach", [V, V, V, L, I, I, I, I, I, U, V, I, U, V])


def fused(epoch):
    _lib.check(attach(peers, _lib.ptr(staging), _lib.ptr(counters), K * 2, R, 1, 0, 0, R // 64, epoch,
                      _lib.ptr(counters), R // 64, epoch, _lib.ptr(err)), "attach")
    return gemm(x, staging, two_cta=True)


def timed(fn, n=5):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(n):
        fn(i)
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3


ep = [0]
def f(i):
    ep[0] += 1
    return fused(ep[0])
y = f(0); ref = gemm(x, shard, two_cta=True); torch.cuda.synchronize()
assert torch.equal(y, ref) and torch.equal(staging, shard) and int(err.item()) == 0
print("K1 local: fused (gather + wait + GEMM) %.1f us, plain GEMM %.1f us, %d MB pulled per launch" %
      (timed(f), timed(lambda i: gemm(x, shard, two_cta=True)), R * K * 2 // 1000000))
