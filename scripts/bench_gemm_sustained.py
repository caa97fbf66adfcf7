#!/usr/bin/env python
"""Sustained (power-capped) GEMM throughput: each variant runs back to back for ~SECS seconds, which is how the
kernels run inside a training step on a 1 kW B200 (the burst numbers of bench_gemm.py are taken before the
power governor reacts). dlti CTA-pair kernel vs torch.matmul (cuBLAS) at the Llama-2-7B shapes, NT and NN,
each measured twice in alternation. Writes gpurun_out/bench_gemm_sustained.json (or $OUT).
  SECS=1.5  STREAMK=-1|0|1 (stream_k argument of the pair kernel)  SHAPES=qkv,o,...  M=4096"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops.gemm import gemm  # noqa: E402

SECS = float(os.environ.get("SECS", 1.5))
SK = int(os.environ.get("STREAMK", -1))


def sustained(fn, nbuf):
    for i in range(3):
        fn(i % nbuf)
    torch.cuda.synchronize()
    s, e, m = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    s.record()
    for i in range(10):
        fn(i % nbuf)
    e.record()
    torch.cuda.synchronize()
    per = s.elapsed_time(e) / 10
    iters = max(20, int(SECS * 1e3 / per))
    half = iters // 2
    for i in range(half):
        fn(i % nbuf)
    m.record()
    for i in range(iters - half):
        fn(i % nbuf)
    e.record()
    torch.cuda.synchronize()
    return m.elapsed_time(e) / (iters - half)      # second half: the power governor has settled


def main():
    M = int(os.environ.get("M", 4096))
    shapes = [("qkv", M, 12288, 4096), ("o", M, 4096, 4096), ("gate_up", M, 22016, 4096),
              ("down", M, 4096, 11008), ("lm_head", M, 32000, 4096)]
    only = os.environ.get("SHAPES")
    if only:
        shapes = [s for s in shapes if s[0] in only.split(",")]
    rows = []
    for name, m, n, k in shapes:
        nbuf = max(2, int(160e6 // (n * k * 2)) + 1)
        a = [torch.randn(m, k, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
        w = [torch.randn(n, k, device="cuda", dtype=torch.bfloat16) for _ in range(nbuf)]
        out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        dy = [torch.randn(m, n, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
        dx = torch.empty(m, k, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * m * n * k
        res = {"shape": name, "M": m, "N": n, "K": k, "secs_per_variant": SECS, "stream_k": SK}
        for rep in range(2):          # alternate so both see the same thermal/power state
            t = sustained(lambda i: torch.matmul(a[i % 2], w[i].t(), out=out), nbuf)
            res[f"cublas_nt_{rep}"] = fl / t / 1e9
            t = sustained(lambda i: gemm(a[i % 2], w[i], out=out, two_cta=True, stream_k=SK), nbuf)
            res[f"dlti_nt_{rep}"] = fl / t / 1e9
            t = sustained(lambda i: torch.matmul(dy[i % 2], w[i], out=dx), nbuf)
            res[f"cublas_nn_{rep}"] = fl / t / 1e9
            t = sustained(lambda i: gemm(dy[i % 2], w[i], b_mn=True, out=dx, two_cta=True, stream_k=SK), nbuf)
            res[f"dlti_nn_{rep}"] = fl / t / 1e9
        rows.append(res)
        print(json.dumps(res), flush=True)
        del a, w, out, dy, dx
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open(os.environ.get("OUT", "gpurun_out/bench_gemm_sustained.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
