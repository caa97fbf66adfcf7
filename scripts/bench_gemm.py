#!/usr/bin/env python
"""Micro-benchmark: dlti tcgen05 GEMM vs torch.matmul (cuBLAS) at the Llama-2-7B shapes.

Timing: CUDA events, 5 warm-up + 20 timed launches, operands rotated over enough distinct
buffers that the working set exceeds the 126 MB L2. Prints TFLOP/s and fraction of the measured
cuBLAS peak in MEASURED_PEAKS.json (SURVEY.md §8 K2 shapes).
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops.gemm import gemm  # noqa: E402


def timeit(fn, nbuf, warm=5, iters=20):
    for i in range(warm):
        fn(i % nbuf)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(iters):
        fn(i % nbuf)
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = peaks.get("bf16_tflops", 1590.0)
    M = int(os.environ.get("M", 4096))
    shapes = [("qkv", M, 12288, 4096), ("o", M, 4096, 4096), ("gate_up", M, 22016, 4096),
              ("down", M, 4096, 11008), ("lm_head", M, 32000, 4096), ("sq8k", 8192, 8192, 8192)]
    rows = []
    for name, m, n, k in shapes:
        nbuf = max(2, int(160e6 // (n * k * 2)) + 1)
        a = [torch.randn(m, k, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
        w = [torch.randn(n, k, device="cuda", dtype=torch.bfloat16) for _ in range(nbuf)]
        out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * m * n * k
        res = {"shape": name, "M": m, "N": n, "K": k}
        for bn in (256, 128):
            t = timeit(lambda i: gemm(a[i % 2], w[i], out=out, block_n=bn), nbuf)
            res[f"dlti_nt_bn{bn}_tflops"] = fl / t / 1e9
        if os.environ.get("BENCH_2CTA", "1") == "1":
            t = timeit(lambda i: gemm(a[i % 2], w[i], out=out, two_cta=True), nbuf)
            res["dlti_nt_2cta_tflops"] = fl / t / 1e9
            t = timeit(lambda i: gemm(a[i % 2], w[i], out=out, two_cta=True, stream_k=0), nbuf)
            res["dlti_nt_2cta_nosk_tflops"] = fl / t / 1e9
        t = timeit(lambda i: torch.matmul(a[i % 2], w[i].t(), out=out), nbuf)
        res["cublas_nt_tflops"] = fl / t / 1e9
        # dX form: dy[m,n] @ W[n,k] -> [m,k]
        dy = [torch.randn(m, n, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
        dx = torch.empty(m, k, device="cuda", dtype=torch.bfloat16)
        t = timeit(lambda i: gemm(dy[i % 2], w[i], b_mn=True, out=dx), nbuf)
        res["dlti_nn_tflops"] = fl / t / 1e9
        if os.environ.get("BENCH_2CTA", "1") == "1":
            t = timeit(lambda i: gemm(dy[i % 2], w[i], b_mn=True, out=dx, two_cta=True), nbuf)
            res["dlti_nn_2cta_tflops"] = fl / t / 1e9
            t = timeit(lambda i: gemm(dy[i % 2], w[i], b_mn=True, out=dx, two_cta=True, stream_k=0), nbuf)
            res["dlti_nn_2cta_nosk_tflops"] = fl / t / 1e9
        t = timeit(lambda i: torch.matmul(dy[i % 2], w[i], out=dx), nbuf)
        res["cublas_nn_tflops"] = fl / t / 1e9
        res["frac_of_measured_peak"] = max(res["dlti_nt_bn256_tflops"], res["dlti_nt_bn128_tflops"]) / peak
        rows.append(res)
        print(json.dumps(res), flush=True)
        del a, w, out, dy, dx
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/bench_gemm.json", "w"), indent=1)


if __name__ == "__main__":
    main()
