#!/usr/bin/env python
"""Numerics check of the CTA-pair GEMM (csrc/gemm2.cu) against the fp32 oracle; run under `timeout`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.ops.gemm import gemm, gemm_ref  # noqa: E402


def main():
    torch.manual_seed(0)
    dev = "cuda"
    worst = 0.0
    cases = [(256, 256, 64, False), (512, 512, 256, False), (4096, 4096, 4096, False), (300, 520, 200, False),
             (4096, 12288, 4096, False), (2560, 2560, 512, False), (2500, 2704, 136, False), (2560, 2560, 512, True), (512, 512, 256, True), (4096, 4096, 11008, True), (1000, 776, 328, True)]
    for M, N, K, b_mn in cases:
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        b = torch.randn((K, N) if b_mn else (N, K), device=dev, dtype=torch.bfloat16)
        r = torch.randn(M, N, device=dev, dtype=torch.bfloat16)
        a2 = torch.randn(M, 64, device=dev, dtype=torch.bfloat16)
        b2 = torch.randn((64, N) if b_mn else (N, 64), device=dev, dtype=torch.bfloat16)
        for ext in (False, True):
            kw = dict(b_mn=b_mn, residual=r, alpha=0.5)
            if ext:
                kw.update(a2=a2, b2=b2)
            y = gemm(a, b, two_cta=True, **kw)
            torch.cuda.synchronize()
            y0 = gemm(a, b, two_cta=True, stream_k=0, **kw)
            y1 = gemm(a, b, two_cta=True, stream_k=1, **kw)
            y2 = gemm(a, b, two_cta=True, stream_k=1, **kw)      # slot re-armed by the previous launch
            torch.cuda.synchronize()
            assert torch.equal(y1, y2), "stream-K not repeatable"
            d = (y1.float() - y0.float()).abs().max().item()
            assert d <= 0.02 * y0.float().abs().max().item(), f"stream-K differs from data-parallel: {d}"
            ref = gemm_ref(a, b, **kw)
            err = ((y.float() - ref.float()).abs().max() / ref.float().abs().max()).item()
            worst = max(worst, err)
            print(f"M={M} N={N} K={K} b_mn={b_mn} ext={ext} rel_err={err:.2e}", flush=True)
            assert err < 1e-2, "mismatch"
        acc = torch.randn(M, N, device=dev, dtype=torch.float32)
        exp = acc + gemm_ref(a, b, b_mn=b_mn, out_dtype=torch.float32)
        gemm(a, b, b_mn=b_mn, out=acc, accumulate=True, two_cta=True)
        err = ((acc - exp).abs().max() / exp.abs().max()).item()
        print(f"   fp32 accumulate rel_err={err:.2e}", flush=True)
        assert err < 1e-2
    print("2CTA GEMM OK worst", worst)


if __name__ == "__main__":
    main()
