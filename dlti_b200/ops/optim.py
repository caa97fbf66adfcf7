"""Optimizer-side ops: gradient sum-of-squares (norm + overflow detection), fused
unscale + clip + AdamW + 16-bit write-back on flat fp32 partitions, batched fp32->16-bit casts.

Replaces torch.optim.AdamW / DeepSpeed FusedAdam (multi_tensor_adam.cu) and the separate
overflow-check / unscale / clip kernels.
"""
from __future__ import annotations

import ctypes as C
import math
import struct

import torch

from . import _lib

_V, _I, _L, _F = C.c_void_p, C.c_int, C.c_longlong, C.c_float

_scratch: dict = {}


def _sumsq_scratch(device) -> torch.Tensor:
    key = str(device)
    s = _scratch.get(key)
    if s is None:
        n = _lib.fn("dlti_sumsq_scratch_floats", [], C.c_int)()
        s = torch.empty(n, dtype=torch.float32, device=device)
        _scratch[key] = s
    return s


def sumsq(g: torch.Tensor, out: torch.Tensor | None = None, accumulate: bool = False) -> torch.Tensor:
    """out[0] (+)= sum(g**2) for a flat fp32 tensor; non-finite grads give a non-finite result."""
    if not _lib.use_native(g):
        v = g.float().pow(2).sum().reshape(1)
        if out is None:
            return v
        out.copy_(out + v if accumulate else v)
        return out
    assert g.dtype == torch.float32 and g.is_contiguous()
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=g.device)
    f = _lib.fn("dlti_sumsq", [_V, _L, _V, _V, _I, _V])
    _lib.check(f(_lib.ptr(g), g.numel(), _lib.ptr(_sumsq_scratch(g.device)), _lib.ptr(out),
                 int(accumulate), _lib.stream_ptr()), "sumsq")
    return out


def adamw_step_ref(p, m, v, g, *, lr, betas, eps, weight_decay, step, gnorm_sq=None, inv_scale=1.0,
                   max_norm=0.0, p16=None, zero_grad=False):
    coef = inv_scale
    if gnorm_sq is not None:
        nsq = float(gnorm_sq)
        if not math.isfinite(nsq):
            return True
        if max_norm > 0:
            coef *= min(1.0, max_norm / (math.sqrt(nsq) * inv_scale + 1e-6))
    b1, b2 = betas
    ge = g * coef
    m.mul_(b1).add_(ge, alpha=1 - b1)
    v.mul_(b2).addcmul_(ge, ge, value=1 - b2)
    bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
    upd = (m / bc1) / ((v / bc2).sqrt() + eps) + weight_decay * p
    p.add_(upd, alpha=-lr)
    if p16 is not None:
        p16.copy_(p)
    if zero_grad:
        g.zero_()
    return False


def adamw_step(p, m, v, g, *, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, step=1,
               gnorm_sq=None, inv_scale=1.0, max_norm=0.0, p16=None, zero_grad=False,
               skipped_out=None):
    """Fused AdamW over flat fp32 (p, m, v, g).

    gnorm_sq: 1-element fp32 tensor holding the global (all-rank) sum of squares of the still
    loss-scaled gradients; non-finite => the step is skipped (fp16 overflow). Gradients are
    multiplied by inv_scale and by the global-norm clip coefficient for max_norm inside the kernel.
    On CUDA the skip decision stays on the device (skipped_out int32[1]); the CPU oracle returns it.
    """
    if not _lib.use_native(p):
        return adamw_step_ref(p, m, v, g, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                              step=step, gnorm_sq=gnorm_sq, inv_scale=inv_scale, max_norm=max_norm,
                              p16=p16, zero_grad=zero_grad)
    for t in (p, m, v, g):
        assert t.dtype == torch.float32 and t.is_contiguous() and t.numel() == p.numel()
    f = _lib.fn("dlti_adamw", [_V, _V, _V, _V, _V, _I, _L, _F, _F, _F, _F, _F, _I, _V, _F, _F, _V,
                               _I, _V])
    _lib.check(f(_lib.ptr(p), _lib.ptr(m), _lib.ptr(v), _lib.ptr(g), _lib.ptr(p16),
                 _lib.dtype_code(p16.dtype) if p16 is not None else 0, p.numel(), lr, betas[0],
                 betas[1], eps, weight_decay, int(step), _lib.ptr(gnorm_sq), float(inv_scale),
                 float(max_norm), _lib.ptr(skipped_out), int(zero_grad), _lib.stream_ptr()), "adamw")
    return None


class MultiCast:
    """One-launch batch of 2-D fp32 -> 16-bit casts with per-job scale (refreshes the padded
    LoRA compute buffers from the flat fp32 master after each optimizer step)."""

    def __init__(self, jobs: list[tuple[torch.Tensor, torch.Tensor, float]]):
        # jobs: (src fp32 2-D view, dst 16-bit 2-D view, scale)
        self.jobs = jobs
        self.native = bool(jobs) and jobs[0][0].is_cuda
        if self.native:
            sz = _lib.fn("dlti_cast_job_bytes", [], C.c_int)()
            assert sz == 48, sz
            raw = bytearray()
            for src, dst, scale in jobs:
                assert src.dtype == torch.float32 and src.stride(1) == 1 and dst.stride(1) == 1
                assert src.shape == dst.shape
                raw += struct.pack("<QQiiqqfi", src.data_ptr(), dst.data_ptr(), src.shape[0],
                                   src.shape[1], src.stride(0), dst.stride(0), float(scale), 0)
            self.table = torch.frombuffer(raw, dtype=torch.uint8).clone().to(jobs[0][0].device)
            self.dst_dtype = jobs[0][1].dtype

    def run(self) -> None:
        if not self.jobs:
            return
        if not self.native:
            for src, dst, scale in self.jobs:
                dst.copy_((src * scale).to(dst.dtype))
            return
        f = _lib.fn("dlti_multi_cast", [_V, _I, _I, _V])
        _lib.check(f(_lib.ptr(self.table), len(self.jobs), _lib.dtype_code(self.dst_dtype),
                     _lib.stream_ptr()), "multi_cast")
