"""Fused softmax cross-entropy (forward + in-place backward) over 16-bit logits.

Semantics of HF LlamaForCausalLM's loss (SP:transformers/models/llama/modeling_llama.py:487-491):
fp32 log-softmax, labels == ignore_index (-100) contribute neither loss nor gradient. The caller
shifts labels by one and chooses the normalisation (token-normalised by default).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

_V, _I, _L = C.c_void_p, C.c_int, C.c_longlong


def cross_entropy_ref(logits, labels, gscale=None, ignore_index=-100, want_grad=True):
    lf = logits.float()
    lse = torch.logsumexp(lf, dim=-1)
    valid = labels != ignore_index
    safe = labels.clamp_min(0)
    picked = lf.gather(-1, safe.unsqueeze(-1)).squeeze(-1)
    loss_rows = torch.where(valid, lse - picked, torch.zeros_like(lse))
    dlogits = None
    if want_grad:
        p = torch.softmax(lf, dim=-1)
        p.scatter_add_(-1, safe.unsqueeze(-1), -torch.ones_like(picked).unsqueeze(-1))
        g = gscale.float().reshape(()) if torch.is_tensor(gscale) else float(gscale)
        dlogits = (p * g * valid.unsqueeze(-1)).to(logits.dtype)
    return loss_rows, dlogits


def cross_entropy(logits, labels, gscale=None, ignore_index=-100, want_grad=True, inplace=True):
    """Returns (loss_rows fp32 [T], dlogits or None). dlogits = (softmax - onehot) * gscale, where
    gscale is a 1-element fp32 device tensor (read on device: no host sync). With inplace=True
    the gradient overwrites `logits`."""
    if not _lib.use_native(logits):
        return cross_entropy_ref(logits, labels, gscale, ignore_index, want_grad)
    assert logits.dim() == 2 and logits.stride(1) == 1 and labels.dtype == torch.int64
    T, V = logits.shape
    loss_rows = torch.empty(T, dtype=torch.float32, device=logits.device)
    dl = None
    if want_grad:
        assert torch.is_tensor(gscale) and gscale.dtype == torch.float32 and gscale.is_cuda
        dl = logits if inplace else torch.empty_like(logits)
    f = _lib.fn("dlti_cross_entropy", [_V, _V, _V, _V, _V, _L, _I, _L, _L, _L, _I, _V])
    _lib.check(f(_lib.ptr(logits), _lib.ptr(dl), _lib.ptr(labels), _lib.ptr(loss_rows),
                 _lib.ptr(gscale), T, V, logits.stride(0), dl.stride(0) if dl is not None else 0,
                 ignore_index, _lib.dtype_code(logits.dtype), _lib.stream_ptr()), "cross_entropy")
    return loss_rows, dl
