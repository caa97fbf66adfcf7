"""RMSNorm / RoPE / SwiGLU / dropout / embedding ops.

Each op has an explicit forward and backward (the model's autograd functions call them
directly: layer-granular recompute) and two implementations with identical
semantics: the sm_100a kernel (csrc/elementwise.cu) for CUDA tensors and a PyTorch oracle for
CPU tensors / tests. Rounding follows HF Llama (SP:transformers/models/llama/modeling_llama.py
:53-67 RMSNorm, :138-168 RoPE, :171-184 MLP).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _debug, _lib
from .rng import dropout_ref

_V, _I, _L, _F, _U = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong


# ------------------------------------------------------------------------------------------
# RMSNorm
# ------------------------------------------------------------------------------------------
def rmsnorm_fwd_ref(x, w, eps, residual=None, drop_p=0.0, seed=0, stream=0, want_drop=False):
    if residual is not None:
        residual.copy_((x.float() + residual.float()).to(x.dtype))
        x = residual
    xf = x.float()
    rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    y = (w.float() * (xf * rstd).to(x.dtype).float()).to(x.dtype)
    ydrop = dropout_ref(y, drop_p, seed, stream) if want_drop else None
    return y, rstd.squeeze(-1), ydrop


def rmsnorm_fwd(x, w, eps, residual=None, drop_p=0.0, seed=0, stream=0, want_drop=False):
    """y = w * T(x * rstd). Returns (y, rstd[fp32 rows], ydrop or None).

    residual: optional residual stream updated IN PLACE (residual += x) and normalised (the
    serving path's fused add+norm). want_drop: also emit dropout(y) for the LoRA branch."""
    if not _lib.use_native(x):
        return rmsnorm_fwd_ref(x, w, eps, residual, drop_p, seed, stream, want_drop)
    d = x.shape[-1]
    rows = x.numel() // d
    assert x.is_contiguous() and w.is_contiguous() and w.dtype == x.dtype
    y = torch.empty_like(x)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    ydrop = torch.empty_like(x) if want_drop else None
    f = _lib.fn("dlti_rmsnorm_fwd", [_V, _V, _V, _V, _V, _V, _L, _I, _F, _F, _U, _U, _I, _V])
    _lib.check(f(_lib.ptr(x), _lib.ptr(residual), _lib.ptr(w), _lib.ptr(y), _lib.ptr(ydrop),
                 _lib.ptr(rstd), rows, d, eps, float(drop_p if want_drop else 0.0), seed, stream,
                 _lib.dtype_code(x.dtype), _lib.stream_ptr()), "rmsnorm_fwd")
    return y, rstd, ydrop


def acc_add_rmsnorm(acc, residual, w, eps):
    """Decode path: consume the fp32 accumulator `acc` [rows, d] of gemm_skinny_acc (re-zeroing it):
    residual += T(acc) in place, returns y = RMSNorm(residual) * w. Bit-identical to skinny_finish followed
    by rmsnorm_fwd(x, w, eps, residual=residual)."""
    rows, d = acc.shape
    assert acc.is_cuda and acc.dtype == torch.float32 and acc.is_contiguous() and residual.is_contiguous()
    y = torch.empty_like(residual)
    f = _lib.fn("dlti_acc_add_rmsnorm", [_V, _V, _V, _V, _L, _I, _F, _I, _V])
    _lib.check(f(_lib.ptr(acc), _lib.ptr(residual), _lib.ptr(w), _lib.ptr(y), rows, d, eps,
                 _lib.dtype_code(residual.dtype), _lib.stream_ptr()), "acc_add_rmsnorm")
    return y


def acc_swiglu(acc, dtype):
    """Decode path: h = silu(gate) * up from the fp32 gate_up accumulator [rows, 2*ff] (re-zeroed)."""
    rows, ff2 = acc.shape
    assert acc.is_cuda and acc.dtype == torch.float32 and acc.is_contiguous()
    h = torch.empty((rows, ff2 // 2), dtype=dtype, device=acc.device)
    f = _lib.fn("dlti_acc_swiglu", [_V, _V, _L, _I, _L, _I, _V])
    _lib.check(f(_lib.ptr(acc), _lib.ptr(h), rows, ff2 // 2, h.stride(0), _lib.dtype_code(dtype), _lib.stream_ptr()),
               "acc_swiglu")
    return h


def rmsnorm_bwd_ref(dy, x, w, rstd, dres=None):
    g = dy.float() * w.float()
    xh = x.float() * rstd.unsqueeze(-1).view(*x.shape[:-1], 1)
    dot = (g * xh).mean(-1, keepdim=True)
    dx = rstd.view(*x.shape[:-1], 1) * (g - xh * dot)
    if dres is not None:
        dx = dx + dres.float()
    return dx.to(x.dtype)


def rmsnorm_bwd(dy, x, w, rstd, dres=None):
    """dx for frozen norm weights; dres (optional) is added (residual-stream gradient)."""
    if not _lib.use_native(x):
        return rmsnorm_bwd_ref(dy, x, w, rstd, dres)
    d = x.shape[-1]
    rows = x.numel() // d
    assert dy.is_contiguous() and x.is_contiguous()
    dx = torch.empty_like(x)
    f = _lib.fn("dlti_rmsnorm_bwd", [_V, _V, _V, _V, _V, _V, _L, _I, _I, _V])
    _lib.check(f(_lib.ptr(dy), _lib.ptr(x), _lib.ptr(w), _lib.ptr(rstd), _lib.ptr(dres),
                 _lib.ptr(dx), rows, d, _lib.dtype_code(x.dtype), _lib.stream_ptr()), "rmsnorm_bwd")
    return dx


# ------------------------------------------------------------------------------------------
# RoPE
# ------------------------------------------------------------------------------------------
_rope_cache: dict = {}


def rope_tables(max_pos: int, hd: int, theta: float, device) -> tuple[torch.Tensor, torch.Tensor]:
    """fp32 cos/sin tables [max_pos, hd/2] (HF computes them in fp32, then casts)."""
    key = (max_pos, hd, float(theta), str(device))
    t = _rope_cache.get(key)
    if t is None:
        inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float32) / hd))
        ang = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv[None, :]
        t = (ang.cos().contiguous().to(device), ang.sin().contiguous().to(device))
        _rope_cache[key] = t
    return t


def rope_ref(buf, pos, cos_t, sin_t, n_rot_heads, hd, backward=False):
    T = buf.shape[0]
    half = hd // 2
    v = buf[:, : n_rot_heads * hd].reshape(T, n_rot_heads, hd).float()
    c = cos_t[pos.long()].to(buf.dtype).float()[:, None, :]
    s = sin_t[pos.long()].to(buf.dtype).float()[:, None, :]
    if backward:
        s = -s
    a, b = v[..., :half], v[..., half:]
    out = torch.cat([a * c - b * s, b * c + a * s], dim=-1).to(buf.dtype)
    buf[:, : n_rot_heads * hd] = out.reshape(T, n_rot_heads * hd)
    return buf


def rope_(buf, pos, cos_t, sin_t, n_rot_heads, hd, backward=False):
    """In-place rotate-half RoPE on the first n_rot_heads heads of each row of buf [T, W]
    (the merged qkv activations: q heads then k heads). pos: int32 [T]."""
    if not _lib.use_native(buf):
        return rope_ref(buf, pos, cos_t, sin_t, n_rot_heads, hd, backward)
    assert buf.dim() == 2 and buf.stride(1) == 1 and pos.dtype == torch.int32
    f = _lib.fn("dlti_rope", [_V, _V, _V, _V, _L, _I, _I, _L, _I, _I, _V])
    _lib.check(f(_lib.ptr(buf), _lib.ptr(pos), _lib.ptr(cos_t), _lib.ptr(sin_t), buf.shape[0],
                 n_rot_heads, hd, buf.stride(0), int(backward), _lib.dtype_code(buf.dtype),
                 _lib.stream_ptr()), "rope")
    return buf


# ------------------------------------------------------------------------------------------
# SwiGLU
# ------------------------------------------------------------------------------------------
def swiglu_fwd_ref(gu):
    ff = gu.shape[-1] // 2
    g, u = gu[..., :ff].float(), gu[..., ff:].float()
    return (torch.nn.functional.silu(g).to(gu.dtype).float() * u).to(gu.dtype)


def swiglu_fwd(gu):
    """h = silu(gate) * up for gu = [T, 2*ff] (gate columns first)."""
    if not _lib.use_native(gu):
        return swiglu_fwd_ref(gu)
    assert gu.dim() == 2 and gu.stride(1) == 1
    ff = gu.shape[1] // 2
    h = torch.empty((gu.shape[0], ff), dtype=gu.dtype, device=gu.device)
    f = _lib.fn("dlti_swiglu_fwd", [_V, _V, _L, _I, _L, _L, _I, _V])
    _lib.check(f(_lib.ptr(gu), _lib.ptr(h), gu.shape[0], ff, gu.stride(0), h.stride(0),
                 _lib.dtype_code(gu.dtype), _lib.stream_ptr()), "swiglu_fwd")
    return h


def swiglu_bwd_ref(dh, gu):
    ff = gu.shape[-1] // 2
    g, u, d = gu[..., :ff].float(), gu[..., ff:].float(), dh.float()
    sig = torch.sigmoid(g)
    dg = d * u * sig * (1 + g * (1 - sig))
    du = d * g * sig
    return torch.cat([dg, du], dim=-1).to(gu.dtype)


def swiglu_bwd(dh, gu):
    if not _lib.use_native(gu):
        return swiglu_bwd_ref(dh, gu)
    assert gu.stride(1) == 1 and dh.stride(1) == 1
    ff = gu.shape[1] // 2
    dgu = torch.empty_like(gu)
    f = _lib.fn("dlti_swiglu_bwd", [_V, _V, _V, _L, _I, _L, _L, _I, _V])
    _lib.check(f(_lib.ptr(dh), _lib.ptr(gu), _lib.ptr(dgu), gu.shape[0], ff, dh.stride(0),
                 gu.stride(0), _lib.dtype_code(gu.dtype), _lib.stream_ptr()), "swiglu_bwd")
    return dgu


# ------------------------------------------------------------------------------------------
# Dropout
# ------------------------------------------------------------------------------------------
def dropout(x, p, seed, stream):
    """Counter-based dropout over a 2-D tensor; mask index = row * cols + col."""
    if p <= 0.0:
        return x
    if not _lib.use_native(x):
        return dropout_ref(x, p, seed, stream)
    assert x.dim() == 2 and x.stride(1) == 1
    y = torch.empty((x.shape[0], x.shape[1]), dtype=x.dtype, device=x.device)
    f = _lib.fn("dlti_dropout", [_V, _V, _L, _I, _L, _L, _F, _U, _U, _I, _V])
    _lib.check(f(_lib.ptr(x), _lib.ptr(y), x.shape[0], x.shape[1], x.stride(0), y.stride(0),
                 float(p), seed, stream, _lib.dtype_code(x.dtype), _lib.stream_ptr()), "dropout")
    return y


def add_dropout(a, b, p, seed, stream):
    """a + dropout(b): merges the LoRA-branch input gradient (which passes back through the
    dropout mask) into the base-path gradient."""
    if not _lib.use_native(a):
        return (a.float() + dropout_ref(b, p, seed, stream).float()).to(a.dtype) if p > 0 else a + b
    assert a.is_contiguous() and b.is_contiguous() and a.shape == b.shape
    y = torch.empty_like(a)
    cols = a.shape[-1]
    f = _lib.fn("dlti_add_dropout", [_V, _V, _V, _L, _I, _F, _U, _U, _I, _V])
    _lib.check(f(_lib.ptr(a), _lib.ptr(b), _lib.ptr(y), a.numel() // cols, cols, float(p), seed,
                 stream, _lib.dtype_code(a.dtype), _lib.stream_ptr()), "add_dropout")
    return y


# ------------------------------------------------------------------------------------------
# Embedding
# ------------------------------------------------------------------------------------------
def embedding(ids, table, vocab_start: int = 0):
    """out[t] = table[ids[t] - vocab_start], zero rows for ids outside this vocab shard. `table` may be a
    parallel.fused.PeerRows (vocabulary-sharded table read in place on its owner ranks)."""
    if hasattr(table, "lookup"):
        return table.lookup(ids)
    if vocab_start == 0:
        _debug.check_range("embedding", "ids", ids, 0, table.shape[0])
    if not _lib.use_native(table):
        local = ids - vocab_start
        ok = (local >= 0) & (local < table.shape[0])
        out = table[local.clamp(0, table.shape[0] - 1)]
        return out * ok.unsqueeze(-1).to(out.dtype)
    assert ids.dtype == torch.int64 and ids.is_contiguous() and table.is_contiguous()
    out = torch.empty((ids.numel(), table.shape[1]), dtype=table.dtype, device=table.device)
    f = _lib.fn("dlti_embedding", [_V, _V, _V, _L, _I, _L, _L, _I, _V])
    _lib.check(f(_lib.ptr(ids), _lib.ptr(table), _lib.ptr(out), ids.numel(), table.shape[1],
                 vocab_start, table.shape[0], _lib.dtype_code(table.dtype), _lib.stream_ptr()),
               "embedding")
    return out.view(*ids.shape, table.shape[1])
