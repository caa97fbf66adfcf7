"""Causal varlen flash attention (fwd + bwd) on token-major merged qkv.

qkv: [T, (H + 2*KV) * hd]  (q heads | k heads | v heads per token row, RoPE applied)
cu_seqlens: int32 [B+1] sequence boundaries inside the T rows.

CUDA: tcgen05 kernels (csrc/attention.cu, head_dim 128). CPU: fp32 PyTorch oracle (also handles
other head dims for the tiny plumbing config). Replaces SDPA / flash-attn.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _debug, _lib

_V, _I, _L, _F = C.c_void_p, C.c_int, C.c_longlong, C.c_float


def _split(qkv, H, KV, hd):
    T = qkv.shape[0]
    q = qkv[:, : H * hd].reshape(T, H, hd)
    k = qkv[:, H * hd:(H + KV) * hd].reshape(T, KV, hd)
    v = qkv[:, (H + KV) * hd:].reshape(T, KV, hd)
    return q, k, v


def attention_fwd_ref(qkv, cu_seqlens, H, KV, hd, scale):
    q, k, v = _split(qkv.float(), H, KV, hd)
    T = qkv.shape[0]
    out = torch.zeros(T, H, hd, dtype=torch.float32, device=qkv.device)
    lse = torch.zeros(H, T, dtype=torch.float32, device=qkv.device)
    G = H // KV
    cu = cu_seqlens.tolist()
    for b in range(len(cu) - 1):
        s, e = cu[b], cu[b + 1]
        L = e - s
        qq = q[s:e].transpose(0, 1)                       # [H, L, hd]
        kk = k[s:e].transpose(0, 1).repeat_interleave(G, 0)
        vv = v[s:e].transpose(0, 1).repeat_interleave(G, 0)
        sc = qq @ kk.transpose(1, 2) * scale
        mask = torch.ones(L, L, dtype=torch.bool, device=qkv.device).tril()
        sc = sc.masked_fill(~mask, float("-inf"))
        lse[:, s:e] = torch.logsumexp(sc, dim=-1)
        out[s:e] = (torch.softmax(sc, dim=-1) @ vv).transpose(0, 1)
    return out.reshape(T, H * hd).to(qkv.dtype), lse


def attention_bwd_ref(dout, qkv, out, lse, cu_seqlens, H, KV, hd, scale):
    q, k, v = _split(qkv.float(), H, KV, hd)
    T = qkv.shape[0]
    do = dout.float().reshape(T, H, hd)
    o = out.float().reshape(T, H, hd)
    dq = torch.zeros_like(q)
    dk = torch.zeros_like(k)
    dv = torch.zeros_like(v)
    G = H // KV
    cu = cu_seqlens.tolist()
    for b in range(len(cu) - 1):
        s, e = cu[b], cu[b + 1]
        L = e - s
        qq = q[s:e].transpose(0, 1)
        kk = k[s:e].transpose(0, 1).repeat_interleave(G, 0)
        vv = v[s:e].transpose(0, 1).repeat_interleave(G, 0)
        dd = do[s:e].transpose(0, 1)
        oo = o[s:e].transpose(0, 1)
        sc = qq @ kk.transpose(1, 2) * scale
        mask = torch.ones(L, L, dtype=torch.bool, device=qkv.device).tril()
        p = torch.exp(sc - lse[:, s:e].unsqueeze(-1)).masked_fill(~mask, 0.0)
        dvh = p.transpose(1, 2) @ dd
        dp = dd @ vv.transpose(1, 2)
        D = (dd * oo).sum(-1, keepdim=True)
        ds = p * (dp - D) * scale
        dq[s:e] = (ds @ kk).transpose(0, 1)
        dkh = ds.transpose(1, 2) @ qq
        dk[s:e] = dkh.reshape(KV, G, L, hd).sum(1).transpose(0, 1)
        dv[s:e] = dvh.reshape(KV, G, L, hd).sum(1).transpose(0, 1)
    return torch.cat([dq.reshape(T, -1), dk.reshape(T, -1), dv.reshape(T, -1)], dim=1).to(qkv.dtype)


def attention_fwd(qkv, cu_seqlens, max_seqlen, H, KV, hd, scale=None, want_lse=True, out=None):
    """Returns (out [T, H*hd], lse [H, T] fp32 or None). `out`: optional contiguous [T, H*hd] destination
    (e.g. a row slice of a larger buffer)."""
    scale = scale if scale is not None else 1.0 / math.sqrt(hd)
    _debug.check_cu_seqlens("attention_fwd", cu_seqlens, qkv.shape[0])
    if not _lib.use_native(qkv):
        o, lse = attention_fwd_ref(qkv, cu_seqlens, H, KV, hd, scale)
        if out is not None:
            out.copy_(o)
            o = out
        return o, lse
    assert qkv.is_contiguous() and cu_seqlens.dtype == torch.int32 and cu_seqlens.is_cuda
    T = qkv.shape[0]
    if out is None:
        out = torch.empty((T, H * hd), dtype=qkv.dtype, device=qkv.device)
    assert out.is_contiguous() and out.shape == (T, H * hd) and out.dtype == qkv.dtype
    lse = torch.empty((H, T), dtype=torch.float32, device=qkv.device) if want_lse else None
    f = _lib.fn("dlti_attn_fwd", [_V, _V, _V, _V, _I, _I, _L, _I, _I, _I, _F, _I, _V])
    _lib.check(f(_lib.ptr(qkv), _lib.ptr(out), _lib.ptr(lse), _lib.ptr(cu_seqlens),
                 cu_seqlens.numel() - 1, int(max_seqlen), T, H, KV, hd, float(scale),
                 _lib.dtype_code(qkv.dtype), _lib.stream_ptr()), "attn_fwd")
    return out, lse


def attention_bwd(dout, qkv, out, lse, cu_seqlens, max_seqlen, H, KV, hd, scale=None):
    """Returns dqkv [T, (H+2KV)*hd] (gradients w.r.t. the post-RoPE q, k and v)."""
    scale = scale if scale is not None else 1.0 / math.sqrt(hd)
    if not _lib.use_native(qkv):
        return attention_bwd_ref(dout, qkv, out, lse, cu_seqlens, H, KV, hd, scale)
    assert dout.is_contiguous() and qkv.is_contiguous() and out.is_contiguous()
    T = qkv.shape[0]
    dqkv = torch.empty_like(qkv)
    D = torch.empty((H, T), dtype=torch.float32, device=qkv.device)
    f = _lib.fn("dlti_attn_bwd", [_V, _V, _V, _V, _V, _V, _V, _I, _I, _L, _I, _I, _I, _F, _I, _V])
    _lib.check(f(_lib.ptr(dout), _lib.ptr(qkv), _lib.ptr(out), _lib.ptr(lse), _lib.ptr(D),
                 _lib.ptr(dqkv), _lib.ptr(cu_seqlens), cu_seqlens.numel() - 1, int(max_seqlen), T,
                 H, KV, hd, float(scale), _lib.dtype_code(qkv.dtype), _lib.stream_ptr()), "attn_bwd")
    return dqkv
