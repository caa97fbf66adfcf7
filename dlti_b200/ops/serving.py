"""Serving ops: fused RoPE + paged KV write (K13), paged-KV decode attention (K12), sampling (K14).
CUDA kernels in csrc/serving.cu; PyTorch oracles for CPU / tests."""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _debug, _lib
from .elementwise import rope_ref

_V, _I, _L, _F, _U = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong


def rope_cache_ref(qkv, pos, cos_t, sin_t, slots, k_cache, v_cache, H, KV, hd):
    rope_ref(qkv, pos, cos_t, sin_t, H + KV, hd)
    T = qkv.shape[0]
    k = qkv[:, H * hd:(H + KV) * hd].reshape(T, KV, hd)
    v = qkv[:, (H + KV) * hd:].reshape(T, KV, hd)
    ok = slots >= 0
    kc = k_cache.view(-1, KV, hd)
    vc = v_cache.view(-1, KV, hd)
    kc[slots[ok]] = k[ok]
    vc[slots[ok]] = v[ok]
    return qkv


def rope_cache_(qkv, pos, cos_t, sin_t, slots, k_cache, v_cache, H, KV, hd):
    """In place: rotate q,k of every token row; write rotated k and v to the paged cache at `slots`
    (int64 [T], -1 = skip). Caches are [num_blocks, block_size, KV, hd]."""
    _debug.check_range("rope_cache", "slots", slots, 0, k_cache.shape[0] * k_cache.shape[1], allow=-1)
    _debug.check_range("rope_cache", "positions", pos, 0, cos_t.shape[0])
    if not _lib.use_native(qkv):
        return rope_cache_ref(qkv, pos, cos_t, sin_t, slots, k_cache, v_cache, H, KV, hd)
    assert qkv.stride(1) == 1 and pos.dtype == torch.int32 and slots.dtype == torch.int64
    f = _lib.fn("dlti_rope_cache", [_V, _V, _V, _V, _V, _V, _V, _L, _I, _I, _I, _L, _I, _V])
    _lib.check(f(_lib.ptr(qkv), _lib.ptr(pos), _lib.ptr(cos_t), _lib.ptr(sin_t), _lib.ptr(slots),
                 _lib.ptr(k_cache), _lib.ptr(v_cache), qkv.shape[0], H, KV, hd, qkv.stride(0),
                 _lib.dtype_code(qkv.dtype), _lib.stream_ptr()), "rope_cache")
    return qkv


def acc_rope_cache(acc, dtype, pos, cos_t, sin_t, slots, k_cache, v_cache, H, KV, hd):
    """Decode path: consume the fp32 qkv accumulator [T, (H+2KV)*hd] of gemm_skinny_acc (re-zeroing it) and
    return the 16-bit qkv with q,k rotated; rotated k and v also go to the paged cache at `slots`. Bit-identical
    to skinny_finish + rope_cache_."""
    T, W = acc.shape
    _debug.check_range("acc_rope_cache", "slots", slots, 0, k_cache.shape[0] * k_cache.shape[1], allow=-1)
    _debug.check_range("acc_rope_cache", "positions", pos, 0, cos_t.shape[0])
    assert acc.is_cuda and acc.dtype == torch.float32 and acc.is_contiguous() and W == (H + 2 * KV) * hd
    assert pos.dtype == torch.int32 and slots.dtype == torch.int64
    qkv = torch.empty((T, W), dtype=dtype, device=acc.device)
    f = _lib.fn("dlti_acc_rope_cache", [_V, _V, _V, _V, _V, _V, _V, _V, _L, _I, _I, _I, _L, _I, _V])
    _lib.check(f(_lib.ptr(acc), _lib.ptr(qkv), _lib.ptr(pos), _lib.ptr(cos_t), _lib.ptr(sin_t), _lib.ptr(slots),
                 _lib.ptr(k_cache), _lib.ptr(v_cache), T, H, KV, hd, qkv.stride(0), _lib.dtype_code(dtype),
                 _lib.stream_ptr()), "acc_rope_cache")
    return qkv


def paged_decode_ref(q, k_cache, v_cache, block_tables, ctx_lens, H, KV, hd, scale):
    B = q.shape[0]
    bs = k_cache.shape[1]
    out = torch.zeros(B, H * hd, dtype=torch.float32, device=q.device)
    G = H // KV
    kc, vc = k_cache.view(-1, KV, hd).float(), v_cache.view(-1, KV, hd).float()
    for b in range(B):
        n = int(ctx_lens[b])
        if n == 0:
            continue
        t = torch.arange(n, device=q.device)
        slots = block_tables[b, (t // bs).long()].long() * bs + (t % bs)
        k = kc[slots].transpose(0, 1).repeat_interleave(G, 0)      # [H, n, hd]
        v = vc[slots].transpose(0, 1).repeat_interleave(G, 0)
        qq = q[b, : H * hd].float().view(H, 1, hd)
        p = torch.softmax(qq @ k.transpose(1, 2) * scale, dim=-1)
        out[b] = (p @ v).reshape(H * hd)
    return out.to(q.dtype)


_ws: dict = {}
_ws_retired: list = []


def paged_decode(q, k_cache, v_cache, block_tables, ctx_lens, H, KV, hd, scale=None, out=None,
                 max_ctx: int | None = None):
    """q: [B, >= H*hd] (row b = the single new token of sequence b; head h at columns h*hd..).
    Returns out [B, H*hd]."""
    scale = scale if scale is not None else 1.0 / math.sqrt(hd)
    _debug.check_paged("paged_decode", block_tables, ctx_lens, k_cache.shape[0], k_cache.shape[1])
    if not _lib.use_native(q):
        o = paged_decode_ref(q, k_cache, v_cache, block_tables, ctx_lens, H, KV, hd, scale)
        if out is not None:
            out.copy_(o)
            o = out
        return o
    B = q.shape[0]
    if out is None:
        out = torch.empty((B, H * hd), dtype=q.dtype, device=q.device)
    bs = k_cache.shape[1]
    mc = max_ctx if max_ctx is not None else block_tables.shape[1] * bs
    # enough CTAs (128 threads each, ~4 resident per SM) to fill the machine: short contexts are latency-bound,
    # so small batches are cut into more KV splits
    sms = 148
    splits = max(1, min(16, -(-(4 * sms) // max(1, B * H)), (mc + 255) // 256))
    need = _lib.fn("dlti_paged_decode_workspace_floats", [_I, _I, _I], C.c_int)(B, H, splits)
    key = str(q.device)
    ws = _ws.get(key)
    if ws is None or ws.numel() < need:
        # a CUDA graph captured with the smaller workspace keeps its pointer: the old buffer must stay alive
        # (replacing it made the graphs of the small decode buckets fault once a larger bucket had been captured)
        if ws is not None:
            _ws_retired.append(ws)
        ws = torch.empty(max(need, 1 << 20), dtype=torch.float32, device=q.device)
        _ws[key] = ws
    f = _lib.fn("dlti_paged_decode", [_V, _L, _V, _V, _V, _I, _V, _I, _I, _I, _I, _I, _F, _I, _V, _V, _L, _I, _V])
    _lib.check(f(_lib.ptr(q), q.stride(0), _lib.ptr(k_cache), _lib.ptr(v_cache), _lib.ptr(block_tables),
                 block_tables.shape[1], _lib.ptr(ctx_lens), B, bs, H, KV, hd, float(scale), splits,
                 _lib.ptr(ws), _lib.ptr(out), out.stride(0), _lib.dtype_code(q.dtype), _lib.stream_ptr()),
               "paged_decode")
    return out


def sample_ref(logits, temperature, top_k, top_p, generator=None):
    """Oracle for the kept-set semantics: returns (greedy tokens, keep mask [B,V]) — the sampled
    token of the kernel must lie inside the mask; greedy rows must match exactly."""
    lf = logits.float()
    B, V = lf.shape
    greedy = lf.argmax(-1)
    keep = torch.ones(B, V, dtype=torch.bool, device=lf.device)
    for b in range(B):
        t = float(temperature[b])
        if t <= 0:
            keep[b] = False
            keep[b, greedy[b]] = True
            continue
        p = torch.softmax(lf[b] / t, -1)
        k = int(top_k[b]) if top_k is not None else 0
        tp = float(top_p[b]) if top_p is not None else 1.0
        sp, si = p.sort(descending=True)
        m = torch.ones(V, dtype=torch.bool, device=lf.device)
        sk = sp
        if 0 < k < V:                        # top-k first ...
            m &= p >= sp[k - 1]
            sk = sp[:k]
        if tp < 1.0:                         # ... then the nucleus of the renormalised top-k distribution (vLLM / HF)
            cum = (sk / sk.sum()).cumsum(0)
            n = int((cum < tp).sum().item()) + 1
            m &= p >= sk[min(n, sk.numel()) - 1]
        keep[b] = m
    return greedy, keep


def sample(logits, temperature=None, top_k=None, top_p=None, seeds=None, step: int = 0, want_logprob=False,
           out=None):
    """logits [B, V] -> token ids int64 [B] (+ log-prob at temperature 1). Per-row device tensors:
    temperature fp32 (<= 0 = greedy), top_k int32 (0 = off), top_p fp32 (1 = off), seeds uint64.
    out: optional int64 [B] destination for the token ids."""
    if not _lib.use_native(logits):
        lf = logits.float()
        B = lf.shape[0]
        dst, out = out, lf.argmax(-1)
        if temperature is not None:
            for b in range(B):
                t = float(temperature[b])
                if t > 0:
                    _, keep = sample_ref(lf[b:b + 1], temperature[b:b + 1], None if top_k is None else top_k[b:b + 1],
                                         None if top_p is None else top_p[b:b + 1])
                    p = torch.softmax(lf[b] / t, -1) * keep[0]
                    g = torch.Generator().manual_seed(int(seeds[b]) * 1000003 + step if seeds is not None else step)
                    out[b] = torch.multinomial(p / p.sum(), 1, generator=g)[0]
        lp = torch.log_softmax(lf, -1).gather(-1, out.unsqueeze(-1)).squeeze(-1) if want_logprob else None
        if dst is not None:
            dst.copy_(out)
            out = dst
        return out, lp
    assert logits.dim() == 2 and logits.stride(1) == 1
    B, V = logits.shape
    if out is None:
        out = torch.empty(B, dtype=torch.int64, device=logits.device)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == B
    lp = torch.empty(B, dtype=torch.float32, device=logits.device) if want_logprob else None
    f = _lib.fn("dlti_sample", [_V, _L, _I, _I, _V, _V, _V, _V, _U, _V, _V, _I, _V])
    _lib.check(f(_lib.ptr(logits), logits.stride(0), B, V, _lib.ptr(temperature), _lib.ptr(top_k),
                 _lib.ptr(top_p), _lib.ptr(seeds), int(step), _lib.ptr(out), _lib.ptr(lp),
                 _lib.dtype_code(logits.dtype), _lib.stream_ptr()), "sample")
    return out, lp
