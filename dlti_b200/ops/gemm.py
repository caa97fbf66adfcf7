"""GEMM op: tcgen05/TMEM/TMA kernel on CUDA (csrc/gemm.cu), fp32 PyTorch oracle on CPU.

    C = alpha * (op(A)·op(B) + A2·op(B2)) [+ R]      (16-bit output)
    C (+)= alpha * (...)                              (fp32 output)

Operand conventions (match the C ABI, include/dlti_b200.h):
    a_mn=False: A is [M,K];  a_mn=True: A is stored [K,M]   (contraction dim outermost)
    b_mn=False: B is [N,K] (y = x·W^T, the nn.Linear layout);  b_mn=True: B is stored [K,N]

Replaces torch.nn.functional.linear / cuBLAS in the reference stack.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib


def gemm_ref(a, b, *, a_mn=False, b_mn=False, a2=None, b2=None, residual=None, out=None,
             out_dtype=None, alpha=1.0, accumulate=False):
    """fp32 oracle with identical semantics (used on CPU and by the GPU numerics tests)."""
    A = a.float().t() if a_mn else a.float()
    Bm = b.float() if b_mn else b.float().t()
    acc = A @ Bm
    if a2 is not None:
        A2 = a2.float().t() if a_mn else a2.float()
        B2 = b2.float() if b_mn else b2.float().t()
        acc = acc + A2 @ B2
    acc = acc * alpha
    out_dtype = out_dtype or (out.dtype if out is not None else a.dtype)
    if residual is not None:
        acc = acc + residual.float()
    if out is not None:
        if accumulate:
            out.add_(acc.to(out.dtype))
        else:
            out.copy_(acc.to(out.dtype))
        return out
    return acc.to(out_dtype)


def _gemm_fn():
    return _lib.fn("dlti_gemm", [
        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
        C.c_longlong, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int, C.c_int,
        C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_longlong,
        C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p])


_TWO_CTA_DEFAULT = os.environ.get("DLTI_GEMM_2CTA", "1") == "1"
_FUSE_ROPE = os.environ.get("DLTI_FUSE_ROPE", "1") == "1"      # A/B switches for the fused epilogues
_FUSE_SWIGLU = os.environ.get("DLTI_FUSE_SWIGLU", "1") == "1"


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         a2: torch.Tensor | None = None, b2: torch.Tensor | None = None,
         residual: torch.Tensor | None = None, out: torch.Tensor | None = None,
         out_dtype: torch.dtype | None = None, alpha: float = 1.0, accumulate: bool = False,
         split_k: int = 1, block_n: int = 0, max_ctas: int = 0, gather=None, push=None,
         two_cta: bool | None = None, stream_k: int = -1, scatter=None) -> torch.Tensor:
    """two_cta: run the CTA-pair (tcgen05 cta_group::2) kernel of csrc/gemm2.cu; None = the
    DLTI_GEMM_2CTA environment default (on), applied to shapes with at least 64 256x256 tiles.
    stream_k (pair kernel only): -1 automatic (ragged last wave), 0 off, 1 forced.
    gather: optional parallel.fused.GatherSpec — `b` is then the local staging buffer of a row-
    sharded weight and the kernel pulls the shards from their owner ranks while it computes (K1)."""
    if not _lib.use_native(a, b):
        return gemm_ref(a, b, a_mn=a_mn, b_mn=b_mn, a2=a2, b2=b2, residual=residual, out=out,
                        out_dtype=out_dtype, alpha=alpha, accumulate=accumulate)
    assert a.dim() == 2 and b.dim() == 2 and a.dtype == b.dtype
    assert a.stride(1) == 1 and b.stride(1) == 1, "operands must have unit inner stride"
    M, K = (a.shape[1], a.shape[0]) if a_mn else a.shape
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else b.shape
    assert K == Kb, f"K mismatch {K} vs {Kb}"
    K2 = 0
    if a2 is not None:
        assert b2 is not None and a2.stride(1) == 1 and b2.stride(1) == 1 and a2.dtype == a.dtype
        K2 = a2.shape[0] if a_mn else a2.shape[1]
    out_dtype = out_dtype or (out.dtype if out is not None else a.dtype)
    if out is None and push is None:
        assert not accumulate
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    assert push is not None or (out.shape == (M, N) and out.stride(1) == 1 and out.dtype == out_dtype)
    if residual is not None:
        assert residual.shape == (M, N) and residual.stride(1) == 1 and residual.dtype == out.dtype
    if push is not None:
        # K10: the epilogue stores the partial output into every rank's receive slot (peer stores)
        # (peer pointers, local receive buffer, rank, world, rows per slot[, row pitch, slot stride]): the default
        # layout is [world, rows, N] slots; with an explicit pitch / stride the slots can interleave, e.g. column
        # blocks of one [rows, world * N] matrix (pitch world * N, stride N: the LM-head logits gather)
        peer_arr, recv, rank, world, max_tokens = push[:5]
        ldc_push, slot_stride = (push[5], push[6]) if len(push) > 5 else (N, max_tokens * N)
        assert not a_mn and not b_mn and a2 is None and residual is None and M <= max_tokens
        assert recv.dtype == a.dtype and (len(push) > 5 or recv.shape[-1] == N)
        f = _lib.fn("dlti_gemm_push", [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_longlong,
                                       C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                       C.c_int, C.c_longlong, C.c_void_p])
        _lib.check(f(_lib.ptr(a), _lib.ptr(b), M, N, K, a.stride(0), b.stride(0), ldc_push, _lib.dtype_code(a.dtype),
                     int(block_n), peer_arr, rank, world, slot_stride, _lib.stream_ptr()), "dlti_gemm_push")
        return recv
    if gather is not None and hasattr(gather, "attach"):
        # K1 on the CTA-pair kernel: the launch carries a slice of pending weight gathers (and waits on its own
        # operand's counters when that is still being pulled); see parallel.fused.GatherQueue
        assert not a_mn and split_k == 1 and (gather.sm is None or b.data_ptr() == gather.sm.staging.data_ptr())
        two_cta = True
        gather.attach(2.0 * M * N * (K + K2))
    elif gather is not None:
        assert not a_mn and not accumulate and split_k == 1
        assert b.data_ptr() == gather.staging.data_ptr() and b.stride(0) == gather.staging.stride(0)
        f = _lib.fn("dlti_gemm_gather", [
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
            C.c_longlong, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int,
            C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_longlong,
            C.c_int, C.c_int, C.c_float, C.c_int, C.c_int,
            C.c_void_p, C.c_void_p, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p])
        rc = f(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out), _lib.ptr(residual), M, N, K,
               a.stride(0), b.stride(0), out.stride(0), residual.stride(0) if residual is not None else 0,
               int(b_mn), _lib.ptr(a2), _lib.ptr(b2), K2,
               a2.stride(0) if a2 is not None else 0, b2.stride(0) if b2 is not None else 0,
               _lib.dtype_code(a.dtype), _lib.dtype_code(out_dtype), float(alpha), int(block_n), int(max_ctas),
               gather.peer_arr, _lib.ptr(gather.counters), int(gather.epoch), gather.rank, gather.world,
               gather.rows_per_rank, _lib.ptr(gather.err), _lib.stream_ptr())
        _lib.check(rc, "dlti_gemm_gather")
        return out
    if scatter is not None:
        # K10, prefill-sized steps: the epilogue stores every output row into the receive slot of the rank that
        # owns it (peer stores); scatter = (peer pointer array of the slot buffers, world, rank, rows per owner)
        assert not a_mn and split_k == 1 and residual is None and out_dtype != torch.float32 and N % 64 == 0
        two_cta = True
        f = _lib.fn("dlti_gemm2_attach_scatter", [C.c_void_p, C.c_int, C.c_int, C.c_int])
        _lib.check(f(scatter[0], int(scatter[1]), int(scatter[2]), int(scatter[3])), "gemm2_attach_scatter")
    if two_cta is None:
        two_cta = _TWO_CTA_DEFAULT and block_n == 0 and max_ctas == 0 and ((M + 255) // 256) * ((N + 255) // 256) >= 64
    if two_cta and not a_mn and split_k == 1:
        f = _lib.fn("dlti_gemm_2cta", [
            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
            C.c_longlong, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int,
            C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_longlong,
            C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p])
        rc = f(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out), _lib.ptr(residual), M, N, K,
               a.stride(0), b.stride(0), out.stride(0), residual.stride(0) if residual is not None else 0,
               int(b_mn), _lib.ptr(a2), _lib.ptr(b2), K2,
               a2.stride(0) if a2 is not None else 0, b2.stride(0) if b2 is not None else 0,
               _lib.dtype_code(a.dtype), _lib.dtype_code(out_dtype), float(alpha), int(accumulate),
               int(stream_k), _lib.stream_ptr())
        _lib.check(rc, "dlti_gemm_2cta")
        return out
    rc = _gemm_fn()(
        _lib.ptr(a), _lib.ptr(b), _lib.ptr(out), _lib.ptr(residual), M, N, K,
        a.stride(0), b.stride(0), out.stride(0), residual.stride(0) if residual is not None else 0,
        int(a_mn), int(b_mn), _lib.ptr(a2), _lib.ptr(b2), K2,
        a2.stride(0) if a2 is not None else 0, b2.stride(0) if b2 is not None else 0,
        _lib.dtype_code(a.dtype), _lib.dtype_code(out_dtype), float(alpha), int(accumulate),
        int(split_k), int(block_n), int(max_ctas), _lib.stream_ptr())
    _lib.check(rc, "dlti_gemm")
    return out


def gemm_swiglu(x: torch.Tensor, w_gu: torch.Tensor, gather=None) -> torch.Tensor:
    """h = silu(x · Wg^T) * (x · Wu^T) for w_gu = [gate rows ; up rows] ([2*ff, K]): the gate_up projection
    with SwiGLU applied in the GEMM epilogue (csrc/gemm2.cu). Falls back to GEMM + swiglu_fwd when the
    shape is not covered (CPU tensors, ff % 128 != 0, fewer than 64 tiles)."""
    from .elementwise import swiglu_fwd
    M, K = x.shape
    ff = w_gu.shape[0] // 2
    if (not _lib.use_native(x, w_gu) or not _TWO_CTA_DEFAULT or not _FUSE_SWIGLU or ff % 128 or x.stride(1) != 1 or w_gu.stride(1) != 1
            or (((M + 255) // 256) * (ff // 128) < 64 and gather is None)):
        return swiglu_fwd(gemm(x, w_gu, gather=gather))
    h = torch.empty((M, ff), dtype=x.dtype, device=x.device)
    if gather is not None:
        gather.attach(4.0 * M * ff * K)
    f = _lib.fn("dlti_gemm_swiglu_2cta", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                          C.c_longlong, C.c_longlong, C.c_longlong, C.c_int, C.c_void_p])
    _lib.check(f(_lib.ptr(x), _lib.ptr(w_gu), _lib.ptr(h), M, ff, K, x.stride(0), w_gu.stride(0), h.stride(0),
                 _lib.dtype_code(x.dtype), _lib.stream_ptr()), "dlti_gemm_swiglu_2cta")
    return h


def gemm_rope(x: torch.Tensor, w: torch.Tensor, pos: torch.Tensor, cos_t: torch.Tensor, sin_t: torch.Tensor,
              n_rot_heads: int, hd: int, a2: torch.Tensor | None = None, b2: torch.Tensor | None = None,
              gather=None) -> torch.Tensor:
    """qkv = rope(x · W^T + a2 · b2^T): the QKV projection (with the LoRA K-extension) whose epilogue applies
    the rotary embedding to the first n_rot_heads heads. Falls back to GEMM + rope_ when not covered."""
    from .elementwise import rope_
    M, K = x.shape
    N = w.shape[0]
    if (not _lib.use_native(x, w) or not _TWO_CTA_DEFAULT or not _FUSE_ROPE or hd != 128 or N % 128 or x.stride(1) != 1
            or w.stride(1) != 1 or (((M + 255) // 256) * ((N + 255) // 256) < 64 and gather is None)
            or pos.dtype != torch.int32):
        y = gemm(x, w, a2=a2, b2=b2, gather=gather)
        rope_(y, pos, cos_t, sin_t, n_rot_heads, hd)
        return y
    K2 = 0
    if a2 is not None:
        assert b2 is not None and a2.stride(1) == 1 and b2.stride(1) == 1
        K2 = a2.shape[1]
    y = torch.empty((M, N), dtype=x.dtype, device=x.device)
    if gather is not None:
        gather.attach(2.0 * M * N * (K + K2))
    f = _lib.fn("dlti_gemm_rope_2cta", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_longlong,
                                        C.c_longlong, C.c_longlong, C.c_void_p, C.c_void_p, C.c_int, C.c_longlong,
                                        C.c_longlong, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                        C.c_void_p])
    _lib.check(f(_lib.ptr(x), _lib.ptr(w), _lib.ptr(y), M, N, K, x.stride(0), w.stride(0), y.stride(0),
                 _lib.ptr(a2), _lib.ptr(b2), K2, a2.stride(0) if a2 is not None else 0,
                 b2.stride(0) if b2 is not None else 0, _lib.dtype_code(x.dtype), 1.0, _lib.ptr(pos),
                 _lib.ptr(cos_t), _lib.ptr(sin_t), n_rot_heads * hd, _lib.stream_ptr()), "dlti_gemm_rope_2cta")
    return y


def gemm_swiglu_bwd(dy: torch.Tensor, w_down: torch.Tensor, gu: torch.Tensor, gather=None) -> torch.Tensor:
    """d(gate|up) = swiglu_bwd(dy · W_down, gate|up) for w_down = [d_model, ff] (dy is [T, d_model]): the
    down projection's input-gradient GEMM with the SwiGLU backward in its epilogue."""
    from .elementwise import swiglu_bwd
    M, K = dy.shape
    ff = w_down.shape[1]
    if (not _lib.use_native(dy, w_down) or not _TWO_CTA_DEFAULT or not _FUSE_SWIGLU or ff % 32 or dy.stride(1) != 1
            or w_down.stride(1) != 1 or gu.stride(1) != 1
            or (((M + 255) // 256) * ((ff + 255) // 256) < 64 and gather is None)):
        return swiglu_bwd(gemm(dy, w_down, b_mn=True, gather=gather), gu)
    assert w_down.shape[0] == K and gu.shape == (M, 2 * ff) and gu.dtype == dy.dtype
    dgu = torch.empty_like(gu)
    if gather is not None:
        gather.attach(2.0 * M * ff * K)
    f = _lib.fn("dlti_gemm_swiglu_bwd_2cta", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                              C.c_longlong, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int,
                                              C.c_void_p])
    _lib.check(f(_lib.ptr(dy), _lib.ptr(w_down), _lib.ptr(gu), _lib.ptr(dgu), M, ff, K, dy.stride(0),
                 w_down.stride(0), gu.stride(0), dgu.stride(0), _lib.dtype_code(dy.dtype), _lib.stream_ptr()),
               "dlti_gemm_swiglu_bwd_2cta")
    return dgu


_skinny_ws: dict = {}
_skinny_ws_retired: list = []
_SKINNY_SPLIT = int(os.environ.get("DLTI_SKINNY_SPLIT", "0"))   # 0 = stream-K, -1 = tile-major auto split, n = n splits


def _skinny_workspace(device, numel: int) -> torch.Tensor:
    key = str(device)
    ws = _skinny_ws.get(key)
    if ws is None or ws.numel() < numel:
        if ws is not None:
            _skinny_ws_retired.append(ws)      # captured CUDA graphs may still point at it
        ws = torch.zeros(max(numel, 256 * 32768), dtype=torch.float32, device=device)
        _skinny_ws[key] = ws
    return ws


def gemm_skinny_acc(x: torch.Tensor, w: torch.Tensor, split_k: int | None = None) -> torch.Tensor:
    """fp32 accumulator [M, N] of x[M, K] · w[N, K]^T for decode-sized M (<= 256), left in the shared
    zero-on-entry workspace: the swap-AB kernel makes the weight the 128-row tensor-core operand and cuts the
    (tile, K block) space into equal ranges per CTA (stream-K), reducing fp32 partials with atomics. The
    caller must consume the view at once with an op that re-zeroes it (skinny_finish / ops.*.acc_*) — the
    next gemm_skinny* call reuses the same memory."""
    M, K = x.shape
    N = w.shape[0]
    assert _lib.use_native(x, w) and M <= 256 and N % 8 == 0 and x.stride(1) == 1 and w.stride(1) == 1
    ws = _skinny_workspace(x.device, M * N)
    f = _lib.fn("dlti_gemm_skinny", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_longlong,
                                     C.c_longlong, C.c_int, C.c_int, C.c_void_p])
    _lib.check(f(_lib.ptr(x), _lib.ptr(w), _lib.ptr(ws), M, N, K, x.stride(0), w.stride(0),
                 _lib.dtype_code(x.dtype), int(_SKINNY_SPLIT if split_k is None else split_k), _lib.stream_ptr()),
               "dlti_gemm_skinny")
    return ws[:M * N].view(M, N)


def gemm_skinny(x: torch.Tensor, w: torch.Tensor, out: torch.Tensor | None = None, split_k: int | None = None) -> torch.Tensor:
    """y[M, N] = x[M, K] · w[N, K]^T for decode-sized M (<= 256): gemm_skinny_acc + a finish kernel that
    converts the fp32 accumulator and re-zeroes it. Falls back to the regular kernel on CPU tensors or larger M."""
    M, K = x.shape
    N = w.shape[0]
    if not _lib.use_native(x, w) or M > 256 or N % 8:
        return gemm(x, w, out=out)
    acc = gemm_skinny_acc(x, w, split_k)
    if out is None:
        out = torch.empty((M, N), dtype=x.dtype, device=x.device)
    g = _lib.fn("dlti_skinny_finish", [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_longlong, C.c_int, C.c_void_p])
    _lib.check(g(_lib.ptr(acc), _lib.ptr(out), M, N, out.stride(0), _lib.dtype_code(out.dtype), _lib.stream_ptr()),
               "dlti_skinny_finish")
    return out
