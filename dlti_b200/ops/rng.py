"""Counter-based dropout masks (Philox-4x32-10), bit-identical between the CUDA kernels
(csrc/common.cuh::philox4x32) and this NumPy restatement used by the CPU oracle.

A mask is a pure function of (seed, stream, element index), so forward, layer recompute and
backward see the same mask without storing it. The reference relies on
torch.nn.Dropout inside PEFT's lora.Linear (lora_dropout=0.05, reference
training/train_baseline.py:134) and on torch's RNG-state stashing in torch.utils.checkpoint.
"""
from __future__ import annotations

import numpy as np
import torch

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = np.uint64(0x9E3779B9)
_W1 = np.uint64(0xBB67AE85)
_MASK = np.uint64(0xFFFFFFFF)
_S32 = np.uint64(32)


def philox4x32(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox-4x32-10 on uint64 arrays holding 32-bit values."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in (c0, c1, c2, c3))
    k0 = np.uint64(k0)
    k1 = np.uint64(k1)
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> _S32, p0 & _MASK
        hi1, lo1 = p1 >> _S32, p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK, lo1, (hi0 ^ c3 ^ k1) & _MASK, lo0
        k0 = (k0 + _W0) & _MASK
        k1 = (k1 + _W1) & _MASK
    return c0, c1, c2, c3


def dropout_threshold(p: float) -> int:
    t = np.float32(p) * np.float32(4294967296.0)
    return int(min(np.float32(t), np.float32(4294967295.0)))


def keep_mask(rows: int, cols: int, p: float, seed: int, stream: int) -> torch.Tensor:
    """Boolean [rows, cols] keep-mask identical to the device kernels'."""
    n = rows * cols
    e = np.arange(n, dtype=np.uint64)
    idx4 = e >> np.uint64(2)
    lane = (e & np.uint64(3)).astype(np.int64)
    u_idx, inv = np.unique(idx4, return_inverse=True)
    s = np.uint64(stream)
    sd = np.uint64(seed)
    r = philox4x32(u_idx & _MASK, u_idx >> _S32, np.full_like(u_idx, s & _MASK),
                   np.full_like(u_idx, s >> _S32), sd & _MASK, sd >> _S32)
    allr = np.stack(r, axis=1)  # [n_unique, 4]
    bits = allr[inv, lane]
    keep = bits >= np.uint64(dropout_threshold(p))
    return torch.from_numpy(keep.reshape(rows, cols))


def dropout_ref(x: torch.Tensor, p: float, seed: int, stream: int) -> torch.Tensor:
    if p <= 0.0:
        return x.clone()
    rows = x.numel() // x.shape[-1]
    m = keep_mask(rows, x.shape[-1], p, seed, stream).to(x.device).view_as(x)
    return (x.float() * m / (1.0 - p)).to(x.dtype)
