"""Opt-in argument validation for the kernels that index memory through caller-supplied integers
(SURVEY.md §5.2 "race detection / sanitizers"). compute-sanitizer is not usable on the GPU pool this framework is
developed on (runs under it have left GPUs needing a reset), so out-of-bounds accesses are hunted with checks of
our own: `DLTI_DEBUG_CHECKS=1` makes every such op verify its index tensors on the host before launching
(token ids against the vocabulary, KV slots and block-table entries against the cache, context lengths against
the table width, cu_seqlens monotonic and inside the token rows). Each check synchronises the stream, so this is a
debugging mode, not a production one. The matching device-side mode is the `libdlti_b200_debug.so` build
(`python -m dlti_b200.build --debug-checks`), which adds signal-pad sequence checks to the peer-memory barriers.
"""
from __future__ import annotations

import os

import torch

ENABLED = os.environ.get("DLTI_DEBUG_CHECKS", "0") == "1"


class DltiIndexError(IndexError):
    pass


def enable(on: bool = True) -> None:
    global ENABLED
    ENABLED = bool(on)


def check_range(op: str, name: str, t: torch.Tensor | None, lo: int, hi: int, allow: int | None = None) -> None:
    """Every element of `t` must lie in [lo, hi) (or equal `allow`, e.g. the -1 "skip" slot)."""
    if not ENABLED or t is None or t.numel() == 0:
        return
    x = t.detach().to("cpu", torch.int64)
    bad = (x < lo) | (x >= hi)
    if allow is not None:
        bad &= x != allow
    if bool(bad.any()):
        i = int(bad.reshape(-1).nonzero()[0])
        raise DltiIndexError(f"{op}: {name}[{i}] = {int(x.reshape(-1)[i])} outside [{lo}, {hi})"
                             + (f" (and != {allow})" if allow is not None else ""))


def check_cu_seqlens(op: str, cu: torch.Tensor | None, total_rows: int) -> None:
    if not ENABLED or cu is None:
        return
    x = cu.detach().to("cpu", torch.int64)
    if x.numel() < 2 or int(x[0]) != 0 or bool((x[1:] < x[:-1]).any()) or int(x[-1]) > total_rows:
        raise DltiIndexError(f"{op}: cu_seqlens {x.tolist()[:8]}... must start at 0, be non-decreasing and end <= {total_rows}")


def check_paged(op: str, block_tables: torch.Tensor, ctx_lens: torch.Tensor, num_blocks: int, block_size: int) -> None:
    """Block-table entries that a sequence's context actually reaches must name real blocks; a context must fit
    the table width."""
    if not ENABLED:
        return
    bt = block_tables.detach().to("cpu", torch.int64)
    cl = ctx_lens.detach().to("cpu", torch.int64)
    width = bt.shape[1]
    need = (cl + block_size - 1) // block_size
    if bool((cl < 0).any()) or bool((need > width).any()):
        b = int(((cl < 0) | (need > width)).nonzero()[0])
        raise DltiIndexError(f"{op}: ctx_lens[{b}] = {int(cl[b])} does not fit {width} blocks of {block_size}")
    used = torch.arange(width).unsqueeze(0) < need.unsqueeze(1)
    bad = used & ((bt < 0) | (bt >= num_blocks))
    if bool(bad.any()):
        b, j = (int(v) for v in bad.nonzero()[0])
        raise DltiIndexError(f"{op}: block_tables[{b}, {j}] = {int(bt[b, j])} outside [0, {num_blocks})")
