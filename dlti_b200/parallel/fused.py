"""Peer-memory ("fused") communication backend over NVLink 5 / NVSwitch.

A *symmetric arena*: every rank allocates the same sequence of chunks; each chunk is mapped into
every peer's address space (torch symmetric memory — CUDA VMM + handle exchange — when available,
classic CUDA IPC otherwise), so a kernel on rank i can load/store rank j's buffers directly. All
ZeRO-3 weight shards, the flat LoRA master/gradient buffers, gather staging and the signal words
live in the arena.

On top of it (kernels in csrc/comm.cu and the GATHER variant of csrc/gemm.cu):
  barrier()                      signal-word barrier kernel (no NCCL)
  all_gather_rows()              pull-based all-gather of a row-sharded matrix
  gather_spec() / fused GEMM     all-gather fused INTO the GEMM: gather warps of the persistent
                                 GEMM pull remote 64-row units into local staging, bump per-unit
                                 counters, and the TMA producer consumes tiles as units land (K1)
  reduce_scatter_accumulate()    grad_part += sum over peers (peer loads, fixed order)      (K9)
  global_sumsq()                 norm / overflow exchange over peer stores                    (K9)
  adamw_push()                   AdamW on the owned slice + push of new values to all peers   (K9)

This replaces deepspeed.comm / NCCL for the hot collectives; torch.distributed
stays for bootstrap and as the "nccl" baseline backend.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from ..ops import _lib
from .comm import DistInfo

_V, _I, _L, _F, _U = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_uint
MAX_RANKS = 16


class _RawCuda:
    """Minimal __cuda_array_interface__ holder so torch can view foreign device memory."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2}


class SymmetricArena:
    def __init__(self, info: DistInfo, chunk_bytes: int = 2 << 30):
        self.info = info
        self.chunk_bytes = chunk_bytes
        self.chunks: list[dict] = []       # {"local": tensor(uint8), "ptrs": [int]*world, "used": int}
        self.mode = None
        self._keep = []

    # -- chunk creation (collective) ---------------------------------------------------------
    def _new_chunk(self, nbytes: int) -> dict:
        info = self.info
        nbytes = (nbytes + (2 << 20) - 1) // (2 << 20) * (2 << 20)
        if self.mode in (None, "symm"):
            try:
                import torch.distributed._symmetric_memory as symm

                t = symm.empty(nbytes, dtype=torch.uint8, device=info.device)
                hdl = symm.rendezvous(t, dist.group.WORLD)
                ptrs = [int(p) for p in hdl.buffer_ptrs]
                assert ptrs[info.rank] == t.data_ptr()
                self.mode = "symm"
                self._keep.append(hdl)
                mc = 0
                try:
                    mc = int(hdl.multicast_ptr) if hdl.has_multicast_support(info.device.type, info.device.index) else 0
                except Exception:
                    mc = 0
                return {"local": t, "ptrs": ptrs, "used": 0, "multicast": mc}
            except Exception as e:  # pragma: no cover - depends on the box
                if self.mode == "symm":
                    raise
                if info.is_main:
                    print(f"[fused] torch symmetric memory unavailable ({type(e).__name__}: {e}); using CUDA IPC")
                self.mode = "ipc"
        # ---- CUDA IPC fallback: torch allocation + cudaIpc handle exchange ----
        t = torch.empty(nbytes, dtype=torch.uint8, device=info.device)
        base = C.c_void_p()
        size = C.c_ulonglong()
        _lib.check(_lib.fn("dlti_mem_base", [_V, C.POINTER(C.c_void_p), C.POINTER(C.c_ulonglong)])(
            t.data_ptr(), C.byref(base), C.byref(size)), "mem_base")
        handle = (C.c_ubyte * 64)()
        _lib.check(_lib.fn("dlti_ipc_get_handle", [_V, _V])(base.value, handle), "ipc_get_handle")
        mine = (bytes(handle), t.data_ptr() - base.value)
        allh = [None] * info.world_size
        dist.all_gather_object(allh, mine)
        ptrs = []
        for r, (h, off) in enumerate(allh):
            if r == info.rank:
                ptrs.append(t.data_ptr())
                continue
            out = C.c_void_p()
            buf = (C.c_ubyte * 64).from_buffer_copy(h)
            _lib.check(_lib.fn("dlti_ipc_open", [_V, C.POINTER(C.c_void_p)])(buf, C.byref(out)), "ipc_open")
            ptrs.append(out.value + off)
        return {"local": t, "ptrs": ptrs, "used": 0, "multicast": 0}

    def alloc(self, nbytes: int, align: int = 1024) -> tuple[torch.Tensor, list[int]]:
        """Collective: every rank must call with the same sizes in the same order."""
        nbytes = max(align, (nbytes + align - 1) // align * align)
        ch = self.chunks[-1] if self.chunks else None
        if ch is None or ch["used"] + nbytes > ch["local"].numel():
            ch = self._new_chunk(max(self.chunk_bytes, nbytes))
            self.chunks.append(ch)
        off = ch["used"]
        ch["used"] += nbytes
        return ch["local"][off:off + nbytes], [p + off for p in ch["ptrs"]]

    def total_bytes(self) -> int:
        return sum(c["local"].numel() for c in self.chunks)


class GatherSpec:
    """Arguments of the fused all-gather -> GEMM kernel for one row-sharded weight matrix."""

    def __init__(self, peer_ptrs, staging: torch.Tensor, counters: torch.Tensor, rows_per_rank: int,
                 rank: int, world: int, err: torch.Tensor):
        self.peer_arr = (C.c_void_p * MAX_RANKS)(*peer_ptrs, *([None] * (MAX_RANKS - len(peer_ptrs))))
        self.staging, self.counters, self.err = staging, counters, err
        self.rows_per_rank, self.rank, self.world = rows_per_rank, rank, world
        self.epoch = 0
        self.used_epoch = -1      # epoch whose gather has already been launched (see LayerWeights.g)


class StagedMatrix:
    """One row-sharded weight matrix kind of one staging slot (K1 on the CTA-pair GEMM, csrc/gemm2.cu): the
    local full-size staging buffer, its per-64-row-unit counters and the bookkeeping of which job positions of
    the current epoch have already been attached to a launch. `bind()` points it at a layer's shards and starts
    a new epoch; positions map to units as (rot + j) % n_units with rot = this rank's first unit, so at any
    moment the ranks pull from different owners."""

    def __init__(self, fc: "FusedComm", rows_per_rank: int, cols: int, dtype: torch.dtype):
        assert rows_per_rank % 64 == 0
        self.fc = fc
        self.rows_per_rank, self.cols = rows_per_rank, cols
        self.units_per_rank = rows_per_rank // 64
        self.n_units = self.units_per_rank * fc.world
        self.row_bytes = cols * torch.empty(0, dtype=dtype).element_size()
        self.unit_bytes = 64 * self.row_bytes
        self.staging = torch.empty(rows_per_rank * fc.world, cols, dtype=dtype, device=fc.info.device)
        self.counters = torch.zeros(self.n_units, dtype=torch.int32, device=fc.info.device)
        self.epoch = 0
        self.launched = self.n_units          # nothing pending
        self.rot = fc.rank * self.units_per_rank
        self.peer_arr = None
        self.tag = None                       # what is (being) gathered here, e.g. (layer, "w_qkv")

    def bind(self, shard: torch.Tensor, tag) -> None:
        assert self.launched == self.n_units, "previous epoch not fully requested (flush the queue first)"
        assert shard.shape == (self.rows_per_rank, self.cols)
        self.peer_arr = self.fc.peer_array(shard)
        self.epoch += 1
        self.launched = 0
        self.tag = tag

    @property
    def pending(self) -> int:
        return self.n_units - self.launched


class GatherQueue:
    """Pending gather work in consumption order. Every large GEMM of the ZeRO-3 step takes a slice sized to
    its own duration (`take`); a GEMM whose own B operand is not fully requested yet takes all of that
    matrix's remainder and waits on its counters tile by tile (`use`)."""
    _ATTACH = [_V, _V, _V, _L, _I, _I, _I, _I, _I, _U, _V, _I, _U, _V]

    def __init__(self, fc: "FusedComm", pull_gbs: float = 200.0, gemm_tflops: float = 1300.0):
        self.fc = fc
        self.pending: list[StagedMatrix] = []
        self.bytes_per_flop = pull_gbs * 1e9 / (gemm_tflops * 1e12)
        self.bytes_requested = 0

    def push(self, sm: StagedMatrix) -> None:
        if sm.pending and sm not in self.pending:
            self.pending.append(sm)

    def _attach(self, job, wait) -> None:
        f = _lib.fn("dlti_gemm2_attach", self._ATTACH)
        if job is not None:
            sm, j0, n = job
            self.bytes_requested += n * sm.unit_bytes * (self.fc.world - 1) // self.fc.world
            ja = (sm.peer_arr, _lib.ptr(sm.staging), _lib.ptr(sm.counters), sm.row_bytes, sm.rows_per_rank,
                  self.fc.world, sm.rot, j0, n, sm.epoch)
        else:
            ja = (None, None, None, 0, 0, 0, 0, 0, 0, 0)
        wa = (_lib.ptr(wait.counters), wait.n_units, wait.epoch) if wait is not None else (None, 0, 0)
        _lib.check(f(*ja, *wa, _lib.ptr(self.fc.err)), "gemm2_attach")

    def attach(self, consumed: StagedMatrix | None, flops: float) -> None:
        """Call immediately before launching a CTA-pair GEMM whose B operand is `consumed.staging` (or a GEMM
        with a plain B when consumed is None): decides what the launch's gather warps pull."""
        if consumed is not None and consumed.pending:
            # cold start: this launch pulls the rest of its own operand and waits for it tile by tile
            job = (consumed, consumed.launched, consumed.pending)
            consumed.launched = consumed.n_units
            if consumed in self.pending:
                self.pending.remove(consumed)
            self._attach(job, consumed)
            return
        if not self.pending:
            return
        sm = self.pending[0]
        n = max(1, min(sm.pending, int(flops * self.bytes_per_flop / sm.unit_bytes)))
        job = (sm, sm.launched, n)
        sm.launched += n
        if not sm.pending:
            self.pending.pop(0)
        self._attach(job, None)

    def flush(self, sm: StagedMatrix | None = None) -> None:
        """Run what is still pending (of `sm`, or of everything) as stand-alone gather kernels."""
        todo = [m for m in self.pending if sm is None or m is sm]
        for m in todo:
            self._attach((m, m.launched, m.pending), None)
            m.launched = m.n_units
            self.pending.remove(m)
            _lib.check(_lib.fn("dlti_gather_job_run", [_I, _V])(0, _lib.stream_ptr()), "gather_job_run")


class PeerRows:
    """A row-sharded table addressed in place on its owner ranks (K15): `ops.elementwise.embedding` looks
    tokens up with peer loads instead of gathering the table first."""

    def __init__(self, fc: "FusedComm", shard: torch.Tensor, rows: int):
        self.fc, self.shard, self.rows = fc, shard, rows
        self.peer_arr = fc.peer_array(shard)
        self.shape = (rows, shard.shape[1])
        self.dtype, self.device = shard.dtype, shard.device

    def lookup(self, ids: torch.Tensor) -> torch.Tensor:
        assert ids.dtype == torch.int64 and ids.is_contiguous()
        d = self.shard.shape[1]
        out = torch.empty((ids.numel(), d), dtype=self.dtype, device=self.device)
        f = _lib.fn("dlti_embedding_peer", [_V, _V, _V, _L, _I, _L, _L, _I, _I, _V])
        _lib.check(f(_lib.ptr(ids), self.peer_arr, _lib.ptr(out), ids.numel(), d, self.shard.shape[0], self.rows,
                     self.fc.world, _lib.dtype_code(self.dtype), _lib.stream_ptr()), "embedding_peer")
        return out.view(*ids.shape, d)


class ActGather:
    """K11 handle (`gather=` of the GEMM ops): the A operand of the launch is `full` [world * per, d], whose rows
    live on their owner ranks (`own`, a symmetric [per_max, d] buffer). The launch's gather warp pulls them unit by
    unit, own rows first; the A producer waits on fresh unit counters (zeroed pool slice, target = one epoch)."""
    sm = None

    def __init__(self, fc: "FusedComm", own: torch.Tensor, full: torch.Tensor, counters: torch.Tensor, per: int):
        self.fc, self.own, self.full, self.counters, self.per = fc, own, full, counters, per
        self.n_units = fc.world * per // 64
        self.done = False

    def _attach(self, wait: bool) -> None:
        fc = self.fc
        f = _lib.fn("dlti_gemm2_attach", GatherQueue._ATTACH)
        row_bytes = self.full.shape[1] * self.full.element_size()
        _lib.check(f(fc.peer_array(self.own), _lib.ptr(self.full), _lib.ptr(self.counters), row_bytes, self.per,
                     fc.world, fc.rank * (self.per // 64), 0, self.n_units, 1,
                     _lib.ptr(self.counters) if wait else None, self.n_units if wait else 0, 1, _lib.ptr(fc.err)),
                   "gemm2_attach")

    def attach(self, flops: float) -> None:
        if self.done:           # a second GEMM on the same activations: already complete in stream order
            return
        self._attach(True)
        _lib.check(_lib.fn("dlti_gemm2_attach_a_side", [_I])((self.fc.rank * self.per) // 256), "gemm2_attach_a_side")
        self.done = True

    def materialize(self) -> None:
        """Pull the rows with the stand-alone gather kernel (no GEMM follows, e.g. before the LM head)."""
        if not self.done:
            self._attach(False)
            _lib.check(_lib.fn("dlti_gather_job_run", [_I, _V])(0, _lib.stream_ptr()), "gather_job_run")
            self.done = True


class GatherUse:
    """Handle a model passes as `gather=` to the GEMM ops: `b` of that GEMM is `sm.staging` (K1)."""

    def __init__(self, sm: StagedMatrix | None, queue: GatherQueue):
        self.sm, self.queue = sm, queue

    def attach(self, flops: float) -> None:
        self.queue.attach(self.sm, flops)


class FusedComm:
    N_CHANNELS = 32
    MAX_GRAD_RANGES = 256

    def __init__(self, info: DistInfo):
        assert info.distributed and info.device.type == "cuda"
        if info.world_size > MAX_RANKS:
            raise ValueError(f"fused backend supports up to {MAX_RANKS} ranks per node")
        self.info = info
        self.rank, self.world = info.rank, info.world_size
        self.arena = SymmetricArena(info)
        self._ptrs: dict[int, list[int]] = {}
        # signal words [channels, MAX_RANKS] u32 and the norm board [MAX_RANKS] f32
        self.pads = self.alloc_symmetric((self.N_CHANNELS * MAX_RANKS,), torch.int32)
        self.board = self.alloc_symmetric((2 * MAX_RANKS,), torch.float32)
        # K9 overlapped form: "gradients of range s are final" flags, one row of MAX_RANKS words per range
        self.grad_flags = self.alloc_symmetric((self.MAX_GRAD_RANGES * MAX_RANKS,), torch.int32)
        self.pads.zero_()
        self.board.zero_()
        self.grad_flags.zero_()
        self.k9_epoch = 0
        self._k9_done = torch.zeros(1, dtype=torch.int32, device=info.device)
        self._k9_local2 = torch.zeros(2, dtype=torch.float32, device=info.device)
        self.side_stream = torch.cuda.Stream(device=info.device)
        self.err = torch.zeros(1, dtype=torch.int32, device=info.device)
        self.epochs = [0] * self.N_CHANNELS
        self._tmp = torch.zeros(1, dtype=torch.float32, device=info.device)
        torch.cuda.synchronize()
        dist.barrier()

    # ---- allocation --------------------------------------------------------------------------
    def alloc_symmetric(self, shape, dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        esz = torch.empty(0, dtype=dtype).element_size()
        raw, ptrs = self.arena.alloc(max(1, n) * esz)
        t = raw[: n * esz].view(dtype).view(*shape)
        self._ptrs[t.data_ptr()] = ptrs
        return t

    def peer_ptrs(self, t: torch.Tensor) -> list[int]:
        base = self._ptrs.get(t.data_ptr())
        if base is None:
            raise ValueError("tensor was not allocated from the symmetric arena")
        return base

    def _arr(self, ptrs: list[int]):
        return (C.c_void_p * MAX_RANKS)(*ptrs, *([None] * (MAX_RANKS - len(ptrs))))

    def peer_array(self, t: torch.Tensor):
        return self._arr(self.peer_ptrs(t))

    def check_error(self) -> None:
        code = int(self.err.item())
        if code == 2:
            raise RuntimeError("fused backend: signal-pad sequence violation (debug build): a peer's barrier flag ran "
                               "more than one epoch ahead — two call sites share a channel or a rank skipped a barrier")
        if code != 0:
            raise RuntimeError("fused backend: a peer did not arrive within the spin-wait timeout")

    # ---- primitives --------------------------------------------------------------------------
    def barrier(self, channel: int = 0) -> None:
        self.epochs[channel] += 1
        f = _lib.fn("dlti_peer_barrier", [_V, _I, _I, _I, _U, _V, _V])
        _lib.check(f(self.peer_array(self.pads), self.rank, self.world, channel, self.epochs[channel],
                     _lib.ptr(self.err), _lib.stream_ptr()), "peer_barrier")

    def barrier_graph_safe(self, channel: int) -> None:
        """Barrier whose epoch lives on the device: safe to capture in / replay from a CUDA graph. Use a
        channel that is never mixed with the host-epoch `barrier()`."""
        if not hasattr(self, "_dev_epochs"):
            self._dev_epochs = torch.zeros(self.N_CHANNELS, dtype=torch.int32, device=self.info.device)
        f = _lib.fn("dlti_peer_barrier_auto", [_V, _I, _I, _I, _V, _V, _V])
        _lib.check(f(self.peer_array(self.pads), self.rank, self.world, channel,
                     C.c_void_p(self._dev_epochs.data_ptr() + 4 * channel), _lib.ptr(self.err),
                     _lib.stream_ptr()), "peer_barrier_auto")

    def all_gather_rows(self, shard: torch.Tensor, out: torch.Tensor, blocks_per_peer: int = 8) -> None:
        """out [world*rows, cols] <- row blocks of every rank (pull). shard must be symmetric."""
        nbytes = shard.numel() * shard.element_size()
        assert out.numel() * out.element_size() == nbytes * self.world and out.is_contiguous()
        f = _lib.fn("dlti_allgather_rows", [_V, _V, _L, _I, _I, _I, _V])
        _lib.check(f(self.peer_array(shard), _lib.ptr(out), nbytes, self.rank, self.world,
                     blocks_per_peer, _lib.stream_ptr()), "allgather_rows")

    def reduce_scatter_accumulate(self, g_full: torch.Tensor, grad_part: torch.Tensor) -> None:
        """grad_part += sum over ranks of their g_full[lo:hi]. g_full must be symmetric.
        barrier (all grads final) -> peer-read reduce -> barrier (peers may overwrite their grads)."""
        part = grad_part.numel()
        self.barrier(1)
        f = _lib.fn("dlti_rs_accumulate", [_V, _V, _L, _L, _I, _V])
        _lib.check(f(self.peer_array(g_full), _lib.ptr(grad_part), self.rank * part, part, self.world,
                     _lib.stream_ptr()), "rs_accumulate")
        self.barrier(2)

    def global_sumsq(self, g: torch.Tensor, out: torch.Tensor) -> None:
        from ..ops import optim as O

        O.sumsq(g, out=self._tmp)
        ch = 3
        self.epochs[ch] += 1
        f = _lib.fn("dlti_sumsq_exchange", [_V, _V, _V, _I, _I, _I, _U, _V, _V, _V])
        _lib.check(f(_lib.ptr(self._tmp), self.peer_array(self.board), self.peer_array(self.pads),
                     self.rank, self.world, ch, self.epochs[ch], _lib.ptr(out), _lib.ptr(self.err),
                     _lib.stream_ptr()), "sumsq_exchange")

    def adamw_push(self, master_full: torch.Tensor, m, v, g, lo: int, n: int, *, lr, betas, eps, wd, step,
                   gnorm_sq, inv_scale, max_norm, skipped_out) -> None:
        """AdamW on [lo, lo+n) and push the new fp32 values into every rank's master buffer; the
        trailing barrier makes them visible before anyone refreshes its compute copies."""
        f = _lib.fn("dlti_adamw_push", [_V, _V, _V, _V, _L, _L, _I, _I, _F, _F, _F, _F, _F, _I, _V, _F, _F, _V, _V])
        _lib.check(f(self.peer_array(master_full), _lib.ptr(m), _lib.ptr(v), _lib.ptr(g), lo, n, self.rank,
                     self.world, lr, betas[0], betas[1], eps, wd, int(step), _lib.ptr(gnorm_sq),
                     float(inv_scale), float(max_norm), _lib.ptr(skipped_out), _lib.stream_ptr()), "adamw_push")
        self.barrier(4)

    # ---- K9, overlapped form -----------------------------------------------------------------------
    def k9_begin_step(self) -> None:
        """New optimizer step: the epoch that the per-range flags and the step kernel's channels carry."""
        self.k9_epoch += 1

    def grad_range_final(self, slot: int, g_full: torch.Tensor, grad_part: torch.Tensor, glo: int, n: int,
                         part_lo: int, part_n: int, side: bool = True, ctas: int = 16) -> None:
        """This rank's gradients of the flat range [glo, glo + n) (range number `slot`) are final for the current
        step: tell every rank (main stream, i.e. behind the kernels that produced them) and, if part of the
        range belongs to this rank's partition [part_lo, part_lo + part_n), start the reduction of that part on
        the side stream — it waits for the other ranks' flags while the main stream goes on with backward
        (side=False: on the main stream, the form used at the end of backward)."""
        assert slot < self.MAX_GRAD_RANGES
        f = _lib.fn("dlti_grad_ready_signal", [_V, _I, _I, _I, _U, _V])
        _lib.check(f(self.peer_array(self.grad_flags), self.rank, self.world, slot, self.k9_epoch,
                     _lib.stream_ptr()), "grad_ready_signal")
        a, b = max(glo, part_lo), min(glo + n, part_lo + part_n)
        if b <= a:
            return
        with torch.cuda.stream(self.side_stream if side else torch.cuda.current_stream()):
            f = _lib.fn("dlti_rs_wait_accumulate", [_V, _V, _V, _L, _L, _L, _I, _I, _U, _I, _V, _V])
            _lib.check(f(self.peer_array(g_full), _lib.ptr(self.grad_flags), _lib.ptr(grad_part), a, a - part_lo, b - a,
                         self.world, slot, self.k9_epoch, ctas, _lib.ptr(self.err), _lib.stream_ptr()), "rs_wait_accumulate")

    def k9_step(self, master_full: torch.Tensor, m, v, g, lo: int, n: int, *, lr, betas, eps, wd, step, loss_sum,
                out2: torch.Tensor, inv_scale, max_norm, skipped_out) -> None:
        """After backward: joins the side stream, then sum of squares of the reduced partition -> one kernel that
        exchanges {norm, loss}, steps AdamW on the partition and pushes the new values to every rank -> wait for
        everybody's pushes. out2 = [global sum of squares, global loss sum]. The caller recasts the adapters."""
        from ..ops import optim as O

        torch.cuda.current_stream().wait_stream(self.side_stream)
        O.sumsq(g, out=self._k9_local2[0:1])
        self._k9_local2[1:2].copy_(loss_sum.reshape(1))
        f = _lib.fn("dlti_k9_step", [_V, _V, _V, _V, _V, _V, _L, _L, _I, _I, _F, _F, _F, _F, _F, _I, _V, _V, _F, _F, _V,
                                     _V, _I, _I, _U, _V, _V])
        _lib.check(f(self.peer_array(master_full), self.peer_array(self.board), self.peer_array(self.pads),
                     _lib.ptr(m), _lib.ptr(v), _lib.ptr(g), lo, n, self.rank, self.world, lr, betas[0], betas[1], eps,
                     wd, int(step), _lib.ptr(self._k9_local2), _lib.ptr(out2), float(inv_scale), float(max_norm),
                     _lib.ptr(skipped_out), _lib.ptr(self._k9_done), 5, 6, self.k9_epoch, _lib.ptr(self.err),
                     _lib.stream_ptr()), "k9_step")
        f = _lib.fn("dlti_peer_wait", [_V, _I, _I, _U, _V, _V])
        _lib.check(f(_lib.ptr(self.pads), self.world, 6, self.k9_epoch, _lib.ptr(self.err), _lib.stream_ptr()), "peer_wait")

    # ---- fused all-gather -> GEMM ----------------------------------------------------------------
    def gather_spec(self, shard: torch.Tensor, staging: torch.Tensor, counters: torch.Tensor) -> GatherSpec:
        """counters: int32[world*rows/64], zero-initialised once and shared by every shard that is
        gathered into the same staging buffer (epochs keep growing)."""
        rows = shard.shape[0]
        assert rows % 64 == 0, "fused gather needs rows_per_rank % 64 == 0"
        assert counters.numel() == rows * self.world // 64
        return GatherSpec(self.peer_ptrs(shard), staging, counters, rows, self.rank, self.world, self.err)


def tp_reduce_add_norm(recv: torch.Tensor, T: int, residual: torch.Tensor, w: torch.Tensor, eps: float,
                       world: int, max_tokens: int, barrier=None) -> torch.Tensor:
    """residual[:T] += sum over ranks of recv[r, :T] (fixed order); returns RMSNorm(residual) * w.
    recv: [world, max_tokens, d] receive slots filled by every rank's GEMM epilogue (K10).
    barrier = (FusedComm, channel): the kernel itself performs the rank barrier that separates the pushes from
    the reduction (no separate barrier launch)."""
    d = residual.shape[-1]
    y = torch.empty((T, d), dtype=residual.dtype, device=residual.device)
    if barrier is not None:
        fc, channel = barrier
        if not hasattr(fc, "_dev_epochs"):
            fc._dev_epochs = torch.zeros(fc.N_CHANNELS, dtype=torch.int32, device=fc.info.device)
        if not hasattr(fc, "_tp_done"):
            fc._tp_done = torch.zeros(1, dtype=torch.int32, device=fc.info.device)
        f = _lib.fn("dlti_tp_barrier_reduce_add_norm", [_V, _L, _I, _V, _V, _V, _L, _I, _F, _I, _V, _I, _I, _V, _V, _V, _V])
        _lib.check(f(_lib.ptr(recv), max_tokens * d, world, _lib.ptr(residual), _lib.ptr(w), _lib.ptr(y), T, d,
                     float(eps), _lib.dtype_code(residual.dtype), fc.peer_array(fc.pads), fc.rank, channel,
                     C.c_void_p(fc._dev_epochs.data_ptr() + 4 * channel), _lib.ptr(fc._tp_done), _lib.ptr(fc.err),
                     _lib.stream_ptr()), "tp_barrier_reduce_add_norm")
        return y
    f = _lib.fn("dlti_tp_reduce_add_norm", [_V, _L, _I, _V, _V, _V, _L, _I, _F, _I, _V])
    _lib.check(f(_lib.ptr(recv), max_tokens * d, world, _lib.ptr(residual), _lib.ptr(w), _lib.ptr(y), T, d,
                 float(eps), _lib.dtype_code(residual.dtype), _lib.stream_ptr()), "tp_reduce_add_norm")
    return y


def tp_own_reduce_norm(fc, slots: torch.Tensor, per: int, T: int, residual: torch.Tensor, w: torch.Tensor, eps: float,
                       xn_own: torch.Tensor) -> None:
    """Owner side of the prefill-sized row-parallel all-reduce: this rank's rows [rank * per, ...) of the step."""
    d = residual.shape[-1]
    lo = min(T, fc.rank * per)
    rows = min(T, lo + per) - lo
    f = _lib.fn("dlti_tp_own_reduce_norm", [_V, _I, _L, _L, _L, _V, _V, _V, _I, _F, _I, _V])
    _lib.check(f(_lib.ptr(slots), fc.world, per, lo, rows, _lib.ptr(residual), _lib.ptr(w), _lib.ptr(xn_own), d,
                 float(eps), _lib.dtype_code(residual.dtype), _lib.stream_ptr()), "tp_own_reduce_norm")


def tp_pull_reduce_norm_push(fc, partial: torch.Tensor, xn: torch.Tensor, T: int, residual: torch.Tensor,
                             w: torch.Tensor, eps: float) -> None:
    """This rank's share of a prefill-sized row-parallel all-reduce (+ residual add + RMSNorm): rows
    [rank*ceil(T/world), ...) of sum_r partial_r are reduced by pulling from the peers and the normalised rows
    are pushed into every rank's `xn`. partial / xn: symmetric [>= T, d] buffers. Barriers are the caller's."""
    d = residual.shape[-1]
    per = -(-T // fc.world)
    lo = min(T, fc.rank * per)
    rows = min(T, lo + per) - lo
    f = _lib.fn("dlti_tp_pull_reduce_norm_push", [_V, _V, _I, _L, _L, _V, _V, _I, _F, _I, _V])
    _lib.check(f(fc.peer_array(partial), fc.peer_array(xn), fc.world, lo, rows, _lib.ptr(residual), _lib.ptr(w), d,
                 float(eps), _lib.dtype_code(residual.dtype), _lib.stream_ptr()), "tp_pull_reduce_norm_push")
