"""ZeRO stage-3 partitioning of the FROZEN base weights (collective C4).

Every matrix [R, C] is split by rows: rank r owns rows [r*Rp, (r+1)*Rp), Rp = ceil(R/N) (zero
padded). Row-major row blocks mean the gathered matrix is simply the rank-ordered concatenation,
so a gather lands directly in the layout the GEMM consumes — and a fused kernel can address
"the rows of tile n" on their owner without any index table.

Parameters with numel <= stage3_param_persistence_threshold stay replicated (for Llama that is
exactly the RMSNorm weights; every LoRA tensor is also below the threshold — SURVEY D14).

Gather scheduling mirrors DeepSpeed's coordinator at decoder-layer granularity: `prefetch(i)`
starts the gather of layer i on a side stream into one of `n_slots` full-layer staging buffers
(n_slots bounded by stage3_max_live_parameters), `layer(i)` waits for it, `release(i)` frees the
slot. The look-ahead is one whole layer (DeepSpeed's stage3_prefetch_bucket_size counts elements;
one 7B layer is 2e8 elements — we round the window up to a layer so forward GEMMs never wait).

Backends: "nccl" = all_gather_into_tensor per matrix (baseline; gloo on CPU);
          "fused" = peer-pull sm_100a copy kernel over symmetric memory (parallel.fused).
offload_param = "cpu": shards live in pinned host memory and are copied H2D before each gather.
"""
from __future__ import annotations

import os

import torch

from ..config.ds_config import DSConfig
from ..models.config import LlamaConfig
from ..models.llama import LayerWeights
from . import comm
from .comm import DistInfo

_MATS = ("w_qkv", "w_o", "w_gu", "w_down")


def _rows_per_rank(rows: int, world: int, align: int = 1) -> int:
    rp = -(-rows // world)
    return -(-rp // align) * align


class Zero3Weights:
    def __init__(self, cfg: LlamaConfig, ds: DSConfig, info: DistInfo, dtype: torch.dtype,
                 backend: str = "nccl", fused=None):
        self.cfg, self.ds, self.info, self.dtype = cfg, ds, info, dtype
        self.device = info.device
        self.world, self.rank = info.world_size, info.rank
        self.backend, self.fused = backend, fused
        d, ff = cfg.hidden_size, cfg.intermediate_size
        self.shapes = {"w_qkv": (cfg.qkv_size, d), "w_o": (d, cfg.q_size), "w_gu": (2 * ff, d),
                       "w_down": (d, ff)}
        self.offload = ds.offload_param == "cpu"
        # the fused gather works in 64-row units (one UMMA K block / a quarter N tile)
        self.row_align = 64 if (fused is not None and not self.offload) else 1
        self.persist_thr = ds.stage3_param_persistence_threshold
        sdev = "cpu" if self.offload else self.device
        pin = self.offload and ds.pin_memory and torch.cuda.is_available()
        L = cfg.num_hidden_layers

        def shard_buf(rows, cols):
            rp = _rows_per_rank(rows, self.world, self.row_align)
            if fused is not None and not self.offload:
                return fused.alloc_symmetric((rp, cols), dtype)
            t = torch.zeros(rp, cols, dtype=dtype, device=sdev)
            return t.pin_memory() if pin else t

        self.layer_shards = [{m: shard_buf(*self.shapes[m]) for m in _MATS} for _ in range(L)]
        self.norms = [(torch.ones(d, dtype=dtype, device=self.device),
                       torch.ones(d, dtype=dtype, device=self.device)) for _ in range(L)]
        self.final_norm = torch.ones(d, dtype=dtype, device=self.device)
        self.embed_shard = shard_buf(cfg.vocab_size, d)
        self.lm_head_shard = self.embed_shard if cfg.tie_word_embeddings else shard_buf(cfg.vocab_size, d)
        # staging: full-layer slots
        layer_elems = sum(r * c for r, c in self.shapes.values())
        max_live = max(layer_elems, ds.stage3_max_live_parameters)
        self.n_slots = max(1, min(3, max_live // layer_elems)) if self.world > 1 else 0
        self.slots = []
        for _ in range(self.n_slots):
            self.slots.append({m: torch.empty(_rows_per_rank(r, self.world, self.row_align) * self.world, c, dtype=dtype,
                                              device=self.device) for m, (r, c) in self.shapes.items()})
        self.slot_of: dict[int, int] = {}
        self.slot_free_event: list = [None] * self.n_slots
        self.slot_ready_event: list = [None] * self.n_slots
        self.free_slots = list(range(self.n_slots))
        self.stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" and self.world > 1 else None
        # K1 mode: one staging buffer per matrix kind, filled by the consuming GEMM kernel itself
        # DLTI_Z3_GATHER=prefetch: pull the next layer with the stand-alone peer-memory all-gather kernel on a side
        # stream (every GEMM then runs the plain kernels); default: gather inside the first GEMM that needs it (K1)
        self.in_gemm_gather = (fused is not None and not self.offload and self.world > 1
                               and os.environ.get("DLTI_Z3_GATHER", "gemm") != "prefetch")
        if self.in_gemm_gather:
            # Two full-layer staging slots (the layer in use and the one being pulled). prefetch(i) only QUEUES
            # layer i's matrices; the queue is drained by the gather warps of the GEMMs that run meanwhile, a
            # slice per launch sized to the launch's duration (parallel.fused.GatherQueue). The GEMMs of layer i
            # then find their operands complete; whatever is not (first layer of a run) is pulled by the
            # consuming launch itself, which waits on the unit counters tile by tile.
            from .fused import GatherQueue, GatherUse, StagedMatrix
            self.slots, self.n_slots, self.free_slots = [], 0, []
            self.queue = GatherQueue(fused, pull_gbs=float(os.environ.get("DLTI_Z3_PULL_GBS", "200")))
            self.stage_slots = [{m: StagedMatrix(fused, _rows_per_rank(r, self.world, self.row_align), c, dtype)
                                 for m, (r, c) in self.shapes.items()} for _ in range(2)]
            self.slot_layer = [None, None]
            self._GatherUse = GatherUse
            self.head_stage = StagedMatrix(fused, self.lm_head_shard.shape[0], d, dtype)
            self.head_step = -1
        self._h2d_live: dict = {}   # offload on one rank: layer -> (device buffer index, copy-done event)
        self._big = {}          # gathered embed / lm_head staging (allocated on demand)
        self._h2d = {}          # offload: device copies of this rank's shards while in flight
        self.gather_bytes = 0   # bytes received from peers (for roofline reports)

    # ---- initialisation ------------------------------------------------------------------------
    def init_random(self, seed: int = 0) -> None:
        """Random-init (std 0.02). Every matrix is generated from a generator seeded by
        (seed, matrix id) — independent of the world size, so N-rank and 1-rank runs see the same
        model — on the compute device (a 7B layer matrix is a ~180 MB temporary), and only this
        rank's row block is kept."""
        gdev = self.device
        gen = torch.Generator(device=gdev)

        def fill(dst, rows_total, cols, key):
            gen.manual_seed((seed * 1000003 + key) & 0x7FFFFFFFFFFF)
            full = torch.empty(rows_total, cols, dtype=torch.float32 if gdev.type == "cpu" else self.dtype,
                               device=gdev)
            full.normal_(0.0, 0.02, generator=gen)
            rp = dst.shape[0]
            lo = self.rank * rp
            hi = min(rows_total, lo + rp)
            dst.zero_()
            if hi > lo:
                dst[: hi - lo].copy_(full[lo:hi])

        for li, sh in enumerate(self.layer_shards):
            for mi, m in enumerate(_MATS):
                fill(sh[m], *self.shapes[m], key=li * 8 + mi + 16)
        fill(self.embed_shard, self.cfg.vocab_size, self.cfg.hidden_size, key=1)
        if self.lm_head_shard is not self.embed_shard:
            fill(self.lm_head_shard, self.cfg.vocab_size, self.cfg.hidden_size, key=2)

    def load_from_resident(self, w) -> None:
        """Shard a fully materialised ResidentWeights (tests, small models)."""
        def take(dst, full):
            rp = dst.shape[0]
            lo = self.rank * rp
            hi = min(full.shape[0], lo + rp)
            dst.zero_()
            if hi > lo:
                dst[: hi - lo].copy_(full[lo:hi])
        for i, sh in enumerate(self.layer_shards):
            lw = w.layers[i]
            for m in _MATS:
                take(sh[m], getattr(lw, m))
            self.norms[i][0].copy_(lw.ln1)
            self.norms[i][1].copy_(lw.ln2)
        self.final_norm.copy_(w.final_norm)
        take(self.embed_shard, w.embed)
        if self.lm_head_shard is not self.embed_shard:
            take(self.lm_head_shard, w.lm_head)

    def load_from_hf(self, directory) -> None:
        """Each rank reads only the byte ranges of its own row blocks (fixes App. D #6)."""
        from ..models.hf_io import CheckpointReader, hf_layer_names

        rd = CheckpointReader(directory)
        cfg = self.cfg
        q, kv, ff = cfg.q_size, cfg.kv_size, cfg.intermediate_size

        def take(dst, pieces):
            """pieces: list of (name, rows) stacked along dim 0 to form the logical matrix."""
            rp = dst.shape[0]
            lo, hi = self.rank * rp, (self.rank + 1) * rp
            dst.zero_()
            base = 0
            for name, rows in pieces:
                a, b = max(lo, base), min(hi, base + rows)
                if b > a:
                    dst[a - lo:b - lo].copy_(rd.get(name, rows=(a - base, b - base)).to(self.dtype))
                base += rows

        for i, sh in enumerate(self.layer_shards):
            n = hf_layer_names(i)
            take(sh["w_qkv"], [(n["q"], q), (n["k"], kv), (n["v"], kv)])
            take(sh["w_o"], [(n["o"], cfg.hidden_size)])
            take(sh["w_gu"], [(n["gate"], ff), (n["up"], ff)])
            take(sh["w_down"], [(n["down"], cfg.hidden_size)])
            self.norms[i][0].copy_(rd.get(n["ln1"]).to(self.dtype))
            self.norms[i][1].copy_(rd.get(n["ln2"]).to(self.dtype))
        self.final_norm.copy_(rd.get("model.norm.weight").to(self.dtype))
        take(self.embed_shard, [("model.embed_tokens.weight", cfg.vocab_size)])
        if self.lm_head_shard is not self.embed_shard:
            nm = "lm_head.weight" if rd.has("lm_head.weight") else "model.embed_tokens.weight"
            take(self.lm_head_shard, [(nm, cfg.vocab_size)])

    # ---- gathers -------------------------------------------------------------------------------
    def _device_shard(self, key, shard: torch.Tensor) -> torch.Tensor:
        if not self.offload:
            return shard
        t = shard.to(self.device, non_blocking=True)       # H2D of the local shard (D16)
        self._h2d[key] = t
        return t

    def _gather(self, key, shard: torch.Tensor, out: torch.Tensor) -> None:
        src = self._device_shard(key, shard)
        if self.fused is not None and not self.offload:
            self.fused.all_gather_rows(shard, out, blocks_per_peer=int(os.environ.get("DLTI_AG_BLOCKS", max(8, 96 // self.world))))
        else:
            comm.all_gather_into(out.view(-1), src.reshape(-1), self.info)
        self.gather_bytes += out.numel() * out.element_size() * (self.world - 1) // self.world

    def _queue_layer(self, i: int) -> int:
        """Fused K1 mode: make layer i resident-or-queued in a staging slot; returns the slot."""
        if i in self.slot_layer:
            return self.slot_layer.index(i)
        # evict the slot whose layer is farthest from i (an empty slot first): in either direction of travel
        # that is the layer whose GEMMs have all been launched, never the one about to be used
        dist_ = [10 ** 9 if l is None else abs(l - i) for l in self.slot_layer]
        s = 0 if dist_[0] >= dist_[1] else 1
        for sm in self.stage_slots[s].values():    # leftovers of a layer that was queued but never used
            if sm.pending:
                self.queue.flush(sm)
        for m in _MATS:
            sm = self.stage_slots[s][m]
            sm.bind(self.layer_shards[i][m], (i, m))
            self.queue.push(sm)
            self.gather_bytes += sm.n_units * sm.unit_bytes * (self.world - 1) // self.world
        self.slot_layer[s] = i
        return s

    def _h2d_layer(self, i: int) -> None:
        """ZeRO-Offload of parameters on ONE rank (D16): start the host-to-device copy of layer i's (pinned)
        matrices on a copy stream into one of two device-resident layer buffers, one layer ahead of its use, so
        that the PCIe transfer of layer i+1 overlaps the compute of layer i. Buffers are recycled only after the
        compute stream has finished with them (events)."""
        if i in self._h2d_live or not (0 <= i < self.cfg.num_hidden_layers):
            return
        if getattr(self, "_h2d_stream", None) is None:
            self._h2d_stream = torch.cuda.Stream(device=self.device)
            self._h2d_bufs = [{m: torch.empty_like(self.layer_shards[0][m], device=self.device) for m in _MATS}
                              for _ in range(2)]
            self._h2d_free_ev = [None, None]
        # recycle the buffer whose layer is farthest from i (an empty one first): in either direction of travel that
        # is the layer whose kernels have all been launched, never the one about to be used
        tenant = [next((k for k, v in self._h2d_live.items() if v[0] == b), None) for b in (0, 1)]
        far = [10 ** 9 if t is None else abs(t - i) for t in tenant]
        b = 0 if far[0] >= far[1] else 1
        if tenant[b] is not None:
            del self._h2d_live[tenant[b]]
        st = self._h2d_stream
        if self._h2d_free_ev[b] is not None:
            st.wait_event(self._h2d_free_ev[b])
        with torch.cuda.stream(st):
            for m in _MATS:
                self._h2d_bufs[b][m].copy_(self.layer_shards[i][m], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(st)
        self._h2d_live[i] = (b, ev)

    def prefetch(self, i: int) -> None:
        if self.world == 1 and self.offload and self.device.type == "cuda":
            self._h2d_layer(i)
            return
        if self.in_gemm_gather and 0 <= i < self.cfg.num_hidden_layers:
            self._queue_layer(i)
            return
        if self.world == 1 or self.in_gemm_gather or i in self.slot_of or i < 0 or i >= self.cfg.num_hidden_layers:
            return
        if not self.free_slots:
            return                                   # window full (stage3_max_live_parameters)
        s = self.free_slots.pop(0)
        self.slot_of[i] = s
        st = self.stream
        if st is not None:
            if self.slot_free_event[s] is not None:
                st.wait_event(self.slot_free_event[s])
            with torch.cuda.stream(st):
                for m in _MATS:
                    self._gather((i, m), self.layer_shards[i][m], self.slots[s][m])
                ev = torch.cuda.Event()
                ev.record(st)
            self.slot_ready_event[s] = ev
        else:
            for m in _MATS:
                self._gather((i, m), self.layer_shards[i][m], self.slots[s][m])

    def layer(self, i: int) -> LayerWeights:
        if self.world == 1 and self.offload and self.device.type == "cuda":
            self._h2d_layer(i)                      # no-op when prefetch(i) already started it
            b, ev = self._h2d_live[i]
            torch.cuda.current_stream().wait_event(ev)
            r = {m: self._h2d_bufs[b][m][: self.shapes[m][0]] for m in _MATS}
            return LayerWeights(r["w_qkv"], r["w_o"], r["w_gu"], r["w_down"], *self.norms[i])
        if self.world == 1:
            sh = self.layer_shards[i]
            dev = (lambda t: t.to(self.device, non_blocking=True)) if self.offload else (lambda t: t)
            r = {m: dev(sh[m])[: self.shapes[m][0]] for m in _MATS}
            return LayerWeights(r["w_qkv"], r["w_o"], r["w_gu"], r["w_down"], *self.norms[i])
        if self.in_gemm_gather:
            s = self._queue_layer(i)
            st = self.stage_slots[s]
            g = {m: self._GatherUse(st[m], self.queue) for m in _MATS}
            return LayerWeights(st["w_qkv"].staging[: self.shapes["w_qkv"][0]], st["w_o"].staging[: self.shapes["w_o"][0]],
                                st["w_gu"].staging[: self.shapes["w_gu"][0]], st["w_down"].staging[: self.shapes["w_down"][0]],
                                *self.norms[i], gather=g)
        if i not in self.slot_of:
            self.prefetch(i)
            if i not in self.slot_of:
                raise RuntimeError("ZeRO-3: no free gather slot (release() not called?)")
        s = self.slot_of[i]
        if self.slot_ready_event[s] is not None:
            torch.cuda.current_stream().wait_event(self.slot_ready_event[s])
        sl = self.slots[s]
        r = {m: sl[m][: self.shapes[m][0]] for m in _MATS}
        return LayerWeights(r["w_qkv"], r["w_o"], r["w_gu"], r["w_down"], *self.norms[i])

    def release(self, i: int) -> None:
        if self.world == 1 and self.offload and i in self._h2d_live:
            ev = torch.cuda.Event()                 # the compute stream is done with this layer's buffer from here on
            ev.record(torch.cuda.current_stream())
            self._h2d_free_ev[self._h2d_live[i][0]] = ev
            return
        if self.world == 1 or self.in_gemm_gather or i not in self.slot_of:
            return
        s = self.slot_of.pop(i)
        if self.stream is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self.slot_free_event[s] = ev
        for m in _MATS:
            self._h2d.pop((i, m), None)
        self.free_slots.append(s)

    def _gather_big(self, name: str, shard: torch.Tensor) -> torch.Tensor:
        rows = self.cfg.vocab_size
        if self.world == 1:
            return (shard.to(self.device, non_blocking=True) if self.offload else shard)[:rows]
        buf = self._big.get(name)
        if buf is None:
            buf = torch.empty(shard.shape[0] * self.world, shard.shape[1], dtype=self.dtype, device=self.device)
            self._big[name] = buf
        self._gather((name,), shard, buf)
        return buf[:rows]

    def get_embed(self):
        if self.in_gemm_gather:
            # vocabulary-sharded table: the embedding kernel fetches each token's row from its owner rank
            from .fused import PeerRows
            return PeerRows(self.fused, self.embed_shard, self.cfg.vocab_size)
        return self._gather_big("embed", self.embed_shard)

    def prefetch_head(self) -> None:
        """Fused K1 mode: queue the LM head behind the layers already queued (call before the last layer's
        forward GEMMs so that they carry it). Re-gathered once per optimizer step."""
        if self.in_gemm_gather and self.head_stage.launched == self.head_stage.n_units and not self._head_valid:
            self.head_stage.bind(self.lm_head_shard, ("lm_head",))
            self.queue.push(self.head_stage)
            self.gather_bytes += self.head_stage.n_units * self.head_stage.unit_bytes * (self.world - 1) // self.world
            self._head_valid = True

    def release_head(self) -> None:
        self._head_valid = False

    _head_valid = False

    def get_lm_head(self) -> torch.Tensor:
        if self.in_gemm_gather:
            self.prefetch_head()
            return self.head_stage.staging[: self.cfg.vocab_size]
        if self.lm_head_shard is self.embed_shard and "embed" in self._big:
            return self._big["embed"][: self.cfg.vocab_size]
        return self._gather_big("lm_head", self.lm_head_shard)

    def head_gather(self):
        """`gather=` handle for GEMMs whose B operand is the LM head (None outside the fused K1 mode)."""
        return self._GatherUse(self.head_stage, self.queue) if self.in_gemm_gather else None

    def carry(self):
        """`gather=` handle for a large GEMM with a plain B operand: lets it carry queued weight pulls."""
        return self._GatherUse(None, self.queue) if self.in_gemm_gather else None

    def get_final_norm(self) -> torch.Tensor:
        return self.final_norm

    def nbytes_local(self) -> int:
        n = sum(t.numel() * t.element_size() for sh in self.layer_shards for t in sh.values())
        n += self.embed_shard.numel() * self.embed_shard.element_size()
        if self.lm_head_shard is not self.embed_shard:
            n += self.lm_head_shard.numel() * self.lm_head_shard.element_size()
        return n
