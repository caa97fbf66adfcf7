"""Serving engine: request queue -> continuous-batching scheduler -> model runner -> sampler ->
incremental detokenisation (call stack §3.4).

One engine object per rank. Rank 0 is the driver: it owns the scheduler / block manager and, each step,
broadcasts the step's metadata (one message through a shared-memory ring,
N10: csrc/shm_ring.cpp) to the TP workers, which run `worker_loop()` and execute the same forward and the same sampling.
A step is one forward pass over [decode rows | prompt rows]; pure-decode steps replay CUDA graphs
captured per batch-size bucket (S12; the collectives inside are graph-safe: device-side barrier epochs).
By default steps are pipelined: step N+1 is launched before the host has read step N's tokens, which stay
on the device and feed the next step's decode rows (`LLMEngine.step`). vLLM counterparts of the pieces:
scheduler + block manager (SP:vllm/v1/core/sched/scheduler.py), async scheduling, CUDA-graph dispatch.
"""
from __future__ import annotations

import time
import uuid

import numpy as np
import torch

from ..models.config import LlamaConfig
from ..ops import serving as S
from .kv_cache import BlockManager, KVCache
from .model_runner import ModelRunner, TPContext, TPWeights
from .params import Request, SamplingParams, StepOutput
from .scheduler import Batch, Scheduler
from .tokenizer import IncrementalDetokenizer

_BUCKETS = [1, 2, 4, 8, 16, 32, 64, 128, 256, 512]


class StepLayout:
    """Where each per-step array lives inside ONE int64 staging buffer, so a step's metadata reaches the
    device (and the TP workers) as a single copy. T token rows = nd decode rows followed by the rows of npf
    packed prompts; B = nd + npf rows are sampled; block tables are [nd, mb]; src[t] >= 0 means "the input
    token of row t is the token sampled at row src[t] of the previous step" (it is still on the device).
    64-bit sections first, then the 32-bit ones (each padded to a whole word); the tables come last so that
    a prefix copy carries everything but unused table rows."""
    _I64 = ("ids", "slots", "last", "seeds", "pslots", "pdst", "csrc")
    _I32 = ("pos", "src", "ctx", "cu", "ecu", "topk", "temp", "topp", "tables")

    def __init__(self, T: int, nd: int, npf: int, mb: int, ns: int | None = None, P: int = 0):
        """ns: prompts that are sampled in this step (default: all npf; fewer under chunked prefill, where a
        prompt's earlier chunks only fill the cache). P > 0: some prompt rows continue a prompt whose first P
        tokens (over all such prompts) are already in the paged cache: pslots[P] = their cache slots,
        pdst[P] / csrc[T - nd] = where prefix tokens / this step's prompt rows sit in the extended
        [prefix | chunk] attention layout, ecu[npf + 1] = its sequence boundaries."""
        ns = npf if ns is None else ns
        self.T, self.nd, self.npf, self.mb, self.ns, self.P = T, nd, npf, mb, ns, P
        B = nd + ns
        self.n = dict(ids=T, slots=T, last=B, seeds=B, pslots=P, pdst=P, csrc=(T - nd) if P else 0, pos=T, src=T, ctx=nd,
                      cu=npf + 1 if npf else 0, ecu=npf + 1 if P else 0, topk=B, temp=B, topp=B, tables=nd * mb)
        o = 0
        self.off = {}
        for k in self._I64:
            self.off[k] = o
            o += self.n[k]
        for k in self._I32:
            self.off[k] = o
            o += (self.n[k] + 1) // 2
        self.words = o
        self.tables = self.off["tables"]

    def views(self, buf, xp) -> dict:
        """Typed views into `buf` (1-D int64 numpy array or torch tensor; xp = numpy or torch module)."""
        i32, f32 = (np.int32, np.float32) if xp is np else (torch.int32, torch.float32)
        out = {}
        for k in self._I64:
            out[k] = buf[self.off[k]:self.off[k] + self.n[k]]
        for k in self._I32:
            n = self.n[k]
            out[k] = buf[self.off[k]:self.off[k] + (n + 1) // 2].view(f32 if k in ("temp", "topp") else i32)[:n]
        out["tables"] = out["tables"].reshape(self.nd, self.mb) if self.nd * self.mb else out["tables"]
        return out


class PackedStep:
    __slots__ = ("layout", "slot", "max_seqlen", "want_lp", "max_ext")

    def __init__(self, layout: StepLayout, slot, max_seqlen: int, max_ext: int = 0):
        self.layout, self.slot, self.max_seqlen, self.want_lp, self.max_ext = layout, slot, max_seqlen, False, max_ext


class LLMEngine:
    def __init__(self, cfg: LlamaConfig, weights: TPWeights, tp: TPContext, tokenizer, *, info=None,
                 block_size: int = 16, num_blocks: int | None = None, gpu_memory_utilization: float = 0.85,
                 max_num_seqs: int = 256, max_num_batched_tokens: int = 8192, max_model_len: int | None = None,
                 use_cuda_graphs: bool = True, mixed_batches: bool = True, async_scheduling: bool = True,
                 prefill_chunk: int = 0, enable_prefix_caching: bool = False):
        self.cfg, self.w, self.tp, self.tokenizer, self.info = cfg, weights, tp, tokenizer, info
        self.device = weights.embed.device
        self.dtype = weights.embed.dtype
        self.max_model_len = max_model_len or cfg.max_position_embeddings
        self.rank = tp.rank
        self.world = tp.world
        if getattr(tp, "use_fused", False) and prefill_chunk == 0 and self.max_model_len > max_num_batched_tokens:
            # the peer-memory TP buffers hold max_num_batched_tokens rows; without chunking a single prompt may be
            # as long as max_model_len
            raise ValueError(f"max_model_len ({self.max_model_len}) exceeds max_num_batched_tokens "
                             f"({max_num_batched_tokens}): raise the latter or enable chunked prefill")
        if num_blocks is None:
            per_block = KVCache.bytes_per_block(cfg.num_hidden_layers, block_size, weights.KV, cfg.head_dim, self.dtype)
            if self.device.type == "cuda":
                free, total = torch.cuda.mem_get_info(self.device)
                budget = max(0, free - (1 - gpu_memory_utilization) * total - (2 << 30))
                num_blocks = max(64, int(budget // per_block))
                if self.world > 1:                       # every rank must agree
                    import torch.distributed as dist
                    t = torch.tensor([num_blocks], dtype=torch.int64)
                    dist.all_reduce(t, op=dist.ReduceOp.MIN)
                    num_blocks = int(t.item())
            else:
                num_blocks = 256
        self.kv = KVCache(cfg.num_hidden_layers, num_blocks, block_size, weights.KV, cfg.head_dim, self.dtype, self.device)
        self.blocks = BlockManager(num_blocks, block_size, enable_prefix_caching)
        self.sched = Scheduler(self.blocks, max_num_seqs, max_num_batched_tokens, self.max_model_len,
                               mixed=mixed_batches, prefill_chunk=prefill_chunk)
        self.runner = ModelRunner(cfg, weights, self.kv, tp)
        self.max_blocks_per_seq = -(-self.max_model_len // block_size)
        self.requests: dict[str, Request] = {}
        self.detok: dict[str, IncrementalDetokenizer] = {}
        self._text: dict[str, list] = {}
        self.step_count = 0
        self.use_graphs = use_cuda_graphs and self.device.type == "cuda"
        self.graphs: dict[int, tuple] = {}
        self.stats = {"prefill_steps": 0, "decode_steps": 0, "mixed_steps": 0, "prompt_tokens": 0,
                      "generated_tokens": 0, "prefill_time": 0.0, "decode_time": 0.0, "mixed_time": 0.0}
        self._rings: dict = {}
        self._ring_pos: dict = {}
        self.async_scheduling = async_scheduling
        self._ctl = None            # TP control channel (shared-memory ring), opened on first use
        self._inflight: dict | None = None
        cap = max(1024, max_num_seqs)
        self._tok_dev = torch.zeros(cap, dtype=torch.int64, device=self.device)      # last step's samples
        pin = self.device.type == "cuda"
        self._tok_host = [(torch.zeros(cap, dtype=torch.int64, pin_memory=pin),
                           torch.zeros(cap, dtype=torch.float32, pin_memory=pin),
                           torch.cuda.Event() if pin else None) for _ in range(2)]
        self._parity = 0
        self._arange = np.arange(max(self.max_model_len, block_size) + 1, dtype=np.int64)

    # ---- request API (driver) ------------------------------------------------------------------
    def add_request(self, prompt: str | list[int], params: SamplingParams | None = None,
                    request_id: str | None = None) -> str:
        params = params or SamplingParams()
        ids = list(prompt) if not isinstance(prompt, str) else self.tokenizer.encode(prompt)
        if not ids:
            ids = [self.cfg.bos_token_id]
        if len(ids) + params.max_tokens > self.max_model_len:
            params.max_tokens = max(1, self.max_model_len - len(ids))
        rid = request_id or f"cmpl-{uuid.uuid4().hex}"
        req = Request(rid, ids, params, seed=params.seed if params.seed is not None else hash(rid) & 0x7FFFFFFF)
        self.sched.add(req)
        self.requests[rid] = req
        self.detok[rid] = IncrementalDetokenizer(self.tokenizer, params.skip_special_tokens)
        # [text produced so far, characters already handed to the client, cut at a stop string]: stop strings
        # need the whole text (a match may span several tokens, and nothing after it may ever be streamed)
        self._text[rid] = ["", 0, False]
        return rid

    def abort(self, request_id: str) -> None:
        if self.sched.abort(request_id):
            self.requests.pop(request_id, None)
            self.detok.pop(request_id, None)
            self._text.pop(request_id, None)

    def has_unfinished(self) -> bool:
        return self.sched.has_work() or self._inflight is not None

    # ---- step ------------------------------------------------------------------------------------
    def _host_buffer(self, key, words: int):
        """Next pinned host staging buffer (int64 words) of ring `key`; waits for the H2D copy that last read it."""
        ring = self._rings.get(key)
        if ring is None or ring[0][0].numel() < words:
            cap = max(words, 1 << 12) if ring is None else max(words, 2 * ring[0][0].numel())
            pin = self.device.type == "cuda"
            ring = [[torch.zeros(cap, dtype=torch.int64, pin_memory=pin), None] for _ in range(4)]
            self._rings[key] = ring
            self._ring_pos[key] = 0
        i = self._ring_pos[key]
        self._ring_pos[key] = (i + 1) % len(ring)
        slot = ring[i]
        if slot[1] is not None:
            slot[1].synchronize()
        return slot

    def _to_device(self, slot, words: int, dst: torch.Tensor | None = None) -> torch.Tensor:
        host = slot[0][:words]
        if self.device.type != "cuda":
            return host.clone() if dst is None else dst[:words].copy_(host)
        if dst is None:
            dst = torch.empty(words, dtype=torch.int64, device=self.device)
        dst[:words].copy_(host, non_blocking=True)
        if slot[1] is None:
            slot[1] = torch.cuda.Event()
        slot[1].record()
        return dst

    def pack_step(self, nd: int, ids, pos, slots, ctx, tables, cu, max_seqlen: int, sampling=None,
                  src=None, sampled_prompts: int | None = None, prefix=None) -> "PackedStep":
        """Lay one step's metadata out in a pinned host buffer (see StepLayout). ids/pos/slots cover all T
        rows (decode rows first); ctx/tables the nd decode rows; cu the packed prompts (relative to row nd);
        sampling = (temperature, top_k, top_p, seeds) per sampled row or None. sampled_prompts: how many of the
        prompts (the first ones) end in this step; prefix = (pslots, pdst, csrc, ecu, max_ext) when prompt rows
        continue prompts whose earlier tokens are in the cache."""
        T, npf = len(ids), max(0, len(cu) - 1)
        ns = npf if sampled_prompts is None else sampled_prompts
        mb = max((len(t) for t in tables), default=0)
        mb = (mb + 7) & ~7
        lay = StepLayout(T, nd, npf, mb, ns, len(prefix[0]) if prefix is not None else 0)
        slot = self._host_buffer("step", lay.words)
        v = lay.views(slot[0].numpy(), np)
        v["ids"][:] = ids
        v["pos"][:] = pos
        v["slots"][:] = slots
        v["src"][:] = -1
        if src is not None and nd:
            v["src"][:nd] = src
        if nd:
            v["ctx"][:] = ctx
            tab = v["tables"]
            for i, t in enumerate(tables):
                tab[i, :len(t)] = t
        if npf:
            v["cu"][:] = cu
            v["last"][nd:] = np.asarray(cu[1:ns + 1], dtype=np.int64) + (nd - 1)
        if nd:
            v["last"][:nd] = np.arange(nd)
        if prefix is not None:
            v["pslots"][:], v["pdst"][:], v["csrc"][:], v["ecu"][:] = prefix[:4]
        if sampling is not None:
            v["temp"][:], v["topk"][:], v["topp"][:], v["seeds"][:] = sampling
        return PackedStep(lay, slot, int(max_seqlen), int(prefix[4]) if prefix is not None else 0)

    def _pack_batch(self, batch: Batch) -> "PackedStep":
        bs = self.blocks.block_size
        nd = batch.num_decode
        reqs = batch.requests
        ids, pos, slots, ctx, tables, cu, src = [], [], [], [], [], [], []
        for r in reqs[:nd]:
            t = r.sched_tokens - 1                 # position of the row's input token
            if r.pending:                          # sampled by the step in flight: fed on the device
                ids.append(0)
                src.append(r.row)
            else:
                ids.append((r.output_ids or r.prompt_ids)[-1])
                src.append(-1)
            pos.append(t)
            slots.append(r.blocks[t // bs] * bs + t % bs)
            ctx.append(t + 1)
            tables.append(r.blocks)
        max_seqlen = 0
        prefix = None
        if len(reqs) > nd:
            cu = [0]
            ar = self._arange
            sl = [np.asarray(slots, dtype=np.int64)]
            ps = [np.asarray(pos, dtype=np.int64)]
            pslots, pdst, csrc, ecu, max_ext = [], [], [], [0], 0
            for r in reqs[nd:]:
                toks = r.prompt_ids + r.output_ids
                a, n = r.num_computed, r.chunk              # this step prefills tokens [a, a + n) of the sequence
                ids += toks[a:a + n]
                ps.append(ar[a:a + n])
                blk = np.asarray(r.blocks[:-(-(a + n) // bs)], dtype=np.int64)
                allslots = (blk[:, None] * bs + ar[None, :bs]).reshape(-1)
                sl.append(allslots[a:a + n])
                cu.append(cu[-1] + n)
                max_seqlen = max(max_seqlen, n)
                # extended [prefix | chunk] layout for the attention of prompts that are under way
                pslots.append(allslots[:a])
                pdst.append(ecu[-1] + ar[:a])
                csrc.append(ecu[-1] + a + ar[:n])
                ecu.append(ecu[-1] + a + n)
                max_ext = max(max_ext, a + n)
            pos, slots = np.concatenate(ps), np.concatenate(sl)
            if ecu[-1] > cu[-1]:                             # at least one prompt has tokens in the cache already
                prefix = (np.concatenate(pslots), np.concatenate(pdst), np.concatenate(csrc), ecu, max_ext)
        reqs = reqs[:batch.num_sampled]
        sampling = ([r.params.temperature for r in reqs], [r.params.top_k for r in reqs],
                    [r.params.top_p for r in reqs],
                    # random stream keyed by (request seed, index of the token being generated): same seed =>
                    # same text, whatever else the engine is doing
                    [(r.seed * 1000003 + len(r.output_ids) + r.pending) & 0x7FFFFFFFFFFFFFFF for r in reqs])
        return self.pack_step(nd, ids, pos, slots, ctx, tables, cu, max_seqlen, sampling, src,
                              sampled_prompts=batch.num_sampled - nd, prefix=prefix)

    def _ring(self):
        """Shared-memory control ring between the driver and the workers (one node): created lazily by rank 0,
        attached by the others. DLTI_TP_CONTROL=gloo falls back to torch.distributed broadcasts."""
        if self._ctl is None:
            import os
            if os.environ.get("DLTI_TP_CONTROL", "shm") != "shm":
                self._ctl = False
            else:
                from .shm_ring import ShmRing, ring_name
                import torch.distributed as dist
                LLMEngine._ring_serial += 1
                name = ring_name(str(LLMEngine._ring_serial))
                if self.rank == 0:
                    self._ctl = ShmRing(name, -1, self.world - 1)
                    dist.barrier()
                else:
                    dist.barrier()
                    self._ctl = ShmRing(name, self.rank - 1, self.world - 1)
        return self._ctl

    _ring_serial = 0

    def _broadcast(self, step: "PackedStep | None") -> "PackedStep | None":
        """Driver -> workers: header[12] followed by the packed step buffer exactly as the driver staged it — ONE
        message through the shared-memory ring (N10), or two gloo broadcasts with DLTI_TP_CONTROL=gloo."""
        if self.world == 1:
            return step
        ring = self._ring()
        if ring:
            if self.rank == 0:
                if step is None:
                    ring.write(np.array([-1] + [0] * 11, dtype=np.int64))
                    return None
                lay = step.layout
                msg = np.empty(12 + lay.words, dtype=np.int64)
                msg[:12] = [0, lay.T, lay.nd, lay.npf, lay.mb, step.max_seqlen, lay.words, int(step.want_lp), lay.ns,
                            lay.P, step.max_ext, 0]
                msg[12:] = step.slot[0][:lay.words].numpy()
                ring.write(msg)
                return step
            raw = ring.read().view(np.int64)
            kind, T, nd, npf, mb, max_seqlen, words, want_lp, ns, P, max_ext = (int(x) for x in raw[:11])
            if kind < 0:
                return None
            lay = StepLayout(T, nd, npf, mb, ns, P)
            slot = self._host_buffer("step", words)
            slot[0][:words].numpy()[:] = raw[12:12 + words]
            step = PackedStep(lay, slot, max_seqlen, max_ext)
            step.want_lp = bool(want_lp)
            return step
        import torch.distributed as dist
        if self.rank == 0:
            if step is None:
                dist.broadcast(torch.tensor([-1] + [0] * 11, dtype=torch.int64), 0)
                return None
            lay = step.layout
            dist.broadcast(torch.tensor([0, lay.T, lay.nd, lay.npf, lay.mb, step.max_seqlen, lay.words,
                                         int(step.want_lp), lay.ns, lay.P, step.max_ext, 0], dtype=torch.int64), 0)
            dist.broadcast(step.slot[0][:lay.words], 0)
            return step
        hdr = torch.zeros(12, dtype=torch.int64)
        dist.broadcast(hdr, 0)
        kind, T, nd, npf, mb, max_seqlen, words, want_lp, ns, P, max_ext = (int(x) for x in hdr[:11])
        if kind < 0:
            return None
        lay = StepLayout(T, nd, npf, mb, ns, P)
        slot = self._host_buffer("step", words)
        dist.broadcast(slot[0][:words], 0)
        step = PackedStep(lay, slot, max_seqlen, max_ext)
        step.want_lp = bool(want_lp)
        return step

    def _feed(self, ids: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
        """Input ids with the rows that continue a sequence of the step in flight taken from its device-side
        samples (no host round trip between two steps)."""
        return torch.where(src >= 0, self._tok_dev[src.clamp_min(0).long()], ids)

    def _forward_step(self, step: "PackedStep"):
        """Run the forward pass of a packed step. Returns (logits [sampled rows, vocab], device views)."""
        lay = step.layout
        if lay.npf == 0 and self.use_graphs:
            return self._decode_graph(step)
        v = lay.views(self._to_device(step.slot, lay.words), torch)
        logits = self.runner.forward(self._feed(v["ids"], v["src"]), v["pos"], v["slots"], num_decode=lay.nd,
                                     block_tables=v["tables"] if lay.nd else None, ctx_lens=v["ctx"] if lay.nd else None,
                                     cu_seqlens=v["cu"] if lay.npf else None, max_seqlen=step.max_seqlen,
                                     last_idx=v["last"] if lay.npf else None, max_ctx=self.max_model_len,
                                     prefix=(v["pslots"], v["pdst"], v["csrc"], v["ecu"], step.max_ext) if lay.P else None)
        return logits, v

    def _run_step(self, step: "PackedStep", want_lp: bool = False):
        """Forward + sampling on this rank (every TP rank samples the same tokens from the same gathered
        logits, so each keeps its own device-side copy to feed the next step). Returns (tokens, logprobs) on
        the device; tokens is a view of the persistent sample buffer."""
        logits, v = self._forward_step(step)
        B = logits.shape[0]
        if B == 0:
            return self._tok_dev[:0], None
        if B > self._tok_dev.numel():
            raise RuntimeError(f"{B} sampled rows exceed the sample buffer ({self._tok_dev.numel()})")
        toks, lps = S.sample(logits, v["temp"], v["topk"], v["topp"], v["seeds"], step=0, want_logprob=want_lp,
                             out=self._tok_dev[:B])
        return toks, lps

    def _decode_graph(self, step: "PackedStep"):
        lay = step.layout
        B = lay.nd
        bucket = next((b for b in _BUCKETS if b >= B), None)
        if bucket is None:
            raise RuntimeError(f"decode batch {B} exceeds the largest CUDA-graph bucket")
        dev, mbs = self.device, self.max_blocks_per_seq
        g = self.graphs.get(bucket)
        if g is None:
            glay = StepLayout(bucket, bucket, 0, mbs)
            buf = torch.zeros(glay.words, dtype=torch.int64, device=dev)
            st = glay.views(buf, torch)
            st["slots"].fill_(-1)
            st["src"].fill_(-1)
            run = lambda: self.runner.forward(self._feed(st["ids"], st["src"]), st["pos"], st["slots"], num_decode=bucket,
                                              block_tables=st["tables"], ctx_lens=st["ctx"], max_ctx=self.max_model_len)
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):                 # warm-up outside capture (workspaces, TMA maps)
                for _ in range(2):
                    run()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = run()
            g = (graph, glay, buf, st, out)
            self.graphs[bucket] = g
        graph, glay, buf, st, out = g
        # re-lay the compact step into the graph's fixed layout (pinned), one H2D copy, replay
        slot = self._host_buffer(("graph", bucket), glay.words)
        src = lay.views(step.slot[0].numpy(), np)
        dst = glay.views(slot[0].numpy(), np)
        for k in ("ids", "pos", "seeds", "topk", "temp", "topp"):
            dst[k][:B] = src[k]
        dst["src"][:B] = src["src"]
        dst["src"][B:] = -1
        dst["slots"][:B] = src["slots"]
        dst["slots"][B:] = -1
        dst["ctx"][:B] = src["ctx"]
        dst["ctx"][B:] = 0
        w = min(lay.mb, mbs)
        dst["tables"][:B, :w] = src["tables"][:, :w]
        self._to_device(slot, glay.tables + (B * mbs + 1) // 2, buf)
        graph.replay()
        return out[:B], {k: t[:B] for k, t in st.items() if k in ("temp", "topk", "topp", "seeds")}

    def step(self) -> list[StepOutput]:
        """Driver: schedule and launch one model step on all ranks, then hand back the results of the step
        launched by the PREVIOUS call (async_scheduling, the default): while the host reads those tokens,
        updates the requests and streams text, the GPU is already running the next step, whose decode rows
        take their input tokens from the device-side samples. The price is one wasted row when a sequence
        ends on EOS / a stop token (known one step late); ends by length are predicted and cost nothing.
        With async_scheduling=False the step's own results are returned."""
        t0 = time.perf_counter()
        batch = self.sched.schedule()
        cur = self._launch(batch) if batch is not None else None
        if not self.async_scheduling:
            return self._finalize(cur, t0) if cur is not None else []
        prev, self._inflight = self._inflight, cur
        return self._finalize(prev, t0) if prev is not None else []

    def _launch(self, batch: Batch) -> dict:
        reqs = batch.requests[:batch.num_sampled]          # rows that produce a token in this step
        packed = self._pack_batch(batch)
        want_lp = any(r.params.logprobs is not None for r in reqs)
        packed.want_lp = want_lp
        self._broadcast(packed)
        toks, lps = self._run_step(packed, want_lp)
        B = len(reqs)
        host = None
        ev = None
        if B:
            self._parity ^= 1
            host = self._tok_host[self._parity]
            host[0][:B].copy_(toks, non_blocking=True)
            if lps is not None:
                host[1][:B].copy_(lps, non_blocking=True)
            if self.device.type == "cuda":
                ev = host[2]
                ev.record()
        for i, r in enumerate(reqs):
            r.pending += 1
            r.row = i
        for r in batch.requests[batch.num_decode:]:        # prompt rows: their tokens are in the cache now
            r.num_computed += r.chunk
            self.blocks.register(r.blocks, r.prompt_ids + r.output_ids, r.num_computed)   # prefix caching
        for r in batch.requests[batch.num_decode:batch.num_sampled]:
            r.prefill_done = True
        st = self.stats
        st[batch.kind + "_steps"] += 1
        st["prompt_tokens"] += packed.layout.T - batch.num_decode
        return dict(batch=batch, host=host, event=ev, want_lp=lps is not None)

    def _finalize(self, fl: dict, t0: float) -> list[StepOutput]:
        batch, host = fl["batch"], fl["host"]
        reqs = batch.requests[:batch.num_sampled]
        if fl["event"] is not None:
            fl["event"].synchronize()
        toks = host[0][:len(reqs)].tolist() if reqs else []
        lps = host[1][:len(reqs)].tolist() if (fl["want_lp"] and reqs) else None
        self.step_count += 1
        now = time.time()
        outs = []
        eos = self.tokenizer.eos_token_id
        for i, r in enumerate(reqs):
            r.pending -= 1
            if r.status == "finished":           # ended (EOS / stop / abort) after this row had been launched
                continue
            tok = toks[i]
            r.output_ids.append(tok)
            want = lps is not None and r.params.logprobs is not None
            if want:
                r.logprobs.append(lps[i])
            if r.first_token_time is None:
                r.first_token_time = now
            reason = None
            n_out = len(r.output_ids)
            if n_out >= r.params.min_tokens:
                if (tok == eos and not r.params.ignore_eos) or tok in r.params.stop_token_ids:
                    reason = "stop"
            if reason is None and r.params.stop and self._hit_stop_string(r):
                reason = "stop"                     # OpenAI `stop`: the text is cut before the match (text_delta)
            if reason is None and n_out >= r.params.max_tokens:
                reason = "length"
            if reason is None and r.num_tokens >= self.max_model_len:
                reason = "length"
            if reason:
                r.finish_time = now
                self.sched.finish(r, reason)
            outs.append(StepOutput(r.request_id, [tok], reason is not None, reason, [lps[i]] if want else None))
        self.stats[batch.kind + "_time"] += time.perf_counter() - t0
        self.stats["generated_tokens"] += len(outs)
        return outs

    def warmup(self, max_batch: int | None = None) -> int:
        """Driver, before serving: run one dummy decode step per CUDA-graph bucket up to max_batch (default
        max_num_seqs) on all ranks so that every graph is captured (and every workspace sized) before the first
        request — otherwise the first step at each new batch size pays the capture (hundreds of ms of TTFT).
        The dummy rows write no KV (slot -1) and their samples are dropped. Returns the number of buckets."""
        if not self.use_graphs:
            return 0
        top = max_batch or self.sched.max_num_seqs
        n = 0
        for b in _BUCKETS:
            step = self.pack_step(b, [0] * b, [0] * b, [-1] * b, [1] * b, [[0]] * b, [], 0,
                                  ([0.0] * b, [0] * b, [1.0] * b, [0] * b))
            self._broadcast(step)
            self._run_step(step)
            n += 1
            if b >= top:
                break
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        return n

    def shutdown_workers(self) -> None:
        if self.world > 1 and self.rank == 0:
            self._broadcast(None)

    def worker_loop(self) -> None:
        """TP ranks > 0: execute the steps the driver broadcasts until it sends the stop header."""
        assert self.rank != 0
        while True:
            step = self._broadcast(None)
            if step is None:
                return
            self._run_step(step, step.want_lp)

    # ---- text helpers / offline API ---------------------------------------------------------------
    def _advance_text(self, r: Request, final: bool) -> list:
        """Detokenise the request's new output tokens into its text buffer."""
        ids = r.output_ids
        if r.finish_reason == "stop" and ids and not r.params.ignore_eos and ids[-1] == self.tokenizer.eos_token_id:
            ids = ids[:-1]
        st = self._text[r.request_id]
        st[0] += self.detok[r.request_id].step(ids, final=final)
        return st

    def _hit_stop_string(self, r: Request) -> bool:
        """True when the text generated so far contains one of the request's stop strings; the text buffer is
        then cut just before the earliest match (the stop string itself is not returned, as in the OpenAI
        API), so nothing at or after it can reach the client."""
        st = self._advance_text(r, False)
        hits = [i for i in (st[0].find(s, max(0, st[1] - len(s) + 1)) for s in r.params.stop if s) if i >= 0]
        if not hits:
            return False
        st[0] = st[0][:max(min(hits), st[1])]
        st[2] = True
        return True

    def text_delta(self, request_id: str, final: bool = False) -> str:
        """Text of `request_id` not yet handed out. While the request is running and has stop strings, the last
        max(len(stop)) - 1 characters are held back: they could still turn out to be the start of a match."""
        r = self.requests[request_id]
        if not r.params.detokenize and not r.params.stop:
            return ""                               # the client asked for token ids only
        st = self._text[request_id]
        if not st[2]:                               # (a stop-string match has already fixed the final text)
            st = self._advance_text(r, final)
        end = len(st[0])
        if not final and r.status != "finished" and r.params.stop:
            end = max(st[1], end - (max(len(s) for s in r.params.stop) - 1))
        out = st[0][st[1]:end]
        st[1] = end
        return out

    def release(self, request_id: str) -> None:
        self.requests.pop(request_id, None)
        self.detok.pop(request_id, None)
        self._text.pop(request_id, None)

    def generate(self, prompts: list, params: SamplingParams | list[SamplingParams] | None = None) -> list[dict]:
        plist = params if isinstance(params, list) else [params or SamplingParams()] * len(prompts)
        rids = [self.add_request(p, SamplingParams(**vars(sp))) for p, sp in zip(prompts, plist)]
        while self.has_unfinished():
            self.step()
        out = []
        for rid in rids:
            r = self.requests[rid]
            out.append({"request_id": rid, "prompt_token_ids": r.prompt_ids, "token_ids": list(r.output_ids),
                        "text": self.text_delta(rid, final=True), "finish_reason": r.finish_reason,
                        "ttft": (r.first_token_time or r.arrival) - r.arrival})
            self.release(rid)
        return out
