"""Serving engine: request queue -> continuous-batching scheduler -> model runner -> sampler ->
incremental detokenisation (call stack §3.4).

One engine object per rank. Rank 0 is the driver: it owns the scheduler / block manager and, each step,
broadcasts the step's metadata (one int64 tensor over gloo — a shared-memory-speed control message,
N10) to the TP workers, which run `worker_loop()` and execute the same forward. Decode steps replay
CUDA graphs captured per batch-size bucket (S12); the collectives inside are graph-safe (device-side
barrier epochs).
"""
from __future__ import annotations

import time
import uuid

import torch

from ..models.config import LlamaConfig
from ..ops import serving as S
from .kv_cache import BlockManager, KVCache
from .model_runner import ModelRunner, TPContext, TPWeights
from .params import Request, SamplingParams, StepOutput
from .scheduler import Batch, Scheduler
from .tokenizer import IncrementalDetokenizer

_BUCKETS = [1, 2, 4, 8, 16, 32, 64, 128, 256, 512]


class LLMEngine:
    def __init__(self, cfg: LlamaConfig, weights: TPWeights, tp: TPContext, tokenizer, *, info=None,
                 block_size: int = 16, num_blocks: int | None = None, gpu_memory_utilization: float = 0.85,
                 max_num_seqs: int = 256, max_num_batched_tokens: int = 8192, max_model_len: int | None = None,
                 use_cuda_graphs: bool = True):
        self.cfg, self.w, self.tp, self.tokenizer, self.info = cfg, weights, tp, tokenizer, info
        self.device = weights.embed.device
        self.dtype = weights.embed.dtype
        self.max_model_len = max_model_len or cfg.max_position_embeddings
        self.rank = tp.rank
        self.world = tp.world
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
        self.blocks = BlockManager(num_blocks, block_size)
        self.sched = Scheduler(self.blocks, max_num_seqs, max_num_batched_tokens, self.max_model_len)
        self.runner = ModelRunner(cfg, weights, self.kv, tp)
        self.max_blocks_per_seq = -(-self.max_model_len // block_size)
        self.requests: dict[str, Request] = {}
        self.detok: dict[str, IncrementalDetokenizer] = {}
        self.step_count = 0
        self.use_graphs = use_cuda_graphs and self.device.type == "cuda"
        self.graphs: dict[int, tuple] = {}
        self.stats = {"prefill_steps": 0, "decode_steps": 0, "prompt_tokens": 0, "generated_tokens": 0,
                      "prefill_time": 0.0, "decode_time": 0.0}

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
        return rid

    def abort(self, request_id: str) -> None:
        if self.sched.abort(request_id):
            self.requests.pop(request_id, None)
            self.detok.pop(request_id, None)

    def has_unfinished(self) -> bool:
        return self.sched.has_work()

    # ---- step ------------------------------------------------------------------------------------
    def _meta_from_batch(self, batch: Batch) -> dict:
        bs = self.blocks.block_size
        reqs = batch.requests
        if batch.kind == "prefill":
            ids, pos, slots, cu = [], [], [], [0]
            for r in reqs:
                toks = r.prompt_ids + r.output_ids
                n = len(toks)
                ids += toks
                pos += list(range(n))
                slots += [r.blocks[t // bs] * bs + t % bs for t in range(n)]
                cu.append(cu[-1] + n)
            last = [c - 1 for c in cu[1:]]
            return dict(kind=0, ids=ids, pos=pos, slots=slots, cu=cu, last=last,
                        max_seqlen=max(b - a for a, b in zip(cu, cu[1:])), B=len(reqs))
        ids, pos, slots, ctx, tables = [], [], [], [], []
        for r in reqs:
            t = r.num_tokens - 1
            ids.append((r.output_ids or r.prompt_ids)[-1])
            pos.append(t)
            slots.append(r.blocks[t // bs] * bs + t % bs)
            ctx.append(t + 1)
            tables.append(list(r.blocks))
        return dict(kind=1, ids=ids, pos=pos, slots=slots, ctx=ctx, tables=tables, B=len(reqs))

    def _broadcast(self, meta: dict | None) -> dict | None:
        """Driver -> workers. Layout: header[8] then payload."""
        if self.world == 1:
            return meta
        import torch.distributed as dist
        if self.rank == 0:
            if meta is None:
                hdr = torch.tensor([-1, 0, 0, 0, 0, 0, 0, 0], dtype=torch.int64)
                dist.broadcast(hdr, 0)
                return None
            if meta["kind"] == 0:
                payload = meta["ids"] + meta["pos"] + meta["slots"] + meta["cu"] + meta["last"]
                hdr = torch.tensor([0, len(meta["ids"]), meta["B"], meta["max_seqlen"], len(payload), 0, 0, 0], dtype=torch.int64)
            else:
                mb = max(len(t) for t in meta["tables"])
                flat = [b for t in meta["tables"] for b in (t + [0] * (mb - len(t)))]
                payload = meta["ids"] + meta["pos"] + meta["slots"] + meta["ctx"] + flat
                hdr = torch.tensor([1, len(meta["ids"]), meta["B"], mb, len(payload), 0, 0, 0], dtype=torch.int64)
            dist.broadcast(hdr, 0)
            dist.broadcast(torch.tensor(payload, dtype=torch.int64), 0)
            return meta
        hdr = torch.zeros(8, dtype=torch.int64)
        dist.broadcast(hdr, 0)
        kind, T, B, x, n = (int(v) for v in hdr[:5])
        if kind < 0:
            return None
        pl = torch.zeros(n, dtype=torch.int64)
        dist.broadcast(pl, 0)
        pl = pl.tolist()
        if kind == 0:
            return dict(kind=0, ids=pl[:T], pos=pl[T:2 * T], slots=pl[2 * T:3 * T], cu=pl[3 * T:3 * T + B + 1],
                        last=pl[3 * T + B + 1:], max_seqlen=x, B=B)
        tables = [pl[4 * B + i * x:4 * B + (i + 1) * x] for i in range(B)]
        return dict(kind=1, ids=pl[:B], pos=pl[B:2 * B], slots=pl[2 * B:3 * B], ctx=pl[3 * B:4 * B], tables=tables, B=B)

    def _run_meta(self, meta: dict) -> torch.Tensor:
        dev = self.device
        t64 = lambda x: torch.tensor(x, dtype=torch.int64, device=dev)
        t32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
        if meta["kind"] == 0:
            return self.runner.forward("prefill", t64(meta["ids"]), t32(meta["pos"]), t64(meta["slots"]),
                                       cu_seqlens=t32(meta["cu"]), max_seqlen=meta["max_seqlen"],
                                       last_idx=t64(meta["last"]))
        B = meta["B"]
        if self.use_graphs:
            return self._decode_graph(meta)[:B]
        mb = max(len(t) for t in meta["tables"])
        tables = torch.zeros(B, mb, dtype=torch.int32)
        for i, t in enumerate(meta["tables"]):
            tables[i, :len(t)] = torch.tensor(t, dtype=torch.int32)
        return self.runner.forward("decode", t64(meta["ids"]), t32(meta["pos"]), t64(meta["slots"]),
                                   block_tables=tables.to(dev), ctx_lens=t32(meta["ctx"]),
                                   max_ctx=self.max_model_len)

    def _decode_graph(self, meta: dict) -> torch.Tensor:
        B = meta["B"]
        bucket = next((b for b in _BUCKETS if b >= B), None)
        if bucket is None:
            raise RuntimeError(f"decode batch {B} exceeds the largest CUDA-graph bucket")
        dev, mbs = self.device, self.max_blocks_per_seq
        g = self.graphs.get(bucket)
        if g is None:
            st = dict(ids=torch.zeros(bucket, dtype=torch.int64, device=dev),
                      pos=torch.zeros(bucket, dtype=torch.int32, device=dev),
                      slots=torch.full((bucket,), -1, dtype=torch.int64, device=dev),
                      ctx=torch.zeros(bucket, dtype=torch.int32, device=dev),
                      tables=torch.zeros(bucket, mbs, dtype=torch.int32, device=dev))
            run = lambda: self.runner.forward("decode", st["ids"], st["pos"], st["slots"], block_tables=st["tables"],
                                              ctx_lens=st["ctx"], max_ctx=self.max_model_len)
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
            g = (graph, st, out)
            self.graphs[bucket] = g
        graph, st, out = g
        # stage the step's inputs (pinned host -> device copies of a few KB)
        ids = torch.zeros(bucket, dtype=torch.int64); ids[:B] = torch.tensor(meta["ids"], dtype=torch.int64)
        pos = torch.zeros(bucket, dtype=torch.int32); pos[:B] = torch.tensor(meta["pos"], dtype=torch.int32)
        slots = torch.full((bucket,), -1, dtype=torch.int64); slots[:B] = torch.tensor(meta["slots"], dtype=torch.int64)
        ctx = torch.zeros(bucket, dtype=torch.int32); ctx[:B] = torch.tensor(meta["ctx"], dtype=torch.int32)
        tables = torch.zeros(bucket, mbs, dtype=torch.int32)
        for i, t in enumerate(meta["tables"]):
            tables[i, :len(t)] = torch.tensor(t, dtype=torch.int32)
        st["ids"].copy_(ids, non_blocking=True); st["pos"].copy_(pos, non_blocking=True)
        st["slots"].copy_(slots, non_blocking=True); st["ctx"].copy_(ctx, non_blocking=True)
        st["tables"].copy_(tables, non_blocking=True)
        graph.replay()
        return out

    def step(self) -> list[StepOutput]:
        """Driver: schedule, run one model step on all ranks, sample, update requests."""
        batch = self.sched.schedule()
        if batch is None:
            return []
        t0 = time.perf_counter()
        meta = self._meta_from_batch(batch)
        self._broadcast(meta)
        logits = self._run_meta(meta)
        reqs = batch.requests
        dev = self.device
        temp = torch.tensor([r.params.temperature for r in reqs], dtype=torch.float32, device=dev)
        topk = torch.tensor([r.params.top_k for r in reqs], dtype=torch.int32, device=dev)
        topp = torch.tensor([r.params.top_p for r in reqs], dtype=torch.float32, device=dev)
        # random stream keyed by (request seed, index of the token being generated): same seed => same text,
        # whatever else the engine is doing
        seeds = torch.tensor([(r.seed * 1000003 + len(r.output_ids)) & 0x7FFFFFFFFFFFFFFF for r in reqs],
                             dtype=torch.int64, device=dev)
        want_lp = any(r.params.logprobs is not None for r in reqs)
        toks, lps = S.sample(logits, temp, topk, topp, seeds, step=0, want_logprob=want_lp)
        toks = toks.tolist()
        lps = lps.tolist() if lps is not None else None
        self.step_count += 1
        now = time.time()
        outs = []
        eos = self.tokenizer.eos_token_id
        for i, r in enumerate(reqs):
            tok = toks[i]
            r.output_ids.append(tok)
            if lps is not None:
                r.logprobs.append(lps[i])
            if r.first_token_time is None:
                r.first_token_time = now
            reason = None
            n_out = len(r.output_ids)
            if n_out >= r.params.min_tokens:
                if (tok == eos and not r.params.ignore_eos) or tok in r.params.stop_token_ids:
                    reason = "stop"
            if reason is None and n_out >= r.params.max_tokens:
                reason = "length"
            if reason is None and r.num_tokens >= self.max_model_len:
                reason = "length"
            if reason:
                r.finish_time = now
                self.sched.finish(r, reason)
            outs.append(StepOutput(r.request_id, [tok], reason is not None, reason, [lps[i]] if lps else None))
        dt = time.perf_counter() - t0
        if batch.kind == "prefill":
            self.stats["prefill_steps"] += 1
            self.stats["prompt_tokens"] += len(meta["ids"])
            self.stats["prefill_time"] += dt
        else:
            self.stats["decode_steps"] += 1
            self.stats["decode_time"] += dt
        self.stats["generated_tokens"] += len(reqs)
        return outs

    def shutdown_workers(self) -> None:
        if self.world > 1 and self.rank == 0:
            self._broadcast(None)

    def worker_loop(self) -> None:
        """TP ranks > 0: execute the steps the driver broadcasts until it sends the stop header."""
        assert self.rank != 0
        while True:
            meta = self._broadcast(None)
            if meta is None:
                return
            self._run_meta(meta)

    # ---- text helpers / offline API ---------------------------------------------------------------
    def text_delta(self, request_id: str, final: bool = False) -> str:
        r = self.requests[request_id]
        ids = r.output_ids
        if r.finish_reason == "stop" and ids and not r.params.ignore_eos and ids[-1] == self.tokenizer.eos_token_id:
            ids = ids[:-1]
        text = self.detok[request_id].step(ids, final=final)
        for s in r.params.stop:
            if s and s in text:
                text = text[: text.index(s)]
        return text

    def release(self, request_id: str) -> None:
        self.requests.pop(request_id, None)
        self.detok.pop(request_id, None)

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
