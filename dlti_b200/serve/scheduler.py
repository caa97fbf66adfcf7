"""Continuous-batching scheduler (call stack §3.4).

Policy: every engine step advances all running sequences by one token (decode rows) and, in the same
forward pass, admits waiting requests FCFS while KV blocks and the token budget allow (prefill rows,
whole prompts). `mixed=False` restores the prefill-first policy (a step is either prompts or decode
rows). When a running sequence cannot get a block to grow, the most recently admitted running request
is preempted (blocks freed, re-queued at the front; its tokens are recomputed on re-admission).
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass

from .kv_cache import BlockManager
from .params import Request


@dataclass
class Batch:
    kind: str                    # "prefill" | "decode" | "mixed"
    requests: list[Request]      # decode rows first, then prompts that complete in this step, then partial prompts
    num_decode: int = 0
    num_sampled: int = -1        # requests[:num_sampled] get a token from this step (-1 = all)

    def __post_init__(self):
        if self.num_sampled < 0:
            self.num_sampled = len(self.requests)


class Scheduler:
    def __init__(self, blocks: BlockManager, max_num_seqs: int = 256, max_num_batched_tokens: int = 8192,
                 max_model_len: int = 4096, mixed: bool = True, prefill_chunk: int = 0):
        self.blocks = blocks
        self.mixed = mixed           # False = prefill-first steps (no decode rows in a prefill pass)
        # > 0: a prompt contributes at most this many tokens to one step (chunked prefill): bounds how long a
        # step with prompt rows can stall the decode rows, and lets prompts longer than the token budget run.
        self.prefill_chunk = prefill_chunk
        self.max_num_seqs = max_num_seqs
        self.max_num_batched_tokens = max_num_batched_tokens
        self.max_model_len = max_model_len
        self.waiting: deque[Request] = deque()
        self.running: list[Request] = []
        self.num_preemptions = 0

    def add(self, req: Request) -> None:
        if len(req.prompt_ids) + 1 > self.max_model_len:
            raise ValueError(f"prompt of {len(req.prompt_ids)} tokens exceeds max_model_len {self.max_model_len}")
        self.waiting.append(req)

    def abort(self, request_id: str) -> bool:
        for q in (self.waiting, self.running):
            for r in list(q):
                if r.request_id == request_id:
                    q.remove(r)
                    self.blocks.release(r.blocks)
                    r.status, r.finish_reason = "finished", "abort"
                    return True
        return False

    def _take_cached_prefix(self, r: Request) -> int:
        """Prefix caching: start a waiting request's block list with the cached blocks that already hold the
        beginning of its tokens. Returns the number of tokens they cover (0 without prefix caching)."""
        if not self.blocks.prefix_caching or r.blocks:
            return 0
        r.blocks.extend(self.blocks.match_prefix(r.prompt_ids + r.output_ids))
        return len(r.blocks) * self.blocks.block_size

    def _at_limit(self, r: Request) -> bool:
        return (len(r.output_ids) + r.pending >= r.params.max_tokens) or r.sched_tokens >= self.max_model_len

    def has_work(self) -> bool:
        return bool(self.waiting or self.running)

    def schedule(self) -> Batch | None:
        """One engine step. The running sequences each grow by one token (decode rows); waiting prompts that
        fit the remaining token budget and the free KV blocks ride along in the same forward pass (prefill
        rows), so a new prompt never stalls the sequences that are already generating and the decode rows
        share the prefill's weight traffic. Batch.requests = decode rows first, then prefill rows."""
        if self.prefill_chunk > 0:
            return self._schedule_chunked()
        # ---- decode rows: every running sequence needs room for one more token ----
        # (a sequence whose tokens in flight already reach its length limit gets no row: it ends as soon as
        # the host has read them)
        preempted = False
        while self.running:
            ok = True
            for r in self.running:
                if not self._at_limit(r) and not self.blocks.ensure(r.blocks, r.sched_tokens + 1):
                    ok = False
                    break
            if ok:
                break
            victim = self.running.pop()           # preempt the youngest
            self.blocks.release(victim.blocks)
            victim.status = "waiting"
            self.waiting.appendleft(victim)
            self.num_preemptions += 1
            preempted = True
        decode = [r for r in self.running if not self._at_limit(r)]
        # ---- prefill rows: admit waiting prompts FCFS (not in a step that had to preempt) ----
        admitted: list[Request] = []
        budget = self.max_num_batched_tokens - len(decode)
        while self.waiting and not preempted and len(self.running) + len(admitted) < self.max_num_seqs:
            r = self.waiting[0]
            if r.pending:                         # preempted with a sampled token still unread: next step
                break
            n = r.num_tokens                      # prompt (+ already generated tokens if it was preempted)
            have = self._take_cached_prefix(r)    # tokens already in the cache (prefix caching), else 0
            if (n - have > budget and (admitted or decode)) or not self.blocks.ensure(r.blocks, n + 1):
                self.blocks.release(r.blocks)
                break
            self.waiting.popleft()
            r.status = "running"
            r.num_computed, r.chunk, r.prefill_done = have, n - have, True      # the rest of the prompt in one step
            admitted.append(r)
            budget -= n - have
            if budget <= 0:
                break
        if admitted and decode and not self.mixed:
            # prefill-first policy (vLLM v0 style): a step is either all prompts or all decode rows
            self.running.extend(admitted)
            return Batch("prefill", admitted, 0)
        self.running.extend(admitted)
        if not decode and not admitted:
            return None
        kind = "mixed" if decode and admitted else ("prefill" if admitted else "decode")
        return Batch(kind, decode + admitted, len(decode))

    def _schedule_chunked(self) -> Batch | None:
        """schedule() with chunked prefill: prompts enter the cache `prefill_chunk` tokens at a time (and never
        more than the step's remaining token budget); a prompt is sampled in the step that carries its last
        chunk. Prompts already under way are continued before new ones are admitted."""
        preempted = False
        while self.running:
            ok = True
            for r in self.running:
                if r.prefill_done and not self._at_limit(r) and not self.blocks.ensure(r.blocks, r.sched_tokens + 1):
                    ok = False
                    break
            if ok:
                break
            victim = self.running.pop()
            self._unschedule(victim)
            preempted = True
        decode = [r for r in self.running if r.prefill_done and not self._at_limit(r)]
        budget = self.max_num_batched_tokens - (len(decode) if self.mixed else 0)
        complete, partial = [], []

        def take(r: Request) -> bool:
            nonlocal budget
            remaining = r.num_tokens - r.num_computed
            c = min(remaining, self.prefill_chunk, budget)
            if c <= 0:
                return False
            last = c == remaining
            if not self.blocks.ensure(r.blocks, r.num_computed + c + (1 if last else 0)):
                return False
            r.chunk = c
            budget -= c
            (complete if last else partial).append(r)
            return True

        for r in self.running:                                  # prompts under way first
            if not r.prefill_done and not take(r):
                break
        stalled = any(not r.prefill_done and r not in complete and r not in partial for r in self.running)
        admitted = []
        while (self.waiting and not preempted and not stalled and budget > 0
               and len(self.running) + len(admitted) < self.max_num_seqs):
            r = self.waiting[0]
            if r.pending:
                break
            r.num_computed, r.prefill_done = self._take_cached_prefix(r), False
            if not take(r):
                self.blocks.release(r.blocks)
                r.num_computed = 0
                break
            self.waiting.popleft()
            r.status = "running"
            admitted.append(r)
        self.running.extend(admitted)
        prompts = complete + partial
        if prompts and decode and not self.mixed:
            decode = []                                          # prefill-first: no decode rows next to prompt rows
        if not decode and not prompts:
            if stalled and self.running:                        # every block is held by unfinished prompts
                self._unschedule(self.running.pop())
                return self._schedule_chunked()
            return None
        kind = "mixed" if decode and prompts else ("prefill" if prompts else "decode")
        return Batch(kind, decode + prompts, len(decode), len(decode) + len(complete))

    def _unschedule(self, victim: Request) -> None:
        self.blocks.release(victim.blocks)
        victim.status = "waiting"
        victim.num_computed, victim.chunk, victim.prefill_done = 0, 0, False
        self.waiting.appendleft(victim)
        self.num_preemptions += 1

    def finish(self, req: Request, reason: str) -> None:
        req.status, req.finish_reason = "finished", reason
        if req in self.running:
            self.running.remove(req)
        elif req in self.waiting:                 # preempted while its last token was still in flight
            self.waiting.remove(req)
        self.blocks.release(req.blocks)
