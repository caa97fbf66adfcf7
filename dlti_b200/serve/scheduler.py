"""Continuous-batching scheduler (call stack §3.4).

Policy (vLLM v0 style, prefill-first): each engine step is either a PREFILL step — admit waiting
requests FCFS while KV blocks and the token budget allow, whole prompts in one pass — or a DECODE
step over every running sequence (one new token each). When a decode step cannot get a block for a
growing sequence, the most recently admitted running request is preempted (blocks freed, re-queued
at the front; its tokens are recomputed on re-admission).
"""
from __future__ import annotations

from collections import deque
from dataclasses import dataclass

from .kv_cache import BlockManager
from .params import Request


@dataclass
class Batch:
    kind: str                    # "prefill" | "decode"
    requests: list[Request]


class Scheduler:
    def __init__(self, blocks: BlockManager, max_num_seqs: int = 256, max_num_batched_tokens: int = 8192,
                 max_model_len: int = 4096):
        self.blocks = blocks
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

    def has_work(self) -> bool:
        return bool(self.waiting or self.running)

    def schedule(self) -> Batch | None:
        # ---- prefill step: admit as many waiting prompts as fit ----
        admitted: list[Request] = []
        budget = self.max_num_batched_tokens
        while self.waiting and len(self.running) + len(admitted) < self.max_num_seqs:
            r = self.waiting[0]
            n = r.num_tokens                      # prompt (+ already generated tokens if it was preempted)
            if n > budget and admitted:
                break
            if not self.blocks.ensure(r.blocks, n + 1):
                break
            self.waiting.popleft()
            r.status = "running"
            admitted.append(r)
            budget -= n
            if budget <= 0:
                break
        if admitted:
            self.running.extend(admitted)
            return Batch("prefill", admitted)
        if not self.running:
            return None
        # ---- decode step: every running sequence grows by one token ----
        while True:
            ok = True
            for r in self.running:
                if not self.blocks.ensure(r.blocks, r.num_tokens + 1):
                    ok = False
                    break
            if ok:
                break
            victim = self.running.pop()           # preempt the youngest
            self.blocks.release(victim.blocks)
            victim.status = "waiting"
            self.waiting.appendleft(victim)
            self.num_preemptions += 1
            if not self.running:
                return None
        return Batch("decode", list(self.running))

    def finish(self, req: Request, reason: str) -> None:
        req.status, req.finish_reason = "finished", reason
        if req in self.running:
            self.running.remove(req)
        self.blocks.release(req.blocks)
