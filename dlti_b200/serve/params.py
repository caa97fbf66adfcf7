"""Sampling parameters and request/response records for the serving engine (SURVEY.md X2, X3,
App. C.5: the OpenAI /v1/completions knobs plus the vLLM extensions worth keeping)."""
from __future__ import annotations

import time
from dataclasses import dataclass, field


@dataclass
class SamplingParams:
    max_tokens: int = 16
    temperature: float = 1.0
    top_p: float = 1.0
    top_k: int = 0              # 0 / -1 = off
    n: int = 1
    seed: int | None = None
    stop: list[str] = field(default_factory=list)
    stop_token_ids: list[int] = field(default_factory=list)
    ignore_eos: bool = False
    min_tokens: int = 0
    logprobs: int | None = None
    echo: bool = False
    skip_special_tokens: bool = True
    detokenize: bool = True     # False (vLLM's field of the same name): return token ids only, produce no text

    def __post_init__(self):
        if isinstance(self.stop, str):
            self.stop = [self.stop]
        self.stop = list(self.stop or [])
        if self.top_k is None or self.top_k < 0:
            self.top_k = 0
        if self.max_tokens is None:
            self.max_tokens = 16
        if self.temperature < 0 or not (0 < self.top_p <= 1.0) or self.max_tokens < 1 or self.n < 1:
            raise ValueError("invalid sampling parameters")


@dataclass
class Request:
    request_id: str
    prompt_ids: list[int]
    params: SamplingParams
    arrival: float = field(default_factory=time.time)
    output_ids: list[int] = field(default_factory=list)
    logprobs: list[float] = field(default_factory=list)
    blocks: list[int] = field(default_factory=list)
    status: str = "waiting"          # waiting | running | finished
    finish_reason: str | None = None
    first_token_time: float | None = None
    finish_time: float | None = None
    text_emitted: int = 0            # characters already streamed
    seed: int = 0
    pending: int = 0                 # tokens sampled on the device that the host has not read yet (0..2)
    row: int = -1                    # row of this request's sample in the most recently launched step
    num_computed: int = 0            # prompt tokens already prefilled into the KV cache (chunked prefill)
    chunk: int = 0                   # prompt tokens this request contributes to the step being built
    prefill_done: bool = False       # its whole prompt has been launched (chunked prefill bookkeeping)

    @property
    def num_tokens(self) -> int:
        return len(self.prompt_ids) + len(self.output_ids)

    @property
    def sched_tokens(self) -> int:
        """Sequence length as the scheduler must see it: tokens in flight occupy positions and KV slots."""
        return len(self.prompt_ids) + len(self.output_ids) + self.pending


@dataclass
class StepOutput:
    request_id: str
    new_token_ids: list[int]
    finished: bool
    finish_reason: str | None
    logprobs: list[float] | None = None
