"""Synthetic instruction-tuning data (no network on the bench box; BASELINE.json asks for
"synthetic Alpaca-shaped instruction data" of the named shape).

Two modes:
  * fixed_length: every sample is exactly `max_length` tokens — the benchmark shape (seq 2048);
  * Alpaca-shaped: [BOS] [INST] instruction(+input) [/INST] response [EOS] with log-normal
    segment lengths (instruction ~ 20-150 tokens, response ~ 30-500), truncated to max_length,
    matching the reference's `<s>[INST] {q} [/INST] {a}</s>` template
    (reference scripts/prepare_dataset.py:12-25) and its truncation (train_baseline.py:152-158).
Token ids follow a Zipf-like distribution with a learnable bigram structure so the loss can
actually decrease in the plumbing test.
"""
from __future__ import annotations

import numpy as np
import torch


class SyntheticInstructionDataset:
    def __init__(self, num_samples: int, vocab_size: int, max_length: int = 512, fixed_length: bool = False,
                 seed: int = 0, bos_id: int = 1, eos_id: int = 2):
        self.n, self.vocab, self.max_length, self.fixed = num_samples, vocab_size, max_length, fixed_length
        self.seed, self.bos, self.eos = seed, bos_id, eos_id
        self.column_names = ["input_ids", "attention_mask"]

    def __len__(self) -> int:
        return self.n

    def _tokens(self, rng: np.random.Generator, n: int) -> np.ndarray:
        # order-1 structure: next = (a * prev + noise) mod usable — gives the model something to fit
        usable = self.vocab - 3
        out = np.empty(n, dtype=np.int64)
        prev = int(rng.integers(0, usable))
        noise = rng.zipf(1.5, size=n) % 7
        for i in range(n):
            prev = (prev * 31 + 7 + int(noise[i])) % usable
            out[i] = prev + 3
        return out

    def __getitem__(self, i: int) -> dict:
        rng = np.random.default_rng((self.seed << 32) + i)
        if self.fixed:
            ids = rng.integers(3, self.vocab, size=self.max_length, dtype=np.int64)
            ids[0] = self.bos
        else:
            li = int(np.clip(rng.lognormal(3.8, 0.6), 8, 400))
            lr = int(np.clip(rng.lognormal(4.6, 0.7), 8, 1500))
            ids = np.concatenate([[self.bos], self._tokens(rng, li), self._tokens(rng, lr), [self.eos]])
            ids = ids[: self.max_length]
        return {"input_ids": ids.tolist(), "attention_mask": [1] * len(ids)}
