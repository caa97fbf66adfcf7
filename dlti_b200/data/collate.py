"""Causal-LM collator: pad to the longest sequence in the batch with pad_token_id;
labels = input_ids with pad positions set to -100 — HF `DataCollatorForLanguageModeling(mlm=False)`
(SP:transformers/data/data_collator.py:790-792; reference training/train_baseline.py:195-198).

Because the reference sets pad = eos, HF's rule also masks REAL eos tokens out of the loss
(App. D #8). `mask_by="pad_id"` reproduces that; `mask_by="attention"` (default) masks only true
padding.
"""
from __future__ import annotations

import torch


class CausalLMCollator:
    def __init__(self, pad_token_id: int, mask_by: str = "attention", pad_to_multiple_of: int | None = None):
        assert mask_by in ("attention", "pad_id")
        self.pad, self.mask_by, self.mult = pad_token_id, mask_by, pad_to_multiple_of

    def __call__(self, features: list[dict]) -> dict[str, torch.Tensor]:
        n = max(len(f["input_ids"]) for f in features)
        if self.mult:
            n = (n + self.mult - 1) // self.mult * self.mult
        B = len(features)
        ids = torch.full((B, n), self.pad, dtype=torch.long)
        am = torch.zeros((B, n), dtype=torch.long)
        for i, f in enumerate(features):
            t = torch.as_tensor(f["input_ids"], dtype=torch.long)
            ids[i, : t.numel()] = t
            am[i, : t.numel()] = 1
        labels = ids.clone()
        if self.mask_by == "pad_id":
            labels[labels == self.pad] = -100
        else:
            labels[am == 0] = -100
        return {"input_ids": ids, "attention_mask": am, "labels": labels}
