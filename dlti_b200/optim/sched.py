"""Learning-rate schedules.

* LinearDecaySchedule — HF Trainer default (`get_linear_schedule_with_warmup`), used by the
  reference's baseline script (implicit defaults of TrainingArguments, train_baseline.py:172-193).
* WarmupLR — DeepSpeed `lr_schedules.WarmupLR`, configured by the `scheduler` block of every
  ds_config (reference configs/ds_config_zero1.json:16-23): log/linear warm-up from
  warmup_min_lr to warmup_max_lr over warmup_num_steps, then CONSTANT (no decay). DeepSpeed
  quirks kept: warmup_num_steps is clamped to >= 2, and before the first `step()` the rate is
  warmup_min_lr (that is the "Attempting to get learning rate from scheduler before it has
  started" warning visible in reference training/train.ipynb:323).
"""
from __future__ import annotations

import math
import sys


class LRSchedule:
    def __init__(self):
        self.last_step = -1  # number of scheduler.step() calls so far, minus 1

    def lr_at(self, it: int) -> float:
        raise NotImplementedError

    def get_lr(self) -> float:
        return self.lr_at(self.last_step)

    def step(self) -> float:
        self.last_step += 1
        return self.get_lr()

    def state_dict(self) -> dict:
        return {"last_step": self.last_step}

    def load_state_dict(self, sd: dict) -> None:
        self.last_step = int(sd["last_step"])


class LinearDecaySchedule(LRSchedule):
    def __init__(self, base_lr: float, total_steps: int, warmup_steps: int = 0):
        super().__init__()
        self.base_lr, self.total, self.warmup = base_lr, max(1, total_steps), warmup_steps
        self.last_step = 0  # torch LambdaLR semantics: epoch 0 before the first step

    def lr_at(self, it: int) -> float:
        if it < self.warmup:
            return self.base_lr * it / max(1, self.warmup)
        return self.base_lr * max(0.0, (self.total - it) / max(1, self.total - self.warmup))


class ConstantSchedule(LRSchedule):
    def __init__(self, base_lr: float):
        super().__init__()
        self.base_lr = base_lr

    def lr_at(self, it: int) -> float:
        return self.base_lr


class WarmupLR(LRSchedule):
    def __init__(self, warmup_min_lr: float = 0.0, warmup_max_lr: float = 1e-3,
                 warmup_num_steps: int = 1000, warmup_type: str = "log"):
        super().__init__()
        if warmup_type not in ("log", "linear"):
            raise ValueError(f"unknown warmup_type {warmup_type}")
        self.min_lr, self.max_lr = warmup_min_lr, warmup_max_lr
        self.warmup_num_steps = max(2, int(warmup_num_steps))
        self.warmup_type = warmup_type
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        self.warned = False

    def _gamma(self, it: int) -> float:
        if it < self.warmup_num_steps:
            if self.warmup_type == "log":
                return self.inverse_log_warm_up * math.log(it + 1)
            return it / self.warmup_num_steps
        return 1.0

    def lr_at(self, it: int) -> float:
        if it < 0:
            if not self.warned:
                print("Attempting to get learning rate from scheduler before it has started", file=sys.stderr)
                self.warned = True
            return self.min_lr
        return self.min_lr + (self.max_lr - self.min_lr) * self._gamma(it)


def build_scheduler(kind: str, **kw) -> LRSchedule:
    if kind == "linear":
        return LinearDecaySchedule(kw["base_lr"], kw["total_steps"], kw.get("warmup_steps", 0))
    if kind == "constant":
        return ConstantSchedule(kw["base_lr"])
    if kind == "WarmupLR":
        return WarmupLR(kw.get("warmup_min_lr", 0.0), kw["warmup_max_lr"], kw.get("warmup_num_steps", 0),
                        kw.get("warmup_type", "log"))
    raise ValueError(f"unknown scheduler {kind}")
