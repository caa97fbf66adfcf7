"""DeepSpeed-style JSON config: parsing, "auto" resolution and mismatch errors (SURVEY.md D10,
App. B). The same `configs/ds_config_zero{1,2,3}.json` files that drive the reference drive this
framework; every key present there is honoured.

Resolution rules follow HF's `HfTrainerDeepSpeedConfig` (SP:transformers/integrations/
deepspeed.py:105-258): an "auto" value takes the trainer's value; a concrete value that
disagrees with the trainer's is collected and reported in one ValueError. Precision is the one
deliberate difference (App. D #1): it is READ from the config (fp16.enabled / bf16.enabled) and
may be overridden from the CLI (--bf16/--fp16/--fp32), so the reference's ZeRO-1 script/config
mismatch cannot occur.
"""
from __future__ import annotations

import copy
import json
import os
import sys
from pathlib import Path
from typing import Any


class DSConfigError(ValueError):
    pass


def _get(d: dict, path: str, default=None):
    cur = d
    for k in path.split("."):
        if not isinstance(cur, dict) or k not in cur:
            return default
        cur = cur[k]
    return cur


def _set(d: dict, path: str, value) -> None:
    ks = path.split(".")
    for k in ks[:-1]:
        d = d.setdefault(k, {})
    d[ks[-1]] = value


class DSConfig:
    def __init__(self, raw: dict, path: str | None = None):
        self.raw = copy.deepcopy(raw)
        self.cfg = copy.deepcopy(raw)
        self.path = path
        self.mismatches: list[str] = []
        self.resolved = False
        stage = self.zero_stage
        if stage not in (0, 1, 2, 3):
            raise DSConfigError(f"zero_optimization.stage must be 0..3, got {stage}")

    # keys that are parsed for compatibility but have no effect here, and why (reported once at load time)
    ACCEPTED_NOT_USED = {
        "zero_optimization.stage3_prefetch_bucket_size": "the gather look-ahead is one whole decoder layer (queued "
                                                         "behind the running layer's GEMMs)",
        "zero_optimization.stage3_max_reuse_distance": "a gathered layer stays until its staging slot is needed by the "
                                                       "layer after next",
        "zero_optimization.sub_group_size": "the optimizer steps its whole partition in one kernel",
        "zero_optimization.round_robin_gradients": "gradients live in one flat buffer partitioned contiguously",
        "zero_optimization.contiguous_gradients": "the flat gradient buffer is always contiguous",
    }

    def ignored_keys(self) -> list[str]:
        out = [f"{k}: {why}" for k, why in self.ACCEPTED_NOT_USED.items() if _get(self.cfg, k) is not None]
        if self.zero_stage == 3 and _get(self.cfg, "zero_optimization.allgather_bucket_size") is not None:
            out.append("zero_optimization.allgather_bucket_size: stage 3 gathers parameters per layer, not per bucket")
        return out

    @classmethod
    def load(cls, path: str | Path) -> "DSConfig":
        with open(path) as f:
            c = cls(json.load(f), str(path))
        ign = c.ignored_keys()
        if ign and os.environ.get("RANK", "0") == "0" and not os.environ.get("DLTI_QUIET"):
            print(f"[ds_config] {path}: accepted for compatibility, no effect in this implementation:\n  - "
                  + "\n  - ".join(ign), file=sys.stderr)
        return c

    @classmethod
    def baseline(cls) -> "DSConfig":
        """Stage 0 config used by train_baseline.py (no DeepSpeed in the reference)."""
        return cls({"zero_optimization": {"stage": 0}, "gradient_clipping": "auto",
                    "train_batch_size": "auto", "train_micro_batch_size_per_gpu": "auto",
                    "gradient_accumulation_steps": "auto"})

    # ---- "auto" resolution ------------------------------------------------------------------
    def _fill(self, key: str, hf_name: str, hf_val: Any, must_match: bool = True) -> None:
        cur = _get(self.cfg, key)
        if cur is None:
            return
        if cur == "auto":
            _set(self.cfg, key, hf_val)
            return
        if not must_match:
            return
        same = (cur == hf_val) if not isinstance(hf_val, float) else abs(float(cur) - hf_val) <= 1e-12 * max(1.0, abs(hf_val))
        if not same:
            self.mismatches.append(f"- ds {key}={cur} vs hf {hf_name}={hf_val}")

    def resolve(self, *, micro_batch: int, grad_accum: int, world_size: int, lr: float,
                betas: tuple[float, float], eps: float, weight_decay: float, max_grad_norm: float,
                warmup_steps: int, total_steps: int, precision: str | None = None) -> "DSConfig":
        """Fill every "auto", check concrete values against the trainer's, settle precision."""
        self._fill("train_micro_batch_size_per_gpu", "per_device_train_batch_size", micro_batch)
        self._fill("gradient_accumulation_steps", "gradient_accumulation_steps", grad_accum)
        self._fill("train_batch_size", "train_batch_size (calculated)", micro_batch * grad_accum * world_size)
        self._fill("gradient_clipping", "max_grad_norm", max_grad_norm)
        self._fill("optimizer.params.lr", "learning_rate", lr)
        self._fill("optimizer.params.betas", "adam_beta1+adam_beta2", [betas[0], betas[1]])
        self._fill("optimizer.params.eps", "adam_epsilon", eps)
        self._fill("optimizer.params.weight_decay", "weight_decay", weight_decay)
        self._fill("scheduler.params.warmup_min_lr", "", 0, must_match=False)
        self._fill("scheduler.params.warmup_max_lr", "learning_rate", lr)
        self._fill("scheduler.params.warmup_num_steps", "warmup steps", warmup_steps)
        self._fill("scheduler.params.total_num_steps", "num_training_steps (calculated)", total_steps)
        # precision: CLI override > config
        if precision is not None:
            self.cfg.setdefault("fp16", {})["enabled"] = precision == "fp16"
            self.cfg.setdefault("bf16", {})["enabled"] = precision == "bf16"
        else:
            for blk in ("fp16", "bf16"):
                if _get(self.cfg, f"{blk}.enabled") == "auto":
                    _set(self.cfg, f"{blk}.enabled", False)
        if _get(self.cfg, "fp16.enabled") and _get(self.cfg, "bf16.enabled"):
            self.mismatches.append("- fp16.enabled and bf16.enabled are both true")
        for key in ("zero_optimization.reduce_bucket_size", "zero_optimization.stage3_prefetch_bucket_size",
                    "zero_optimization.stage3_param_persistence_threshold"):
            if _get(self.cfg, key) == "auto":
                _set(self.cfg, key, {"zero_optimization.reduce_bucket_size": 5e8,
                                     "zero_optimization.stage3_prefetch_bucket_size": 5e7,
                                     "zero_optimization.stage3_param_persistence_threshold": 1e5}[key])
        if self.mismatches:
            raise DSConfigError(
                "Please correct the following DeepSpeed config values that mismatch TrainingArguments"
                " values:\n" + "\n".join(self.mismatches) +
                "\nThe easiest method is to set these DeepSpeed config values to 'auto'.")
        bs = self.cfg.get("train_batch_size")
        if isinstance(bs, (int, float)) and int(bs) != micro_batch * grad_accum * world_size:
            raise DSConfigError(f"train_batch_size {bs} != micro {micro_batch} x accum {grad_accum} x world {world_size}")
        self.resolved = True
        return self

    # ---- typed accessors --------------------------------------------------------------------
    @property
    def zero(self) -> dict:
        return self.cfg.get("zero_optimization", {}) or {}

    @property
    def zero_stage(self) -> int:
        return int(self.zero.get("stage", 0))

    @property
    def precision(self) -> str:
        if _get(self.cfg, "fp16.enabled") is True:
            return "fp16"
        if _get(self.cfg, "bf16.enabled") is True:
            return "bf16"
        return "fp32"

    @property
    def fp16_block(self) -> dict:
        return self.cfg.get("fp16", {}) or {}

    @property
    def gradient_clipping(self) -> float:
        v = self.cfg.get("gradient_clipping", 0.0)
        return 0.0 if v in (None, "auto") else float(v)

    @property
    def has_optimizer(self) -> bool:
        return "optimizer" in self.cfg

    @property
    def optimizer_params(self) -> dict:
        t = _get(self.cfg, "optimizer.type", "AdamW")
        if t not in ("AdamW", "Adam"):
            raise DSConfigError(f"unsupported optimizer type {t} (AdamW/Adam only)")
        return dict(_get(self.cfg, "optimizer.params", {}) or {}, type=t)

    @property
    def has_scheduler(self) -> bool:
        return "scheduler" in self.cfg

    @property
    def scheduler(self) -> tuple[str, dict]:
        return _get(self.cfg, "scheduler.type", "WarmupLR"), dict(_get(self.cfg, "scheduler.params", {}) or {})

    def _zi(self, key: str, default) -> int:
        return int(float(self.zero.get(key, default)))

    @property
    def reduce_bucket_size(self) -> int:
        return self._zi("reduce_bucket_size", 5e8)

    @property
    def allgather_bucket_size(self) -> int:
        return self._zi("allgather_bucket_size", 5e8)

    @property
    def overlap_comm(self) -> bool:
        return bool(self.zero.get("overlap_comm", False))

    @property
    def reduce_scatter(self) -> bool:
        return bool(self.zero.get("reduce_scatter", True))

    @property
    def allgather_partitions(self) -> bool:
        return bool(self.zero.get("allgather_partitions", True))

    @property
    def contiguous_gradients(self) -> bool:
        return bool(self.zero.get("contiguous_gradients", True))

    @property
    def offload_optimizer(self) -> str:
        return str(_get(self.zero, "offload_optimizer.device", "none") or "none")

    @property
    def offload_param(self) -> str:
        return str(_get(self.zero, "offload_param.device", "none") or "none")

    @property
    def pin_memory(self) -> bool:
        return bool(_get(self.zero, "offload_optimizer.pin_memory", False) or
                    _get(self.zero, "offload_param.pin_memory", False))

    @property
    def stage3_prefetch_bucket_size(self) -> int:
        return self._zi("stage3_prefetch_bucket_size", 5e7)

    @property
    def stage3_param_persistence_threshold(self) -> int:
        return self._zi("stage3_param_persistence_threshold", 1e5)

    @property
    def stage3_max_live_parameters(self) -> int:
        return self._zi("stage3_max_live_parameters", 1e9)

    @property
    def stage3_max_reuse_distance(self) -> int:
        return self._zi("stage3_max_reuse_distance", 1e9)

    @property
    def stage3_gather_16bit_weights_on_model_save(self) -> bool:
        return bool(self.zero.get("stage3_gather_16bit_weights_on_model_save", False))

    @property
    def steps_per_print(self) -> int:
        return int(self.cfg.get("steps_per_print", 10) or 10)

    @property
    def wall_clock_breakdown(self) -> bool:
        return bool(self.cfg.get("wall_clock_breakdown", False))

    def override_no_offload(self) -> None:
        z = self.cfg.setdefault("zero_optimization", {})
        for k in ("offload_optimizer", "offload_param"):
            if k in z:
                z[k] = dict(z[k], device="none")
