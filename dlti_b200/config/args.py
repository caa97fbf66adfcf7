"""TrainingArgs: the slice of HF `TrainingArguments` the reference scripts set or rely on
(reference training/train_baseline.py:172-193, train_deepspeed_zero2.py:261-288).
Defaults are HF's: seed 42, max_grad_norm 1.0, weight_decay 0, AdamW betas (0.9, 0.999) eps 1e-8,
warmup 0, linear schedule (DeepSpeed configs replace it with WarmupLR), logging every 10 steps.
"""
from __future__ import annotations

from dataclasses import asdict, dataclass


@dataclass
class TrainingArgs:
    output_dir: str = "./checkpoints/run"
    num_train_epochs: float = 3
    per_device_train_batch_size: int = 1
    gradient_accumulation_steps: int = 1
    learning_rate: float = 2e-4
    weight_decay: float = 0.0
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_epsilon: float = 1e-8
    max_grad_norm: float = 1.0
    warmup_steps: int = 0
    lr_scheduler_type: str = "linear"
    logging_steps: int = 10
    save_strategy: str = "epoch"      # "epoch" | "steps" | "no"
    save_steps: int = 500
    save_total_limit: int | None = None
    seed: int = 42
    precision: str | None = None      # None = take from the DS config; "bf16" | "fp16" | "fp32"
    gradient_checkpointing: bool = True
    max_steps: int = -1
    deepspeed: str | None = None      # path to a ds_config JSON
    no_offload: bool = False
    loss_norm: str = "tokens"         # "tokens" (App. H) | "microbatch" (pinned transformers 4.45)
    device: str = "cuda"
    backend: str = "fused"            # "fused" (peer-memory kernels) | "nccl" (baseline collectives)
    local_rank: int = -1
    report_to: str = "none"
    logging_dir: str | None = None
    disable_tqdm: bool = True
    profile: bool = False
    dataloader_pin_memory: bool = True

    def to_dict(self) -> dict:
        return asdict(self)
