#!/usr/bin/env python
"""LoRA fine-tuning with ZeRO-1 (optimizer-state partitioning), 1..8 GPUs.

Usage:
    torchrun --nproc_per_node=4 training/train_deepspeed_zero1.py --deepspeed_config configs/ds_config_zero1.json
    python -m dlti_b200.launch --num_gpus=N training/train_deepspeed_zero1.py ...   (deepspeed-launcher style)

Same flags, defaults, output layout and metrics row as the reference entrypoint
(reference training/train_deepspeed_zero1.py). The five stages themselves live in
dlti_b200.engine.cli; add --synthetic --model_config llama2-7b to run without a checkpoint or dataset.
"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from dlti_b200.engine.cli import run_training  # noqa: E402


def main():
    run_training("zero1")


if __name__ == "__main__":
    main()
