#!/usr/bin/env python
"""LoRA fine-tuning with ZeRO-3 (parameters partitioned too; optional CPU offload), 1..8 GPUs.

Usage:
    torchrun --nproc_per_node=8 training/train_deepspeed_zero3.py --deepspeed configs/ds_config_zero3.json --bf16 --no_offload
    python -m dlti_b200.launch --num_gpus=N training/train_deepspeed_zero3.py ...   (deepspeed-launcher style)

Same flags, defaults, output layout and metrics row as the reference entrypoint
(reference training/train_deepspeed_zero3.py). The five stages themselves live in
dlti_b200.engine.cli; add --synthetic --model_config llama2-7b to run without a checkpoint or dataset.
"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from dlti_b200.engine.cli import run_training  # noqa: E402


def main():
    run_training("zero3")


if __name__ == "__main__":
    main()
