#!/usr/bin/env python
"""Single-GPU LoRA fine-tuning without ZeRO (the comparison baseline).

Usage:
    python training/train_baseline.py --dataset_path ./data/glaive_code_full
    python -m dlti_b200.launch --num_gpus=N training/train_baseline.py ...   (deepspeed-launcher style)

Same flags, defaults, output layout and metrics row as the reference entrypoint
(reference training/train_baseline.py). The five stages themselves live in
dlti_b200.engine.cli; add --synthetic --model_config llama2-7b to run without a checkpoint or dataset.
"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from dlti_b200.engine.cli import run_training  # noqa: E402


def main():
    run_training("baseline")


if __name__ == "__main__":
    main()
