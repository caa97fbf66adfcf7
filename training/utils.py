"""Shared helpers for the training entrypoints (importable as `training.utils` and, from inside
training/, as `utils` — the reference imports it both ways: train_baseline.py:24, zero2.py:29)."""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from dlti_b200.utils.metrics import (create_experiment_name, get_zero_stage_from_config,  # noqa: E402,F401
                                     print_metrics_summary, save_training_metrics)

__all__ = ["create_experiment_name", "get_zero_stage_from_config", "save_training_metrics",
           "print_metrics_summary"]
