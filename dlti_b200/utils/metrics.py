"""Experiment naming, ZeRO-stage lookup, metrics CSV and summary block.

Same observable behaviour as the reference's training/utils.py (names `baseline` / `zero{S}_{N}gpu`,
append-with-header CSV at results/training_metrics.csv with the column order
experiment,num_gpus,zero_stage,strategy,training_time_hours,samples_per_second,peak_memory_gb,
final_loss — scripts/compare_training.py indexes by these names), plus appended columns
tokens_per_second, max_rank_step_ms, peak_memory_gb_max_rank, dtype, seq_len.
"""
from __future__ import annotations

import csv
import json
from pathlib import Path

CORE_COLUMNS = ["experiment", "num_gpus", "zero_stage", "strategy", "training_time_hours",
                "samples_per_second", "peak_memory_gb", "final_loss"]
EXTRA_COLUMNS = ["tokens_per_second", "max_rank_step_ms", "peak_memory_gb_max_rank", "dtype", "seq_len"]
RULE = "=" * 70


def create_experiment_name(num_gpus: int, zero_stage: int) -> str:
    """
    >>> create_experiment_name(1, 0)
    'baseline'
    >>> create_experiment_name(2, 2)
    'zero2_2gpu'
    >>> create_experiment_name(4, 3)
    'zero3_4gpu'
    """
    return "baseline" if zero_stage == 0 else f"zero{zero_stage}_{num_gpus}gpu"


def get_zero_stage_from_config(config_path: str) -> int:
    return int(json.loads(Path(config_path).read_text())["zero_optimization"]["stage"])


def save_training_metrics(metrics: dict, csv_path: str = "results/training_metrics.csv") -> None:
    """Append one row; write the header only when the file is new. If an existing file has fewer
    columns (e.g. written by the reference), new keys are dropped so the file stays rectangular."""
    path = Path(csv_path)
    path.parent.mkdir(parents=True, exist_ok=True)
    fields = list(metrics.keys())
    new_file = not path.is_file()
    if not new_file:
        with open(path, newline="") as f:
            header = next(csv.reader(f), None)
        if header:
            fields = header
    with open(path, "a", newline="") as f:
        w = csv.DictWriter(f, fieldnames=fields, extrasaction="ignore", restval="")
        if new_file:
            w.writeheader()
        w.writerow(metrics)
    print(f"\nMetrics saved to {csv_path}")


def print_metrics_summary(metrics: dict) -> None:
    lines = ["", RULE, "TRAINING METRICS", RULE, "",
             f"Experiment: {metrics['experiment']}",
             f"GPUs: {metrics['num_gpus']}",
             f"Strategy: {metrics['strategy']}", "",
             f"Training time: {metrics['training_time_hours']:.2f} hours",
             f"Throughput: {metrics['samples_per_second']:.1f} samples/sec",
             f"Memory/GPU: {metrics['peak_memory_gb']:.2f} GB",
             f"Final loss: {metrics['final_loss']:.4f}"]
    if "tokens_per_second" in metrics:
        lines.append(f"Tokens/sec (all GPUs): {metrics['tokens_per_second']:.0f}")
    if "max_rank_step_ms" in metrics:
        lines.append(f"Device step time (max over ranks): {metrics['max_rank_step_ms']:.1f} ms")
    lines.append(RULE)
    print("\n".join(lines))
