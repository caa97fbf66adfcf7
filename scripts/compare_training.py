#!/usr/bin/env python
"""Compare training runs recorded in results/training_metrics.csv.

Reads the CSV, derives speedup (vs the `baseline` row, else the first row) and per-GPU scaling
efficiency, prints the comparison table and key findings, and — when matplotlib is installed —
writes results/plots/training_comparison.png (2x2: time, speedup, peak memory, efficiency).
Same inputs/outputs as the reference tool (reference scripts/compare_training.py:16-196); additions:
tokens/s and device-step-time columns when present, efficiency grouped per ZeRO stage (App. D #15),
graceful degradation without matplotlib/seaborn (neither is installed on the B200 image).

Usage:
    python scripts/compare_training.py
"""
import os
import sys

import pandas as pd

CSV_PATH = "results/training_metrics.csv"
PLOT_PATH = "results/plots/training_comparison.png"
BAR = "=" * 70


def load_and_calculate(csv_path: str = CSV_PATH) -> pd.DataFrame:
    if not os.path.exists(csv_path):
        print(f"No results found at {csv_path}")
        print("\nRun training experiments first!")
        sys.exit(1)
    df = pd.read_csv(csv_path)
    print(f"\n{BAR}\nLOADED {len(df)} EXPERIMENTS\n{BAR}")
    print("Experiments: " + ", ".join(df["experiment"].astype(str)))
    print()
    base = df.loc[df["experiment"] == "baseline", "training_time_hours"]
    if base.empty:
        print("No baseline experiment found.")
        print("Using first experiment as reference for speedup calculation.")
        ref_time = float(df["training_time_hours"].iloc[0])
    else:
        ref_time = float(base.iloc[0])
        print(f"Baseline time: {ref_time:.2f} hours")
    df["speedup"] = ref_time / df["training_time_hours"]
    df["efficiency_percent"] = df["speedup"] / df["num_gpus"] * 100.0
    if "tokens_per_second" in df.columns:
        one = df[df["num_gpus"] == 1].groupby("zero_stage")["tokens_per_second"].first()
        df["token_scaling_percent"] = [
            (r.tokens_per_second / (one[r.zero_stage] * r.num_gpus) * 100.0) if r.zero_stage in one.index else float("nan")
            for r in df.itertuples()]
    return df


def print_comparison_table(df: pd.DataFrame) -> None:
    print(f"\n{BAR}\nTRAINING COMPARISON\n{BAR}\n")
    # every column of the metrics CSV (as the reference prints them: experiment, strategy, zero_stage, num_gpus,
    # times, samples/s, peak memory, final loss, timestamp) followed by the derived ones
    lead = ["experiment", "strategy", "zero_stage", "num_gpus", "training_time_hours", "training_time_seconds",
            "samples_per_second", "peak_memory_gb", "final_loss", "speedup", "efficiency_percent"]
    cols = [c for c in lead if c in df.columns]
    cols += [c for c in df.columns if c not in cols]
    view = df[cols].copy()
    for c, nd in (("training_time_hours", 2), ("samples_per_second", 1), ("peak_memory_gb", 2), ("speedup", 2),
                  ("efficiency_percent", 1), ("tokens_per_second", 0), ("max_rank_step_ms", 1),
                  ("token_scaling_percent", 1)):
        if c in view.columns:
            view[c] = view[c].round(nd)
    print(view.to_string(index=False))


def print_key_findings(df: pd.DataFrame) -> None:
    print(f"\n{BAR}\nKEY FINDINGS\n{BAR}")
    base = df[df["experiment"] == "baseline"]
    ref_time = float(base["training_time_hours"].iloc[0]) if len(base) else float(df["training_time_hours"].iloc[0])
    for r in df.itertuples():
        if r.experiment == "baseline":
            continue
        print(f"\n{r.experiment}:")
        print(f"  Speedup: {r.speedup:.2f}x on {r.num_gpus} GPU(s)")
        print(f"  Scaling efficiency: {r.efficiency_percent:.1f}%")
        print(f"  Time saved vs reference run: {ref_time - r.training_time_hours:.2f} hours")
        print(f"  Memory per GPU: {r.peak_memory_gb:.2f} GB")
        if hasattr(r, "tokens_per_second") and r.tokens_per_second == r.tokens_per_second:
            print(f"  Tokens/sec (all GPUs): {r.tokens_per_second:,.0f}")


def create_plots(df: pd.DataFrame, out_path: str = PLOT_PATH) -> bool:
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except Exception:
        print("\nmatplotlib is not installed: skipping plots (table above is complete).")
        return False
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    fig, ax = plt.subplots(2, 2, figsize=(14, 10))
    names = df["experiment"].astype(str)
    ax[0, 0].bar(names, df["training_time_hours"]); ax[0, 0].set_title("Training time (hours)")
    ax[0, 1].bar(names, df["speedup"]); ax[0, 1].axhline(1.0, ls="--", c="r"); ax[0, 1].set_title("Speedup vs baseline")
    ax[1, 0].bar(names, df["peak_memory_gb"]); ax[1, 0].set_title("Peak memory per GPU (GB)")
    for stage, grp in df.groupby("zero_stage"):
        g = grp.sort_values("num_gpus")
        ax[1, 1].plot(g["num_gpus"], g["efficiency_percent"], marker="o", label=f"ZeRO-{stage}" if stage else "baseline")
    ax[1, 1].axhline(100.0, ls="--", c="g"); ax[1, 1].set_xlabel("GPUs"); ax[1, 1].set_title("Scaling efficiency (%)")
    ax[1, 1].legend()
    for a in (ax[0, 0], ax[0, 1], ax[1, 0]):
        a.tick_params(axis="x", rotation=45)
    fig.tight_layout()
    fig.savefig(out_path, dpi=300, bbox_inches="tight")
    print(f"\nPlots saved to {out_path}")
    return True


def main() -> None:
    df = load_and_calculate()
    print_comparison_table(df)
    print_key_findings(df)
    create_plots(df)
    print(f"\n{BAR}\nCOMPARISON COMPLETE\n{BAR}\n")


if __name__ == "__main__":
    main()
