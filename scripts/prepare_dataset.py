#!/usr/bin/env python
"""Prepare the instruction dataset used by the training entrypoints.

    python scripts/prepare_dataset.py [--num_samples N] [--output_dir ./data] [--synthetic]

Online (default): downloads glaiveai/glaive-code-assistant with `datasets`, optionally keeps the
first N rows, renders each row with the Llama-2 chat template `<s>[INST] q [/INST] a</s>` and
saves to {output_dir}/glaive_code_{N//1000}k or {output_dir}/glaive_code_full — the behaviour of the
reference tool (reference scripts/prepare_dataset.py:12-155).
Offline (--synthetic, or automatically when the download fails): writes a synthetic
Alpaca-shaped corpus of the same schema (one `text` column) so the pipeline can be exercised on a
box without network access.
"""
import argparse
import os
import sys
from pathlib import Path

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.data.text import format_conversation_for_llama2, format_conversations_for_llama2  # noqa: E402,F401

BAR = "=" * 70
_VERBS = ["Write", "Implement", "Explain", "Refactor", "Debug", "Optimize", "Document", "Test"]
_THINGS = ["a binary search", "an LRU cache", "a REST client", "a tokenizer", "a matrix transpose",
           "a thread pool", "a JSON parser", "a rate limiter", "a trie", "a heap"]
_LANGS = ["Python", "C++", "Rust", "Go", "JavaScript", "CUDA"]


def synthetic_rows(n: int, seed: int = 0):
    import random

    rng = random.Random(seed)
    for i in range(n):
        q = f"{rng.choice(_VERBS)} {rng.choice(_THINGS)} in {rng.choice(_LANGS)} (variant {i})."
        body = "\n".join(f"    step_{j} = compute(step_{j - 1 if j else 0}, {rng.randint(0, 99)})"
                         for j in range(rng.randint(3, 40)))
        yield {"question": q, "answer": f"Here is one way to do it:\n\ndef solution():\n{body}\n    return step_0\n"}


def dataset_dir_name(num_samples) -> str:
    return f"glaive_code_{num_samples // 1000}k" if num_samples else "glaive_code_full"


def prepare_glaive_dataset(num_samples=None, output_dir="./data", synthetic=False):
    from datasets import Dataset

    print(f"{BAR}\nPREPARING GLAIVE-CODE-ASSISTANT DATASET\n{BAR}")
    Path(output_dir).mkdir(parents=True, exist_ok=True)
    ds = None
    if not synthetic:
        print("\n[1/4] Loading dataset from HuggingFace...\nDataset: glaiveai/glaive-code-assistant")
        try:
            from datasets import load_dataset
            ds = load_dataset("glaiveai/glaive-code-assistant", split="train")
            print(f"Loaded {len(ds):,} samples")
        except Exception as e:  # no network on the bench box
            print(f"Could not download ({type(e).__name__}); falling back to the synthetic corpus")
    if ds is None:
        n = num_samples or 10000
        print(f"\n[1/4] Generating {n:,} synthetic instruction/response pairs...")
        ds = Dataset.from_list(list(synthetic_rows(n)))
    if num_samples:
        print(f"\n[2/4] Subsampling to {num_samples:,} samples...")
        ds = ds.select(range(min(num_samples, len(ds))))
    else:
        print("\n[2/4] Using full dataset")
    print("\n[3/4] Formatting conversations...")
    import time
    t0 = time.time()
    ds = ds.map(format_conversations_for_llama2, batched=True, batch_size=4096, remove_columns=ds.column_names,
                desc="Formatting")
    print(f"Formatted {len(ds):,} examples in {time.time() - t0:.2f} s ({len(ds) / max(1e-9, time.time() - t0):,.0f} examples/s)")
    print("Sample:\n" + "-" * 70 + "\n" + ds[0]["text"][:500] + "\n" + "-" * 70)
    out = os.path.join(output_dir, dataset_dir_name(num_samples))
    print(f"\n[4/4] Saving to {out}...")
    ds.save_to_disk(out)
    size_mb = sum(f.stat().st_size for f in Path(out).rglob("*") if f.is_file()) / 1e6
    print(f"Saved {len(ds):,} samples ({size_mb:.1f} MB)\n{BAR}\nDATASET READY\n{BAR}")
    print(f"\nUse with: --dataset_path {out}")
    return ds


def main():
    ap = argparse.ArgumentParser(description="Prepare Glaive-Code-Assistant dataset")
    ap.add_argument("--num_samples", type=int, default=None, help="Number of samples (default: full dataset)")
    ap.add_argument("--output_dir", type=str, default="./data", help="Output directory")
    ap.add_argument("--synthetic", action="store_true", help="Skip the download; write a synthetic corpus")
    a = ap.parse_args()
    if prepare_glaive_dataset(a.num_samples, a.output_dir, a.synthetic) is None:
        sys.exit(1)


if __name__ == "__main__":
    main()
