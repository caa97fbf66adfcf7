#!/usr/bin/env python
"""Merge a PEFT LoRA adapter into its base model and write a servable HF Llama checkpoint
(W' = W + alpha/r * B A per target module). Output layout: config.json + (sharded) safetensors +
tokenizer files — the `merged_model/` directory the reference's .gitignore:20 hints at.

    python scripts/merge_adapter.py --base <hf_dir> --adapter checkpoints/zero3_8gpu/final --out merged_model
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--base", required=True)
    ap.add_argument("--adapter", required=True)
    ap.add_argument("--out", default="merged_model")
    ap.add_argument("--device", default="cpu")
    a = ap.parse_args()
    from dlti_b200.models.merge import merge_checkpoint

    out = merge_checkpoint(a.base, a.adapter, a.out, device=a.device)
    print(f"merged model written to {out}")


if __name__ == "__main__":
    main()
