"""Shared driver behind training/train_baseline.py and training/train_deepspeed_zero{1,2,3}.py.

Each reference script is ~300 lines of the same five stages ([1/5] tokenizer, [2/5] model, [3/5]
LoRA, [4/5] dataset, [5/5] trainer) with different argparse defaults; here the
stages live once and the four entrypoints only declare their defaults. Flags, defaults, output
directory rules, banners, the metrics row and the final hint line follow the reference
(training/train_baseline.py:27-271, train_deepspeed_zero1.py:29-315, zero2.py:37-392,
zero3.py:35-330). Added optional flags: --synthetic, --model_config, --max_length, --max_steps,
--bf16/--fp16/--fp32, --device, --no_offload, --seed, --backend, --num_samples, --loss_norm,
--no_gradient_checkpointing, --mask_eos_like_reference.
"""
from __future__ import annotations

import argparse
import os
import sys
import time
from dataclasses import dataclass
from pathlib import Path

import torch

from ..config.args import TrainingArgs
from ..data.collate import CausalLMCollator
from ..data.synthetic import SyntheticInstructionDataset
from ..models.config import resolve_config
from ..models.lora import LoraConfig, print_trainable_parameters
from ..parallel import comm
from ..utils.metrics import (RULE, create_experiment_name, get_zero_stage_from_config,
                             print_metrics_summary, save_training_metrics)
from .trainer import Trainer


@dataclass
class ScriptSpec:
    kind: str                       # "baseline" | "zero1" | "zero2" | "zero3"
    description: str
    output_dir: str | None
    batch_size: int
    grad_accum: int
    ds_flag: str | None             # "--deepspeed_config" | "--deepspeed" | None
    ds_default: str | None
    ds_required: bool
    has_resume: bool
    save_strategy: str
    save_steps: int
    save_total_limit: int
    writes_metrics: bool


SPECS = {
    "baseline": ScriptSpec("baseline", "Baseline LoRA fine-tuning (single GPU, no ZeRO)",
                           "./checkpoints/baseline_1gpu", 1, 16, None, None, False, False, "epoch", 500, 2, True),
    "zero1": ScriptSpec("zero1", "Fine-tune with ZeRO-1 (optimizer-state partitioning)",
                        "./checkpoints/deepspeed_zero1", 1, 16, "--deepspeed_config",
                        "./configs/ds_config_zero1.json", False, True, "steps", 100, 3, True),   # metrics row added (reference zero1 wrote none)
    "zero2": ScriptSpec("zero2", "Fine-tune with ZeRO-2 (+ gradient partitioning)", None, 1, 1,
                        "--deepspeed_config", "./configs/ds_config_zero2.json", False, True, "steps", 100, 3, True),
    "zero3": ScriptSpec("zero3", "Fine-tune with ZeRO-3 (+ parameter partitioning, optional CPU offload)",
                        None, 2, 4, "--deepspeed", None, True, False, "epoch", 500, 2, True),
}


def build_parser(spec: ScriptSpec) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description=spec.description)
    p.add_argument("--model_name", type=str, default="meta-llama/Llama-2-7b-hf", help="Base model")
    p.add_argument("--dataset_path", type=str, default="./data/glaive_code_full", help="Preprocessed dataset path")
    p.add_argument("--output_dir", type=str, default=spec.output_dir, help="Checkpoint directory")
    p.add_argument("--num_train_epochs", type=int, default=3, help="Number of epochs")
    p.add_argument("--per_device_train_batch_size", type=int, default=spec.batch_size, help="Batch size per GPU")
    p.add_argument("--gradient_accumulation_steps", type=int, default=spec.grad_accum, help="Gradient accumulation")
    p.add_argument("--learning_rate", type=float, default=2e-4, help="Learning rate")
    p.add_argument("--lora_r", type=int, default=16, help="LoRA rank")
    if spec.ds_flag:
        p.add_argument(spec.ds_flag, dest="ds_config", type=str, default=spec.ds_default,
                       required=spec.ds_required, help="ZeRO (DeepSpeed-format) JSON config")
    if spec.has_resume:
        p.add_argument("--resume_from_checkpoint", action="store_true", help="Resume from latest checkpoint")
    if spec.kind != "baseline":
        p.add_argument("--local_rank", type=int, default=-1, help="Local rank for distributed training")
    # ---- additions (all optional; reference invocations keep working) ----
    p.add_argument("--synthetic", action="store_true", help="Synthetic Alpaca-shaped data + random-init weights")
    p.add_argument("--model_config", type=str, default=None, help="tiny | llama2-7b | llama2-70b | <hf dir>")
    p.add_argument("--max_length", type=int, default=512, help="Truncation length (reference hard-codes 512)")
    p.add_argument("--fixed_length", action="store_true", help="Synthetic samples are exactly max_length long")
    p.add_argument("--num_samples", type=int, default=2048, help="Synthetic dataset size")
    p.add_argument("--max_steps", type=int, default=-1)
    g = p.add_mutually_exclusive_group()
    g.add_argument("--bf16", action="store_true")
    g.add_argument("--fp16", action="store_true")
    g.add_argument("--fp32", action="store_true")
    p.add_argument("--device", type=str, default="cuda", choices=["cuda", "cpu"])
    p.add_argument("--no_offload", action="store_true", help="Ignore offload_optimizer / offload_param")
    p.add_argument("--seed", type=int, default=42)
    p.add_argument("--backend", type=str, default="fused", choices=["fused", "nccl"])
    p.add_argument("--loss_norm", type=str, default="tokens", choices=["tokens", "microbatch"])
    p.add_argument("--no_gradient_checkpointing", action="store_true")
    p.add_argument("--mask_eos_like_reference", action="store_true",
                   help="labels masked wherever id == pad id (pad = eos), as HF's collator does")
    p.add_argument("--logging_steps", type=int, default=10)
    p.add_argument("--profile", action="store_true")
    return p


def _rank0_print(is_main, *a, **k):
    if is_main:
        print(*a, **k, flush=True)


def run_training(kind: str, argv: list[str] | None = None) -> dict | None:
    spec = SPECS[kind]
    args = build_parser(spec).parse_args(argv)
    ds_path = getattr(args, "ds_config", None)
    if ds_path and not os.path.exists(ds_path):
        # reference defect App. D #2: zero1 defaulted to ./training/ds_config_zero1.json
        for alt in (os.path.join("configs", os.path.basename(ds_path)),
                    os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(__file__))), "configs",
                                 os.path.basename(ds_path))):
            if os.path.exists(alt):
                ds_path = alt
                break
        else:
            raise FileNotFoundError(f"ZeRO config not found: {ds_path}")
    device = args.device
    if device == "cuda" and not torch.cuda.is_available():
        raise RuntimeError("CUDA required (pass --device cpu for the CPU plumbing path)")
    info = comm.init_distributed(device, getattr(args, "local_rank", -1))
    is_main = info.is_main
    num_gpus = info.world_size                                        # App. D #7: not device_count()
    zero_stage = get_zero_stage_from_config(ds_path) if ds_path else 0
    experiment = create_experiment_name(num_gpus, zero_stage)
    if args.output_dir is None:
        args.output_dir = f"./checkpoints/{experiment}"
    elif kind == "zero1":
        args.output_dir = f"{args.output_dir}_{num_gpus}gpu"

    title = "BASELINE TRAINING" if kind == "baseline" else f"ZeRO-{zero_stage} TRAINING (B200-native)"
    _rank0_print(is_main, "\n" + RULE + f"\n{title}\n" + RULE)
    _rank0_print(is_main, f"\nExperiment: {experiment}\nGPUs: {num_gpus}\nZeRO Stage: {zero_stage}")
    if ds_path:
        _rank0_print(is_main, f"Config: {ds_path}")
    _rank0_print(is_main, f"Output: {args.output_dir}\n")
    if is_main and device == "cuda":
        for i in range(torch.cuda.device_count()):
            print(f"  GPU {i}: {torch.cuda.get_device_name(i)}")       # App. D #11 fixed

    # [1/5] tokenizer
    _rank0_print(is_main, "[1/5] Loading tokenizer...")
    tokenizer = None
    model_dir = args.model_name if os.path.isdir(args.model_name) else None
    cfg = resolve_config(args.model_config or args.model_name)
    pad_id = cfg.pad_token_id if cfg.pad_token_id is not None else cfg.eos_token_id   # pad = eos
    if not args.synthetic:
        from ..data.text import load_tokenizer
        tokenizer = load_tokenizer(args.model_name)
        pad_id = tokenizer.pad_token_id
    _rank0_print(is_main, "Tokenizer loaded" if tokenizer else "Synthetic token stream (no tokenizer needed)")

    # [2/5] model config (weights are materialised by the Trainer: resident or ZeRO-3 shards)
    _rank0_print(is_main, "[2/5] Loading model...")
    if not args.synthetic and model_dir is None:
        raise FileNotFoundError(f"model '{args.model_name}' is not a local HF directory; there is no network "
                                f"access — pass a local path or --synthetic --model_config <preset>")
    _rank0_print(is_main, f"Model: {cfg.name} ({cfg.num_params():,} parameters)")

    # [3/5] LoRA — fixed hyper-parameters of the reference (App. A.2)
    _rank0_print(is_main, f"[3/5] Applying LoRA (r={args.lora_r})...")
    lcfg = LoraConfig(r=args.lora_r, lora_alpha=2 * args.lora_r, lora_dropout=0.05,
                      target_modules=("q_proj", "k_proj", "v_proj", "o_proj"), bias="none",
                      base_model_name_or_path=args.model_name)

    # [4/5] dataset
    _rank0_print(is_main, "[4/5] Loading dataset...")
    if args.synthetic:
        dataset = SyntheticInstructionDataset(args.num_samples, cfg.vocab_size, args.max_length,
                                              fixed_length=args.fixed_length, seed=args.seed,
                                              bos_id=cfg.bos_token_id, eos_id=cfg.eos_token_id)
    else:
        if not os.path.exists(args.dataset_path):
            raise FileNotFoundError(f"Dataset not found: {args.dataset_path}")
        from ..data.text import load_text_dataset, tokenize_dataset
        dataset = tokenize_dataset(load_text_dataset(args.dataset_path), tokenizer, args.max_length,
                                   desc="Tokenizing" if is_main else None)
    _rank0_print(is_main, f"Dataset ready: {len(dataset):,} samples")
    collator = CausalLMCollator(pad_id, mask_by="pad_id" if args.mask_eos_like_reference else "attention")

    # [5/5] trainer
    _rank0_print(is_main, "[5/5] Configuring training...")
    precision = "bf16" if args.bf16 else "fp16" if args.fp16 else "fp32" if args.fp32 else None
    if precision is None and kind == "baseline":
        precision = "fp16" if device == "cuda" else "fp32"            # reference baseline: fp16=True
    if device == "cpu" and precision is None:
        precision = "fp32"
    targs = TrainingArgs(
        output_dir=args.output_dir, num_train_epochs=args.num_train_epochs,
        per_device_train_batch_size=args.per_device_train_batch_size,
        gradient_accumulation_steps=args.gradient_accumulation_steps, learning_rate=args.learning_rate,
        logging_steps=args.logging_steps, save_strategy=spec.save_strategy, save_steps=spec.save_steps,
        save_total_limit=spec.save_total_limit, seed=args.seed, precision=precision,
        gradient_checkpointing=not args.no_gradient_checkpointing, max_steps=args.max_steps,
        deepspeed=ds_path, no_offload=args.no_offload, loss_norm=args.loss_norm, device=device,
        backend=args.backend, local_rank=getattr(args, "local_rank", -1),
        logging_dir=f"{args.output_dir}/logs", profile=args.profile)
    trainer = Trainer(cfg, lcfg, targs, dataset, collator, model_path=model_dir, info=info,
                      tokenizer=tokenizer)
    if is_main:
        print_trainable_parameters(cfg, trainer.lora)
        print("LoRA applied")
        eff = args.per_device_train_batch_size * args.gradient_accumulation_steps * num_gpus
        print(f"Effective batch size: {eff}")
        print(f"Precision: {trainer.precision} | ZeRO stage {trainer.stage} | backend "
              f"{'fused peer-memory kernels' if trainer.fused else 'collectives (baseline)'}")

    _rank0_print(is_main, "\n" + RULE + f"\nSTARTING TRAINING: {experiment.upper()}\n" + RULE + "\n")
    resume = bool(getattr(args, "resume_from_checkpoint", False))
    if device == "cuda":
        torch.cuda.reset_peak_memory_stats()
    t0 = time.time()
    res = trainer.train(resume_from_checkpoint=True if resume else None)
    training_time = time.time() - t0

    _rank0_print(is_main, "\n" + RULE + "\nSAVING MODEL\n" + RULE)
    final_dir = f"{args.output_dir}/final"
    trainer.save_model(final_dir)                    # all ranks enter (App. D #3)
    _rank0_print(is_main, f"Model saved to {final_dir}")
    metrics = None
    if is_main:
        n = len(dataset)
        epochs_done = args.num_train_epochs if args.max_steps <= 0 else None
        samples = (n * args.num_train_epochs if epochs_done is not None else
                   trainer.global_step * args.per_device_train_batch_size *
                   args.gradient_accumulation_steps * num_gpus)
        metrics = {
            "experiment": experiment, "num_gpus": num_gpus, "zero_stage": zero_stage,
            "strategy": "pytorch_lora" if kind == "baseline" else f"deepspeed_zero{zero_stage}",
            "training_time_hours": training_time / 3600,
            "samples_per_second": samples / training_time,
            "peak_memory_gb": res["peak_memory_gb"],
            "final_loss": res["final_loss"] if res["final_loss"] is not None else 0.0,
            "tokens_per_second": res["tokens_per_second"], "max_rank_step_ms": res["max_rank_step_ms"],
            "peak_memory_gb_max_rank": res["peak_memory_gb_max_rank"], "dtype": trainer.precision,
            "seq_len": args.max_length,
        }
        if spec.writes_metrics:
            print("\n" + RULE + "\nCOLLECTING METRICS\n" + RULE)
            print_metrics_summary(metrics)
            save_training_metrics(metrics)
        print("\n" + RULE + f"\n{experiment.upper()} TRAINING COMPLETE!\n" + RULE)
        print(f"\nTime: {training_time / 3600:.2f} hours")
        print(f"Throughput: {metrics['samples_per_second']:.1f} samples/sec")
        print(f"Memory: {metrics['peak_memory_gb']:.2f} GB")
        print(f"\nCheckpoints: {args.output_dir}")
        if spec.writes_metrics:
            print("Metrics: results/training_metrics.csv")
            print("\nRun 'python scripts/compare_training.py' to see comparison\n")
    comm.barrier(info)
    return metrics
