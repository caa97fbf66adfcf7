"""Checkpoint save / rotate / resume.

Directory contract kept from HF Trainer + DeepSpeed so the reference's resume scan
(reference training/train_deepspeed_zero2.py:302-315: pick the highest `checkpoint-*`) works:

  {output_dir}/checkpoint-{global_step}/
      adapter_config.json, adapter_model.safetensors        (PEFT layout; trainable params only)
      trainer_state.json                                     (global_step, epoch, log_history, ...)
      rng_state.pth | rng_state_{rank}.pth                   (python, numpy, cpu, cuda)
      training_args.bin
      ZeRO stage 0 : optimizer.pt, scheduler.pt, scaler.pt
      ZeRO stage>=1: latest  +  global_step{N}/zero_pp_rank_{r}_mp_rank_00_optim_states.pt (per rank)
                                 global_step{N}/mp_rank_00_model_states.pt (scheduler, counters)

Byte compatibility with DeepSpeed's .pt payloads is not a goal (App. C.3); names and layout are.
Fix of reference defect App. D #4: the resume directory is discovered on rank 0 and broadcast.
"""
from __future__ import annotations

import json
import os
import random
import re
import shutil
from pathlib import Path

import numpy as np
import torch

_CKPT_RE = re.compile(r"^checkpoint-(\d+)$")


def list_checkpoints(output_dir: str | Path) -> list[Path]:
    d = Path(output_dir)
    if not d.is_dir():
        return []
    found = [(int(m.group(1)), p) for p in d.iterdir() if p.is_dir() and (m := _CKPT_RE.match(p.name))]
    return [p for _, p in sorted(found)]


def latest_checkpoint(output_dir: str | Path) -> Path | None:
    c = list_checkpoints(output_dir)
    return c[-1] if c else None


def rotate_checkpoints(output_dir: str | Path, save_total_limit: int | None) -> list[Path]:
    """Delete the oldest checkpoints beyond the limit (SP:transformers/trainer.py:3092-3099)."""
    if not save_total_limit or save_total_limit <= 0:
        return []
    ck = list_checkpoints(output_dir)
    doomed = ck[: max(0, len(ck) - save_total_limit)]
    for p in doomed:
        shutil.rmtree(p, ignore_errors=True)
    return doomed


def rng_state() -> dict:
    st = {"python": random.getstate(), "numpy": np.random.get_state(), "cpu": torch.get_rng_state()}
    if torch.cuda.is_available():
        st["cuda"] = torch.cuda.get_rng_state()
    return st


def set_rng_state(st: dict) -> None:
    random.setstate(st["python"])
    np.random.set_state(st["numpy"])
    torch.set_rng_state(st["cpu"])
    if "cuda" in st and torch.cuda.is_available():
        try:
            torch.cuda.set_rng_state(st["cuda"])
        except Exception:
            pass


def save_trainer_state(path: Path, state: dict) -> None:
    with open(path / "trainer_state.json", "w") as f:
        json.dump(state, f, indent=2)


def load_trainer_state(path: Path) -> dict:
    with open(Path(path) / "trainer_state.json") as f:
        return json.load(f)


def rng_file(path: Path, rank: int, world: int) -> Path:
    return path / ("rng_state.pth" if world == 1 else f"rng_state_{rank}.pth")


def ds_tag_dir(path: Path, global_step: int) -> Path:
    return path / f"global_step{global_step}"


def optim_file(tag_dir: Path, rank: int) -> Path:
    return tag_dir / f"zero_pp_rank_{rank}_mp_rank_00_optim_states.pt"


def write_latest(path: Path, global_step: int) -> None:
    (path / "latest").write_text(f"global_step{global_step}")


def read_latest(path: Path) -> str | None:
    p = Path(path) / "latest"
    return p.read_text().strip() if p.exists() else None
