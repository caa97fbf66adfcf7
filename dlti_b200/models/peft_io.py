"""PEFT-compatible adapter IO: adapter_config.json + adapter_model.safetensors (SURVEY.md D5,
App. C.1/C.2). Format consumers: PEFT itself and vLLM's LoRA loader
(SP:vllm/lora/utils.py:155-195, SP:vllm/lora/peft_helper.py:28-58).

Reference call site: `trainer.save_model(final_dir)` on a PeftModel
(reference training/train_baseline.py:227, train_deepspeed_zero3.py:281).
"""
from __future__ import annotations

import json
from pathlib import Path

import torch
from safetensors.torch import load_file, save_file

from .lora import LoraConfig, LoraParams


def adapter_config_dict(lcfg: LoraConfig) -> dict:
    return {
        "alpha_pattern": {}, "auto_mapping": None,
        "base_model_name_or_path": lcfg.base_model_name_or_path, "bias": lcfg.bias,
        "fan_in_fan_out": False, "inference_mode": True, "init_lora_weights": True,
        "layer_replication": None, "layers_pattern": None, "layers_to_transform": None,
        "loftq_config": {}, "lora_alpha": lcfg.lora_alpha, "lora_dropout": lcfg.lora_dropout,
        "megatron_config": None, "megatron_core": "megatron.core", "modules_to_save": None,
        "peft_type": "LORA", "r": lcfg.r, "rank_pattern": {}, "revision": None,
        "target_modules": sorted(lcfg.target_modules), "task_type": lcfg.task_type,
        "use_dora": False, "use_rslora": False,
    }


def save_adapter(lora_state: dict[str, torch.Tensor], lcfg: LoraConfig, directory: str | Path) -> None:
    d = Path(directory)
    d.mkdir(parents=True, exist_ok=True)
    json.dump(adapter_config_dict(lcfg), open(d / "adapter_config.json", "w"), indent=2)
    sd = {k: v.detach().to(torch.float32).contiguous().cpu() for k, v in lora_state.items()}
    save_file(sd, str(d / "adapter_model.safetensors"), metadata={"format": "pt"})


def load_adapter_config(directory: str | Path) -> LoraConfig:
    c = json.load(open(Path(directory) / "adapter_config.json"))
    if c.get("peft_type", "LORA") != "LORA":
        raise ValueError(f"unsupported peft_type {c.get('peft_type')}")
    for k in ("use_dora", "use_rslora"):
        if c.get(k):
            raise ValueError(f"{k} adapters are not supported")
    tm = c["target_modules"]
    tm = [tm] if isinstance(tm, str) else list(tm)
    return LoraConfig(r=c["r"], lora_alpha=c["lora_alpha"], lora_dropout=c.get("lora_dropout", 0.0),
                      target_modules=tuple(tm), bias=c.get("bias", "none"),
                      task_type=c.get("task_type", "CAUSAL_LM"),
                      base_model_name_or_path=c.get("base_model_name_or_path") or "")


def load_adapter_state(directory: str | Path) -> dict[str, torch.Tensor]:
    return load_file(str(Path(directory) / "adapter_model.safetensors"))


def load_adapter_into(lora: LoraParams, directory: str | Path) -> None:
    lora.load_state_dict(load_adapter_state(directory))
