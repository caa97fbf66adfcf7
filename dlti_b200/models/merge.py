"""Adapter merge: W' = W + (alpha/r) * B @ A for every target module, written back in the HF Llama
layout so the server (ours or any HF-compatible one) can load it (the reference only
hints at it through `.gitignore:20` `merged_model/`)."""
from __future__ import annotations

from pathlib import Path

import torch

from .config import LlamaConfig
from .hf_io import load_resident, resident_state_dict, save_hf_checkpoint
from .lora import LoraConfig
from .peft_io import load_adapter_config, load_adapter_state


def merge_into_weights(w, adapter_state: dict[str, torch.Tensor], lcfg: LoraConfig) -> None:
    """In-place merge into a ResidentWeights (fp32 accumulate, cast back to the weight dtype)."""
    cfg: LlamaConfig = w.cfg
    q, kv = cfg.q_size, cfg.kv_size
    rows = {"q_proj": (0, q), "k_proj": (q, q + kv), "v_proj": (q + kv, q + 2 * kv)}
    s = lcfg.scaling

    def find(li, mod, which):
        for k, v in adapter_state.items():
            if f"layers.{li}.self_attn.{mod}.lora_{which}" in k:
                return v
        return None

    for li, l in enumerate(w.layers):
        for mod in lcfg.target_modules:
            A, B = find(li, mod, "A"), find(li, mod, "B")
            if A is None or B is None:
                continue
            delta = (B.to(torch.float32) @ A.to(torch.float32)) * s
            if mod == "o_proj":
                tgt = l.w_o
            else:
                r0, r1 = rows[mod]
                tgt = l.w_qkv[r0:r1]
            tgt.copy_((tgt.to(torch.float32) + delta.to(tgt.device)).to(tgt.dtype))


def merge_checkpoint(base_dir: str | Path, adapter_dir: str | Path, out_dir: str | Path,
                     device="cpu", dtype: torch.dtype | None = None) -> Path:
    cfg, w = load_resident(base_dir, device, dtype)
    lcfg = load_adapter_config(adapter_dir)
    merge_into_weights(w, load_adapter_state(adapter_dir), lcfg)
    save_hf_checkpoint(cfg, resident_state_dict(w), out_dir)
    for fn in ("tokenizer.json", "tokenizer.model", "tokenizer_config.json", "special_tokens_map.json"):
        src = Path(base_dir) / fn
        if src.exists():
            (Path(out_dir) / fn).write_bytes(src.read_bytes())
    return Path(out_dir)
