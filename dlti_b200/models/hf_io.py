"""HF Llama checkpoint IO: config.json + (sharded) safetensors.

Reads with `safetensors.safe_open`, which memory-maps and slices: under ZeRO-3 / TP each rank
reads only the byte ranges of its own shard (fixes reference defect App. D #6: every rank loading
the full 13.5 GB model on the host, reference training/train_deepspeed_zero3.py:164-167).

In memory q/k/v and gate/up are merged ([q|k|v, d] and [gate|up, d]); on disk the HF names are
kept: model.layers.{i}.self_attn.{q,k,v,o}_proj.weight, model.layers.{i}.mlp.{gate,up,down}_proj
.weight, input_layernorm, post_attention_layernorm, model.embed_tokens, model.norm, lm_head.
"""
from __future__ import annotations

import json
from pathlib import Path

import torch
from safetensors import safe_open
from safetensors.torch import save_file

from .config import LlamaConfig

_DTYPES = {"bfloat16": torch.bfloat16, "float16": torch.float16, "float32": torch.float32}


def dtype_from_str(s: str) -> torch.dtype:
    return _DTYPES[str(s).replace("torch.", "")]


class CheckpointReader:
    """Name -> tensor access over one or many safetensors shards."""

    def __init__(self, directory: str | Path):
        self.dir = Path(directory)
        idx = self.dir / "model.safetensors.index.json"
        if idx.exists():
            self.weight_map = json.load(open(idx))["weight_map"]
        else:
            single = self.dir / "model.safetensors"
            if not single.exists():
                raise FileNotFoundError(f"no safetensors checkpoint in {self.dir}")
            with safe_open(single, framework="pt") as f:
                self.weight_map = {k: "model.safetensors" for k in f.keys()}
        self._open: dict[str, object] = {}

    def _file(self, name: str):
        fn = self.weight_map[name]
        if fn not in self._open:
            self._open[fn] = safe_open(self.dir / fn, framework="pt")
        return self._open[fn]

    def has(self, name: str) -> bool:
        return name in self.weight_map

    def get(self, name: str, rows: tuple[int, int] | None = None,
            cols: tuple[int, int] | None = None) -> torch.Tensor:
        """Whole tensor or a [rows, cols] slice (only those bytes are read)."""
        sl = self._file(name).get_slice(name)
        if rows is None and cols is None:
            return sl[:]
        if len(sl.get_shape()) == 1:
            return sl[rows[0]:rows[1]]
        r = rows or (0, sl.get_shape()[0])
        c = cols or (0, sl.get_shape()[1])
        return sl[r[0]:r[1], c[0]:c[1]]


def hf_layer_names(i: int) -> dict[str, str]:
    p = f"model.layers.{i}."
    return {"q": p + "self_attn.q_proj.weight", "k": p + "self_attn.k_proj.weight",
            "v": p + "self_attn.v_proj.weight", "o": p + "self_attn.o_proj.weight",
            "gate": p + "mlp.gate_proj.weight", "up": p + "mlp.up_proj.weight",
            "down": p + "mlp.down_proj.weight", "ln1": p + "input_layernorm.weight",
            "ln2": p + "post_attention_layernorm.weight"}


def load_resident(directory: str | Path, device, dtype: torch.dtype | None = None):
    """HF checkpoint directory -> ResidentWeights."""
    from .llama import ResidentWeights

    cfg = LlamaConfig.load(directory)
    dtype = dtype or dtype_from_str(cfg.torch_dtype)
    rd = CheckpointReader(directory)
    w = ResidentWeights(cfg, device, dtype)
    cp = lambda dst, src: dst.copy_(src.to(dtype))
    cp(w.embed, rd.get("model.embed_tokens.weight"))
    cp(w.final_norm, rd.get("model.norm.weight"))
    if not cfg.tie_word_embeddings:
        cp(w.lm_head, rd.get("lm_head.weight") if rd.has("lm_head.weight") else rd.get("model.embed_tokens.weight"))
    q, kv, ff = cfg.q_size, cfg.kv_size, cfg.intermediate_size
    for i, l in enumerate(w.layers):
        n = hf_layer_names(i)
        cp(l.w_qkv[:q], rd.get(n["q"]))
        cp(l.w_qkv[q:q + kv], rd.get(n["k"]))
        cp(l.w_qkv[q + kv:], rd.get(n["v"]))
        cp(l.w_o, rd.get(n["o"]))
        cp(l.w_gu[:ff], rd.get(n["gate"]))
        cp(l.w_gu[ff:], rd.get(n["up"]))
        cp(l.w_down, rd.get(n["down"]))
        cp(l.ln1, rd.get(n["ln1"]))
        cp(l.ln2, rd.get(n["ln2"]))
    return cfg, w


def resident_state_dict(w) -> dict[str, torch.Tensor]:
    """ResidentWeights -> HF-named CPU tensors (q/k/v and gate/up split back out)."""
    cfg = w.cfg
    q, kv, ff = cfg.q_size, cfg.kv_size, cfg.intermediate_size
    c = lambda t: t.detach().to("cpu").contiguous().clone()
    sd = {"model.embed_tokens.weight": c(w.embed), "model.norm.weight": c(w.final_norm)}
    if not cfg.tie_word_embeddings:
        sd["lm_head.weight"] = c(w.lm_head)
    for i, l in enumerate(w.layers):
        n = hf_layer_names(i)
        sd[n["q"]], sd[n["k"]], sd[n["v"]] = c(l.w_qkv[:q]), c(l.w_qkv[q:q + kv]), c(l.w_qkv[q + kv:])
        sd[n["o"]] = c(l.w_o)
        sd[n["gate"]], sd[n["up"]], sd[n["down"]] = c(l.w_gu[:ff]), c(l.w_gu[ff:]), c(l.w_down)
        sd[n["ln1"]], sd[n["ln2"]] = c(l.ln1), c(l.ln2)
    return sd


def save_hf_checkpoint(cfg: LlamaConfig, state_dict: dict[str, torch.Tensor], directory: str | Path,
                       max_shard_bytes: int = 5 * 1024 ** 3) -> list[str]:
    """Write config.json + model-0000i-of-0000N.safetensors + model.safetensors.index.json
    (single `model.safetensors` when everything fits in one shard)."""
    d = Path(directory)
    d.mkdir(parents=True, exist_ok=True)
    cfg.save(d)
    shards: list[dict[str, torch.Tensor]] = [{}]
    size = 0
    for k in sorted(state_dict, key=_hf_order):
        t = state_dict[k]
        nb = t.numel() * t.element_size()
        if size + nb > max_shard_bytes and shards[-1]:
            shards.append({})
            size = 0
        shards[-1][k] = t
        size += nb
    files = []
    if len(shards) == 1:
        save_file(shards[0], str(d / "model.safetensors"), metadata={"format": "pt"})
        files.append("model.safetensors")
    else:
        weight_map = {}
        total = 0
        for i, sh in enumerate(shards):
            fn = f"model-{i + 1:05d}-of-{len(shards):05d}.safetensors"
            save_file(sh, str(d / fn), metadata={"format": "pt"})
            files.append(fn)
            for k, t in sh.items():
                weight_map[k] = fn
                total += t.numel() * t.element_size()
        json.dump({"metadata": {"total_size": total}, "weight_map": weight_map},
                  open(d / "model.safetensors.index.json", "w"), indent=2)
    return files


def _hf_order(name: str):
    parts = name.split(".")
    if "layers" in parts:
        return (1, int(parts[parts.index("layers") + 1]), name)
    return (0 if "embed" in name else 2, 0, name)
