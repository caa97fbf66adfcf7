"""Llama architecture config + presets, HF `config.json` compatible.

Reference: the model is whatever `AutoModelForCausalLM.from_pretrained(args.model_name)` loads
(reference training/train_baseline.py:122-126), i.e. transformers' LlamaConfig.
"""
from __future__ import annotations

import json
from dataclasses import asdict, dataclass, field
from pathlib import Path


@dataclass
class LlamaConfig:
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 32
    vocab_size: int = 32000
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    max_position_embeddings: int = 4096
    bos_token_id: int = 1
    eos_token_id: int = 2
    pad_token_id: int | None = None
    tie_word_embeddings: bool = False
    torch_dtype: str = "bfloat16"
    name: str = "llama"

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    @property
    def q_size(self) -> int:
        return self.num_attention_heads * self.head_dim

    @property
    def kv_size(self) -> int:
        return self.num_key_value_heads * self.head_dim

    @property
    def qkv_size(self) -> int:
        return self.q_size + 2 * self.kv_size

    def params_per_layer(self) -> int:
        d, ff = self.hidden_size, self.intermediate_size
        return d * self.qkv_size + self.q_size * d + 3 * d * ff + 2 * d

    def num_params(self) -> int:
        d = self.hidden_size
        n = self.num_hidden_layers * self.params_per_layer() + self.vocab_size * d + d
        if not self.tie_word_embeddings:
            n += self.vocab_size * d
        return n

    # ---- HF config.json ------------------------------------------------------------------
    def to_hf_dict(self) -> dict:
        return {
            "architectures": ["LlamaForCausalLM"],
            "model_type": "llama",
            "hidden_size": self.hidden_size,
            "intermediate_size": self.intermediate_size,
            "num_hidden_layers": self.num_hidden_layers,
            "num_attention_heads": self.num_attention_heads,
            "num_key_value_heads": self.num_key_value_heads,
            "vocab_size": self.vocab_size,
            "rms_norm_eps": self.rms_norm_eps,
            "rope_theta": self.rope_theta,
            "max_position_embeddings": self.max_position_embeddings,
            "bos_token_id": self.bos_token_id,
            "eos_token_id": self.eos_token_id,
            "pad_token_id": self.pad_token_id,
            "hidden_act": "silu",
            "tie_word_embeddings": self.tie_word_embeddings,
            "torch_dtype": self.torch_dtype,
            "attention_bias": False,
            "mlp_bias": False,
            "initializer_range": 0.02,
            "use_cache": True,
        }

    @classmethod
    def from_hf_dict(cls, d: dict, name: str = "llama") -> "LlamaConfig":
        if d.get("hidden_act", "silu") != "silu":
            raise ValueError(f"unsupported hidden_act {d.get('hidden_act')}")
        nh = d["num_attention_heads"]
        return cls(
            hidden_size=d["hidden_size"], intermediate_size=d["intermediate_size"],
            num_hidden_layers=d["num_hidden_layers"], num_attention_heads=nh,
            num_key_value_heads=d.get("num_key_value_heads") or nh, vocab_size=d["vocab_size"],
            rms_norm_eps=d.get("rms_norm_eps", 1e-5), rope_theta=d.get("rope_theta", 10000.0),
            max_position_embeddings=d.get("max_position_embeddings", 4096),
            bos_token_id=d.get("bos_token_id", 1), eos_token_id=d.get("eos_token_id", 2),
            pad_token_id=d.get("pad_token_id"), tie_word_embeddings=d.get("tie_word_embeddings", False),
            torch_dtype=d.get("torch_dtype") or d.get("dtype") or "bfloat16", name=name)

    def save(self, directory: str | Path) -> None:
        Path(directory).mkdir(parents=True, exist_ok=True)
        with open(Path(directory) / "config.json", "w") as f:
            json.dump(self.to_hf_dict(), f, indent=2)

    @classmethod
    def load(cls, directory: str | Path) -> "LlamaConfig":
        with open(Path(directory) / "config.json") as f:
            return cls.from_hf_dict(json.load(f), name=Path(directory).name)


PRESETS: dict[str, LlamaConfig] = {
    # BASELINE.json config 1: tiny random-init Llama (2 layers, d=256) for CPU plumbing.
    # 2 heads x 128 so the same config also runs through the sm_100a attention kernel.
    "tiny": LlamaConfig(hidden_size=256, intermediate_size=704, num_hidden_layers=2,
                        num_attention_heads=2, num_key_value_heads=2, vocab_size=1024,
                        max_position_embeddings=2048, name="tiny"),
    "tiny-gqa": LlamaConfig(hidden_size=512, intermediate_size=1408, num_hidden_layers=2,
                            num_attention_heads=4, num_key_value_heads=2, vocab_size=1024,
                            max_position_embeddings=2048, name="tiny-gqa"),
    "tiny-mha8": LlamaConfig(hidden_size=1024, intermediate_size=2816, num_hidden_layers=2,
                             num_attention_heads=8, num_key_value_heads=8, vocab_size=1024,
                             max_position_embeddings=2048, name="tiny-mha8"),      # TP tests up to 8 ranks
    "tiny-2h": LlamaConfig(hidden_size=256, intermediate_size=704, num_hidden_layers=2,
                           num_attention_heads=2, num_key_value_heads=2, vocab_size=1024,
                           max_position_embeddings=2048, name="tiny-2h"),  # at TP=2: the per-rank shapes of tiny-mha8 at TP=8
    "llama2-7b": LlamaConfig(name="llama2-7b"),
    "llama2-13b": LlamaConfig(hidden_size=5120, intermediate_size=13824, num_hidden_layers=40,
                              num_attention_heads=40, num_key_value_heads=40, name="llama2-13b"),
    "llama2-70b": LlamaConfig(hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                              num_attention_heads=64, num_key_value_heads=8, name="llama2-70b"),
}
_ALIASES = {
    "meta-llama/Llama-2-7b-hf": "llama2-7b",
    "meta-llama/Llama-2-13b-hf": "llama2-13b",
    "meta-llama/Llama-2-70b-hf": "llama2-70b",
}


def resolve_config(name_or_path: str) -> LlamaConfig:
    """Preset name, HF hub id of a known Llama-2 size, or a directory with config.json."""
    import copy

    if name_or_path in PRESETS:
        return copy.deepcopy(PRESETS[name_or_path])
    if name_or_path in _ALIASES:
        return copy.deepcopy(PRESETS[_ALIASES[name_or_path]])
    p = Path(name_or_path)
    if (p / "config.json").exists():
        return LlamaConfig.load(p)
    raise ValueError(f"unknown model config '{name_or_path}' (presets: {sorted(PRESETS)})")
