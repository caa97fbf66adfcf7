"""LoRA adapters: configuration, flat fp32 parameter layout, packed 16-bit compute buffers.

Behaviour reproduced from PEFT 0.13 `LoraConfig` / `lora.Linear` as the reference uses them
(reference training/train_baseline.py:131-140): target q/k/v/o projections,
rank r, alpha = 2r (scaling alpha/r), dropout 0.05 on the LoRA-branch input, A ~ Kaiming-
uniform(a=sqrt 5), B = 0, base weights frozen, adapter master weights kept in fp32.

B200-first design:
  * all trainable parameters live in ONE flat fp32 buffer (and one flat fp32 gradient buffer) —
    this is the unit ZeRO partitions, reduces and steps (no flatten/unflatten copies);
  * the GEMM kernel consumes 16-bit *packed* copies laid out for the fused K-extension
    (csrc/gemm.cu): per layer A_qkv_c [Rq, d] (rows: A_q | A_k | A_v | 0), B_qkv_c [qkv, Rq]
    (block-diagonal, pre-multiplied by alpha/r), A_o_c [Ro, q_size], B_o_c [d, Ro]; they are
    refreshed from the flat master by one batched cast launch after each optimizer step.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch

from ..ops.optim import MultiCast
from .config import LlamaConfig

TARGETS = ("q_proj", "k_proj", "v_proj", "o_proj")


@dataclass
class LoraConfig:
    r: int = 16
    lora_alpha: int = 32
    lora_dropout: float = 0.05
    target_modules: tuple[str, ...] = TARGETS
    bias: str = "none"
    task_type: str = "CAUSAL_LM"
    base_model_name_or_path: str = ""

    def __post_init__(self):
        self.target_modules = tuple(self.target_modules)
        bad = [t for t in self.target_modules if t not in TARGETS]
        if bad:
            raise ValueError(f"unsupported LoRA target modules {bad}; supported: {TARGETS}")
        if self.bias != "none":
            raise ValueError("only bias='none' is supported (as in the reference)")
        if self.r % 8 or self.r <= 0:
            raise ValueError("lora r must be a positive multiple of 8")

    @property
    def scaling(self) -> float:
        return self.lora_alpha / self.r


def _pad64(n: int) -> int:
    return max(64, (n + 63) // 64 * 64)


@dataclass
class _Slot:
    name: str          # e.g. "layers.3.self_attn.q_proj.lora_A"
    offset: int
    shape: tuple[int, int]

    @property
    def numel(self) -> int:
        return self.shape[0] * self.shape[1]


class LoraParams:
    """Flat fp32 master/grad buffers + packed compute buffers for every layer."""

    def __init__(self, cfg: LlamaConfig, lcfg: LoraConfig, device, compute_dtype: torch.dtype,
                 seed: int = 42, pad_to: int = 1, alloc=None):
        """alloc: optional callable (shape, dtype) -> tensor used for the flat master / gradient
        buffers (the fused backend passes its symmetric-arena allocator so peers can address them)."""
        self.cfg, self.lcfg = cfg, lcfg
        self.device = torch.device(device)
        self.compute_dtype = compute_dtype
        r, d = lcfg.r, cfg.hidden_size
        self.in_out = {"q_proj": (d, cfg.q_size), "k_proj": (d, cfg.kv_size),
                       "v_proj": (d, cfg.kv_size), "o_proj": (cfg.q_size, d)}
        self.slots: dict[str, _Slot] = {}
        off = 0
        for li in range(cfg.num_hidden_layers):
            for mod in TARGETS:                     # A blocks first (q, k, v adjacent), then B blocks
                if mod in lcfg.target_modules:
                    n = f"layers.{li}.self_attn.{mod}.lora_A"
                    self.slots[n] = _Slot(n, off, (r, self.in_out[mod][0]))
                    off += self.slots[n].numel
            for mod in TARGETS:
                if mod in lcfg.target_modules:
                    n = f"layers.{li}.self_attn.{mod}.lora_B"
                    self.slots[n] = _Slot(n, off, (self.in_out[mod][1], r))
                    off += self.slots[n].numel
        self.numel = off
        self.padded_numel = (off + pad_to - 1) // pad_to * pad_to
        if alloc is not None:
            self.master = alloc((self.padded_numel,), torch.float32)
            self.grad = alloc((self.padded_numel,), torch.float32)
            self.master.zero_()
            self.grad.zero_()
        else:
            self.master = torch.zeros(self.padded_numel, dtype=torch.float32, device=self.device)
            self.grad = torch.zeros(self.padded_numel, dtype=torch.float32, device=self.device)
        self.reset_parameters(seed)

        # packed compute buffers
        self.n_qkv_targets = sum(m in lcfg.target_modules for m in ("q_proj", "k_proj", "v_proj"))
        self.Rq = _pad64(3 * r)
        self.Ro = _pad64(r)
        L = cfg.num_hidden_layers
        cd = compute_dtype
        self.A_qkv_c = torch.zeros(L, self.Rq, d, dtype=cd, device=self.device)
        self.B_qkv_c = torch.zeros(L, cfg.qkv_size, self.Rq, dtype=cd, device=self.device)
        self.A_o_c = torch.zeros(L, self.Ro, cfg.q_size, dtype=cd, device=self.device)
        self.B_o_c = torch.zeros(L, d, self.Ro, dtype=cd, device=self.device)
        self._build_cast_jobs()
        self.refresh_compute()

    # ---- views ---------------------------------------------------------------------------
    def view(self, name: str, buf: torch.Tensor | None = None) -> torch.Tensor:
        s = self.slots[name]
        b = self.master if buf is None else buf
        return b[s.offset:s.offset + s.numel].view(*s.shape)

    def layer_range(self, li: int) -> tuple[int, int]:
        """[lo, hi) of decoder layer li's adapter tensors in the flat buffers (they are contiguous)."""
        cache = self.__dict__.setdefault("_layer_ranges", {})
        if li not in cache:
            ss = [s for n, s in self.slots.items() if n.startswith(f"layers.{li}.")]
            cache[li] = (min(s.offset for s in ss), max(s.offset + s.numel for s in ss)) if ss else (0, 0)
        return cache[li]

    def has(self, li: int, mod: str) -> bool:
        return f"layers.{li}.self_attn.{mod}.lora_A" in self.slots

    def a_view(self, li: int, mod: str, buf=None):
        return self.view(f"layers.{li}.self_attn.{mod}.lora_A", buf)

    def b_view(self, li: int, mod: str, buf=None):
        return self.view(f"layers.{li}.self_attn.{mod}.lora_B", buf)

    def qkv_col_offset(self, mod: str) -> int:
        """Column (rank) offset of a q/k/v module inside the packed [.., Rq] buffers."""
        return ("q_proj", "k_proj", "v_proj").index(mod) * self.lcfg.r

    def qkv_row_range(self, mod: str) -> tuple[int, int]:
        c = self.cfg
        return {"q_proj": (0, c.q_size), "k_proj": (c.q_size, c.q_size + c.kv_size),
                "v_proj": (c.q_size + c.kv_size, c.qkv_size)}[mod]

    # ---- init ----------------------------------------------------------------------------
    def reset_parameters(self, seed: int) -> None:
        """PEFT init: A ~ kaiming_uniform(a=sqrt(5)) => U(-1/sqrt(in), 1/sqrt(in)); B = 0.
        Generated on CPU from a seeded generator so every rank/device gets identical values."""
        g = torch.Generator(device="cpu").manual_seed(seed)
        self.master.zero_()
        for n, s in self.slots.items():
            if n.endswith("lora_A"):
                bound = 1.0 / math.sqrt(s.shape[1])
                w = (torch.rand(s.shape, generator=g, dtype=torch.float32) * 2 - 1) * bound
                self.view(n).copy_(w)

    # ---- packed compute copies -------------------------------------------------------------
    def _build_cast_jobs(self) -> None:
        jobs = []
        s = self.lcfg.scaling
        r = self.lcfg.r
        for li in range(self.cfg.num_hidden_layers):
            for mod in ("q_proj", "k_proj", "v_proj"):
                if not self.has(li, mod):
                    continue
                c0 = self.qkv_col_offset(mod)
                r0, r1 = self.qkv_row_range(mod)
                jobs.append((self.a_view(li, mod), self.A_qkv_c[li, c0:c0 + r, :], 1.0))
                jobs.append((self.b_view(li, mod), self.B_qkv_c[li, r0:r1, c0:c0 + r], s))
            if self.has(li, "o_proj"):
                jobs.append((self.a_view(li, "o_proj"), self.A_o_c[li, :r, :], 1.0))
                jobs.append((self.b_view(li, "o_proj"), self.B_o_c[li, :, :r], s))
        self._cast = MultiCast(jobs)

    def refresh_compute(self) -> None:
        """fp32 master -> 16-bit packed buffers (B pre-scaled by alpha/r). One launch."""
        self._cast.run()

    def zero_grad(self) -> None:
        self.grad.zero_()

    # ---- bookkeeping -----------------------------------------------------------------------
    def trainable_numel(self) -> int:
        return self.numel

    def state_dict(self) -> dict[str, torch.Tensor]:
        """PEFT key names (adapter name stripped): base_model.model.model.<slot>.weight"""
        return {f"base_model.model.model.{n}.weight": self.view(n).detach().clone().cpu()
                for n in self.slots}

    def load_state_dict(self, sd: dict[str, torch.Tensor], strict: bool = True) -> None:
        seen = set()
        for k, v in sd.items():
            n = k
            for pre in ("base_model.model.model.", "base_model.model.", "model."):
                if n.startswith(pre):
                    n = n[len(pre):]
                    break
            n = n.replace(".default.weight", "").replace(".weight", "")
            if n not in self.slots:
                if strict:
                    raise KeyError(f"unexpected adapter tensor {k}")
                continue
            if tuple(v.shape) != self.slots[n].shape:
                raise ValueError(f"{k}: shape {tuple(v.shape)} != {self.slots[n].shape}")
            self.view(n).copy_(v.to(torch.float32))
            seen.add(n)
        missing = set(self.slots) - seen
        if strict and missing:
            raise KeyError(f"missing adapter tensors: {sorted(missing)[:4]} ...")
        self.refresh_compute()


def print_trainable_parameters(cfg: LlamaConfig, lora: LoraParams) -> str:
    """Same line PEFT prints (reference notebook training/train.ipynb:307)."""
    t = lora.trainable_numel()
    a = cfg.num_params() + t
    line = f"trainable params: {t:,} || all params: {a:,} || trainable%: {100 * t / a:.4f}"
    print(line)
    return line
