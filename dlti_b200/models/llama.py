"""Llama causal LM with LoRA adapters — explicit forward/backward over fused ops.

What the reference gets from `transformers.LlamaForCausalLM` + `peft.get_peft_model` + autograd +
`torch.utils.checkpoint` (reference training/train_baseline.py:122-141,181)
is restated here as a hand-scheduled training step:

  * activations are token-major [T, width] (sequences packed, described by cu_seqlens);
  * q/k/v and gate/up projections are merged GEMMs; LoRA is fused into the base GEMM as a
    K-extension (ops.gemm a2/b2); residual adds ride in GEMM epilogues;
  * base weights are frozen (no dW): backward computes only dX through W (MN-major operand of
    the same weight buffer) plus the small dA/dB GEMMs, accumulated in fp32 straight into the
    flat gradient buffer that ZeRO partitions;
  * layer-granular recompute: forward keeps only each layer's input; backward re-runs the layer
    forward (same dropout masks: counter-based RNG) and then its backward.

Weights come from a *provider* (`ResidentWeights` here, `parallel.zero3.Zero3Weights` for
ZeRO-3) exposing layer(i)/embed/final_norm/lm_head plus prefetch/release hooks.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

from ..ops import attention as A
from ..ops import elementwise as E
from ..ops.gemm import gemm, gemm_rope, gemm_swiglu, gemm_swiglu_bwd
from ..ops.loss import cross_entropy
from .config import LlamaConfig
from .lora import LoraConfig, LoraParams


@dataclass
class LayerWeights:
    w_qkv: torch.Tensor    # [q+2kv, d]   rows: q | k | v
    w_o: torch.Tensor      # [d, q]
    w_gu: torch.Tensor     # [2*ff, d]    rows: gate | up
    w_down: torch.Tensor   # [d, ff]
    ln1: torch.Tensor      # [d]
    ln2: torch.Tensor      # [d]
    gather: dict | None = None   # fused ZeRO-3: name -> GatherSpec (weights are staging buffers the
                                 # GEMM itself fills from the owner ranks, K1)

    def g(self, name: str):
        """`gather=` handle for a GEMM whose B operand is `name` (fused ZeRO-3: the launch carries a slice of
        the queued weight pulls and, if its own operand is still incomplete, waits on it tile by tile)."""
        return self.gather[name] if self.gather else None


class ResidentWeights:
    """Frozen base weights fully resident on one device (ZeRO stages 0-2, single-GPU serving)."""

    def __init__(self, cfg: LlamaConfig, device, dtype: torch.dtype):
        self.cfg, self.device, self.dtype = cfg, torch.device(device), dtype
        d, ff, L = cfg.hidden_size, cfg.intermediate_size, cfg.num_hidden_layers
        mk = lambda *s: torch.empty(*s, dtype=dtype, device=self.device)
        self.embed = mk(cfg.vocab_size, d)
        self.final_norm = mk(d)
        self.lm_head = self.embed if cfg.tie_word_embeddings else mk(cfg.vocab_size, d)
        self.layers = [LayerWeights(mk(cfg.qkv_size, d), mk(d, cfg.q_size), mk(2 * ff, d),
                                    mk(d, ff), mk(d), mk(d)) for _ in range(L)]

    @classmethod
    def random_init(cls, cfg: LlamaConfig, device, dtype, seed: int = 0) -> "ResidentWeights":
        """HF initializer_range=0.02 normal init, norms = 1 (random-init weights of the named
        architecture: there is no network for real checkpoints on the bench box)."""
        w = cls(cfg, device, dtype)
        g = torch.Generator(device=w.device).manual_seed(seed)
        for t in w.all_matrices():
            t.normal_(0.0, 0.02, generator=g)
        for t in w.all_norms():
            t.fill_(1.0)
        return w

    def all_matrices(self):
        yield self.embed
        if self.lm_head is not self.embed:
            yield self.lm_head
        for l in self.layers:
            yield from (l.w_qkv, l.w_o, l.w_gu, l.w_down)

    def all_norms(self):
        yield self.final_norm
        for l in self.layers:
            yield from (l.ln1, l.ln2)

    # provider interface --------------------------------------------------------------------
    def layer(self, i: int) -> LayerWeights:
        return self.layers[i]

    def prefetch(self, i: int) -> None:  # ZeRO-3 hook (no-op when resident)
        pass

    def release(self, i: int) -> None:
        pass

    def prefetch_head(self) -> None:
        pass

    def release_head(self) -> None:
        pass

    def head_gather(self):
        return None

    def get_embed(self):
        return self.embed

    def get_final_norm(self):
        return self.final_norm

    def get_lm_head(self):
        return self.lm_head

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in list(self.all_matrices()) + list(self.all_norms()))


@dataclass
class PackedBatch:
    """One micro-batch, sequences concatenated along T."""
    input_ids: torch.Tensor        # int64 [T]
    labels: torch.Tensor           # int64 [T]  already shifted: label of token t = id of t+1, -100 = ignore
    positions: torch.Tensor        # int32 [T]
    cu_seqlens: torch.Tensor       # int32 [B+1]
    max_seqlen: int
    n_valid: int                   # number of labels != -100 (host-side count)

    def to(self, device, non_blocking=True) -> "PackedBatch":
        f = lambda t: t.to(device, non_blocking=non_blocking)
        return PackedBatch(f(self.input_ids), f(self.labels), f(self.positions), f(self.cu_seqlens),
                           self.max_seqlen, self.n_valid)

    @property
    def n_tokens(self) -> int:
        return int(self.input_ids.numel())


def pack_batch(input_ids: torch.Tensor, labels: torch.Tensor, attention_mask: torch.Tensor | None,
               pin: bool = False) -> PackedBatch:
    """[B, S] padded batch (HF collator layout) -> packed, label-shifted token stream (CPU)."""
    B, S = input_ids.shape
    if attention_mask is None:
        attention_mask = torch.ones_like(input_ids)
    lens = attention_mask.sum(1).to(torch.int64)
    shifted = torch.full_like(labels, -100)
    shifted[:, :-1] = labels[:, 1:]
    ar = torch.arange(S).unsqueeze(0)
    shifted = torch.where(ar < (lens - 1).unsqueeze(1), shifted, torch.full_like(shifted, -100))
    keep = ar < lens.unsqueeze(1)                       # right padding (HF default for Llama)
    ids = input_ids[keep].contiguous().to(torch.int64)
    lab = shifted[keep].contiguous().to(torch.int64)
    pos = ar.expand(B, S)[keep].contiguous().to(torch.int32)
    cu = torch.zeros(B + 1, dtype=torch.int32)
    cu[1:] = lens.cumsum(0).to(torch.int32)
    out = PackedBatch(ids, lab, pos, cu, int(lens.max().item()), int((lab != -100).sum().item()))
    if pin and torch.cuda.is_available():
        out = PackedBatch(ids.pin_memory(), lab.pin_memory(), pos.pin_memory(), cu.pin_memory(),
                          out.max_seqlen, out.n_valid)
    return out


class LoraLlama:
    """Frozen Llama + LoRA with a hand-written training step and a plain forward."""

    def __init__(self, cfg: LlamaConfig, weights, lora: LoraParams | None = None,
                 gradient_checkpointing: bool = True, dropout_seed: int = 42):
        self.cfg, self.w, self.lora = cfg, weights, lora
        self.gradient_checkpointing = gradient_checkpointing
        self.dropout_seed = dropout_seed
        self.micro_step = 0          # advances once per forward_backward (dropout stream id)
        self.training = True
        self.scale = 1.0 / math.sqrt(cfg.head_dim)
        self._rope = None
        # called as grad_hook(flat_offset, numel) when a decoder layer's adapter gradients are final for the
        # accumulation window (ZeRO: the reduce-scatter of that range can start while backward continues)
        self.grad_hook = None

    # ---- helpers ---------------------------------------------------------------------------
    def _rope_tables(self, device):
        if self._rope is None or self._rope[0].device != torch.device(device):
            self._rope = E.rope_tables(self.cfg.max_position_embeddings, self.cfg.head_dim,
                                       self.cfg.rope_theta, device)
        return self._rope

    def _drop(self) -> float:
        return self.lora.lcfg.lora_dropout if (self.lora is not None and self.training) else 0.0

    def _stream(self, li: int, site: int) -> int:
        return (self.micro_step << 20) | (li << 4) | site

    # ---- one decoder layer -----------------------------------------------------------------
    def _layer_fwd(self, li: int, lw: LayerWeights, x: torch.Tensor, b: PackedBatch, save: bool,
                   need_out: bool = True):
        """need_out=False (recompute inside backward): stop after gate/up — SwiGLU and the down
        projection only feed the layer output, which backward never reads."""
        cfg, lo = self.cfg, self.lora
        p = self._drop()
        H, KV, hd = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        cos, sin = self._rope_tables(x.device)
        use_qkv = lo is not None and lo.n_qkv_targets > 0
        use_o = lo is not None and lo.has(li, "o_proj")

        xn, rstd1, xn_drop = E.rmsnorm_fwd(x, lw.ln1, cfg.rms_norm_eps, drop_p=p,
                                           seed=self.dropout_seed, stream=self._stream(li, 0),
                                           want_drop=use_qkv and p > 0)
        if xn_drop is None:
            xn_drop = xn
        t_qkv = gemm(xn_drop, lo.A_qkv_c[li]) if use_qkv else None            # [T, Rq]
        # RoPE applied in the projection's epilogue
        qkv = gemm_rope(xn, lw.w_qkv, b.positions, cos, sin, H + KV, hd, a2=t_qkv,
                        b2=lo.B_qkv_c[li] if use_qkv else None, gather=lw.g("w_qkv"))
        attn, lse = A.attention_fwd(qkv, b.cu_seqlens, b.max_seqlen, H, KV, hd, self.scale)
        attn_drop = E.dropout(attn, p, self.dropout_seed, self._stream(li, 1)) if use_o else attn
        t_o = gemm(attn_drop, lo.A_o_c[li]) if use_o else None                 # [T, Ro]
        x2 = gemm(attn, lw.w_o, a2=t_o, b2=lo.B_o_c[li] if use_o else None, residual=x, gather=lw.g("w_o"))
        xn2, rstd2, _ = E.rmsnorm_fwd(x2, lw.ln2, cfg.rms_norm_eps)
        if need_out and not save:
            # forward of a checkpointed layer: nothing downstream needs gate|up itself, so SwiGLU runs in
            # the GEMM epilogue and the [T, 2*ff] tensor is never written
            gu = None
            h = gemm_swiglu(xn2, lw.w_gu, gather=lw.g("w_gu"))
        else:
            gu = gemm(xn2, lw.w_gu, gather=lw.g("w_gu"))
            h = E.swiglu_fwd(gu) if need_out else None
        if need_out:
            x3 = gemm(h, lw.w_down, residual=x2, gather=lw.g("w_down"))
        else:
            x3 = None
        if not save:
            return x3, None
        saved = dict(x=x, rstd1=rstd1, xn=xn, xn_drop=xn_drop, t_qkv=t_qkv, qkv=qkv, attn=attn,
                     lse=lse, attn_drop=attn_drop, t_o=t_o, x2=x2, rstd2=rstd2, xn2=xn2, gu=gu)
        return x3, saved

    def _layer_bwd(self, li: int, lw: LayerWeights, dx3: torch.Tensor, s: dict, b: PackedBatch):
        cfg, lo = self.cfg, self.lora
        p = self._drop()
        H, KV, hd = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        cos, sin = self._rope_tables(dx3.device)
        r = lo.lcfg.r if lo is not None else 0
        sc = lo.lcfg.scaling if lo is not None else 1.0
        use_qkv = lo is not None and lo.n_qkv_targets > 0
        use_o = lo is not None and lo.has(li, "o_proj")
        T = dx3.shape[0]
        splitk = max(1, min(16, T // 512))

        # ---- MLP ----
        dgu = gemm_swiglu_bwd(dx3, lw.w_down, s["gu"], gather=lw.g("w_down"))   # SwiGLU backward in the epilogue
        dxn2 = gemm(dgu, lw.w_gu, b_mn=True, gather=lw.g("w_gu"))                               # [T, d]
        del dgu
        dx2 = E.rmsnorm_bwd(dxn2, s["x2"], lw.ln2, s["rstd2"], dres=dx3)
        del dxn2
        # ---- o_proj (+LoRA) ----
        if use_o:
            u_o = gemm(dx2, lo.B_o_c[li], b_mn=True)                       # [T, Ro] = s * dx2 · B_o
            if p > 0:
                dattn = gemm(dx2, lw.w_o, b_mn=True, gather=lw.g("w_o"))
                dl = gemm(u_o, lo.A_o_c[li], b_mn=True)
                dattn = E.add_dropout(dattn, dl, p, self.dropout_seed, self._stream(li, 1))
                del dl
            else:
                dattn = gemm(dx2, lw.w_o, b_mn=True, a2=u_o, b2=lo.A_o_c[li], gather=lw.g("w_o"))
            gB = lo.b_view(li, "o_proj", lo.grad)
            gemm(dx2, s["t_o"][:, :r], a_mn=True, b_mn=True, out=gB, accumulate=True, alpha=sc,
                 split_k=splitk, block_n=64)
            gA = lo.a_view(li, "o_proj", lo.grad)
            gemm(u_o[:, :r], s["attn_drop"], a_mn=True, b_mn=True, out=gA, accumulate=True,
                 split_k=splitk, block_n=128)
        else:
            dattn = gemm(dx2, lw.w_o, b_mn=True, gather=lw.g("w_o"))
        # ---- attention ----
        dqkv = A.attention_bwd(dattn, s["qkv"], s["attn"], s["lse"], b.cu_seqlens, b.max_seqlen,
                               H, KV, hd, self.scale)
        del dattn
        E.rope_(dqkv, b.positions, cos, sin, H + KV, hd, backward=True)
        # ---- qkv (+LoRA) ----
        if use_qkv:
            u_qkv = gemm(dqkv, lo.B_qkv_c[li], b_mn=True)                  # [T, Rq]
            if p > 0:
                dxn = gemm(dqkv, lw.w_qkv, b_mn=True, gather=lw.g("w_qkv"))
                dl = gemm(u_qkv, lo.A_qkv_c[li], b_mn=True)
                dxn = E.add_dropout(dxn, dl, p, self.dropout_seed, self._stream(li, 0))
                del dl
            else:
                dxn = gemm(dqkv, lw.w_qkv, b_mn=True, a2=u_qkv, b2=lo.A_qkv_c[li], gather=lw.g("w_qkv"))
            first = None
            for mod in ("q_proj", "k_proj", "v_proj"):
                if not lo.has(li, mod):
                    continue
                c0 = lo.qkv_col_offset(mod)
                r0, r1 = lo.qkv_row_range(mod)
                gemm(dqkv[:, r0:r1], s["t_qkv"][:, c0:c0 + r], a_mn=True, b_mn=True,
                     out=lo.b_view(li, mod, lo.grad), accumulate=True, alpha=sc, split_k=splitk,
                     block_n=64)
                first = mod if first is None else first
            # dA for q,k,v in ONE GEMM when the three A blocks are adjacent in the flat buffer
            mods = [m for m in ("q_proj", "k_proj", "v_proj") if lo.has(li, m)]
            if len(mods) == 3:
                s0 = lo.slots[f"layers.{li}.self_attn.q_proj.lora_A"]
                gA = lo.grad[s0.offset:s0.offset + 3 * r * cfg.hidden_size].view(3 * r, cfg.hidden_size)
                gemm(u_qkv[:, :3 * r], s["xn_drop"], a_mn=True, b_mn=True, out=gA, accumulate=True,
                     split_k=splitk, block_n=128)
            else:
                for mod in mods:
                    c0 = lo.qkv_col_offset(mod)
                    gemm(u_qkv[:, c0:c0 + r], s["xn_drop"], a_mn=True, b_mn=True,
                         out=lo.a_view(li, mod, lo.grad), accumulate=True, split_k=splitk, block_n=128)
        else:
            dxn = gemm(dqkv, lw.w_qkv, b_mn=True, gather=lw.g("w_qkv"))
        del dqkv
        dx = E.rmsnorm_bwd(dxn, s["x"], lw.ln1, s["rstd1"], dres=dx2)
        return dx

    # ---- full passes -----------------------------------------------------------------------
    def forward_hidden(self, b: PackedBatch) -> torch.Tensor:
        """Final-norm hidden states [T, d] (no KV cache; eval / scoring path)."""
        x = E.embedding(b.input_ids, self.w.get_embed())
        L = self.cfg.num_hidden_layers
        self.w.prefetch(0)
        for li in range(L):
            if li + 1 < L:
                self.w.prefetch(li + 1)
            x, _ = self._layer_fwd(li, self.w.layer(li), x, b, save=False)
            self.w.release(li)
        xf, _, _ = E.rmsnorm_fwd(x, self.w.get_final_norm(), self.cfg.rms_norm_eps)
        return xf

    def forward_logits(self, b: PackedBatch) -> torch.Tensor:
        xf = self.forward_hidden(b)
        logits = gemm(xf, self.w.get_lm_head(), gather=self.w.head_gather())
        self.w.release_head()
        return logits

    def forward_backward(self, b: PackedBatch, gscale: torch.Tensor | float) -> torch.Tensor:
        """One micro-batch: accumulates LoRA gradients of  gscale * sum_t CE_t  into lora.grad and
        returns sum_t CE_t (fp32 scalar tensor on the device, unscaled).

        gscale = loss_scale / (number of target tokens in the whole accumulated global batch): the
        token-normalised form. A 1-element fp32 device tensor on CUDA."""
        return self.forward_backward_multi([b], [gscale])

    def forward_backward_multi(self, batches: list[PackedBatch], gscales: list) -> torch.Tensor:
        """A whole gradient-accumulation window, scheduled LAYER-MAJOR: for each decoder layer the
        weights are fetched once and every micro-batch runs through it (forward), and again once
        for recompute + backward in reverse layer order. Per-micro-batch math, dropout streams and
        the accumulated gradient are exactly those of running the micro-batches one after another;
        what changes is that a ZeRO-3 layer is gathered 2x per optimizer step instead of 2x per
        micro-batch. The price is keeping every micro-batch's layer inputs (L x T x d each), which
        180 GB of HBM absorbs easily (7B, 4 x 4096 tokens: 4.3 GB).

        Returns the sum over all micro-batches of sum_t CE_t (unscaled, device scalar)."""
        assert self.lora is not None and len(batches) == len(gscales)
        cfg, L, nb = self.cfg, self.cfg.num_hidden_layers, len(batches)
        ckpt = self.gradient_checkpointing
        base_step = self.micro_step
        xs = [E.embedding(b.input_ids, self.w.get_embed()) for b in batches]
        inputs = [[None] * L for _ in range(nb)]
        saves = [[None] * L for _ in range(nb)]
        self.w.prefetch(0)
        for li in range(L):
            if li + 1 < L:
                self.w.prefetch(li + 1)
            lw = self.w.layer(li)
            if li == L - 1:
                self.w.prefetch_head()      # the last layer's GEMMs carry the LM head's shards (fused ZeRO-3)
            for k, b in enumerate(batches):
                self.micro_step = base_step + k
                inputs[k][li] = xs[k]
                xs[k], saves[k][li] = self._layer_fwd(li, lw, xs[k], b, save=not ckpt)
            self.w.release(li)
        # ---- head per micro-batch: final norm -> LM head -> CE (in-place gradient) -> dX ----
        lm = self.w.get_lm_head()
        hg = self.w.head_gather()
        fnorm = self.w.get_final_norm()
        loss_sum = None
        dxs = []
        for k, b in enumerate(batches):
            xf, rstd_f, _ = E.rmsnorm_fwd(xs[k], fnorm, cfg.rms_norm_eps)
            logits = gemm(xf, lm, gather=hg)
            loss_rows, dlogits = cross_entropy(logits, b.labels, gscales[k], inplace=True)
            ls = loss_rows.sum()
            loss_sum = ls if loss_sum is None else loss_sum + ls
            dxf = gemm(dlogits, lm, b_mn=True, gather=hg)
            del logits, dlogits
            dxs.append(E.rmsnorm_bwd(dxf, xs[k], fnorm, rstd_f))
            xs[k] = None
            del dxf, xf
        self.w.release_head()
        # ---- layers in reverse, recomputing each layer's forward ----
        self.w.prefetch(L - 1)
        for li in reversed(range(L)):
            if li > 0:
                self.w.prefetch(li - 1)
            lw = self.w.layer(li)
            for k, b in enumerate(batches):
                self.micro_step = base_step + k
                sv = saves[k][li]
                if sv is None:
                    _, sv = self._layer_fwd(li, lw, inputs[k][li], b, save=True, need_out=False)
                dxs[k] = self._layer_bwd(li, lw, dxs[k], sv, b)
                saves[k][li] = None
                inputs[k][li] = None
            self.w.release(li)
            if self.grad_hook is not None:
                lo_, hi_ = self.lora.layer_range(li)
                if hi_ > lo_:
                    self.grad_hook(lo_, hi_ - lo_)
        self.micro_step = base_step + nb
        return loss_sum
