"""Inference forward pass over the paged KV cache, single GPU or Megatron-style tensor parallel
(§2.7 S1-S12).

TP layout per rank (tp = world size): q/k/v heads and gate/up columns are column-parallel, o_proj and
down_proj are row-parallel, the LM head is vocab-parallel; the embedding table is replicated (262 MB
at 7B: on 180 GB parts that removes collective C8 entirely). The two row-parallel GEMMs of a layer
produce partial sums that are combined by `TPContext.reduce_add_norm`: on the "fused" backend the
GEMM epilogue pushes its partial tile straight into every peer's receive slot over NVLink (K10) and
one kernel sums the slots in rank order, adds the residual and applies RMSNorm (the all-gather side of
K11 is implicit: every rank ends up with the full normalised activations); on the "nccl" backend it is
NCCL all-reduce + the fused add+norm kernel.

A step's token rows are [decode rows | prompt rows] (either part may be empty): the GEMMs, norms and the
RoPE + KV-cache-write kernel run over all rows at once, the decode rows go through the paged-KV split-KV
attention kernel and the packed prompts through the varlen tcgen05 flash-attention kernel.
"""
from __future__ import annotations

import math
import os

import torch

from ..models.config import LlamaConfig
from ..ops import attention as A
from ..ops import elementwise as E
from ..ops import serving as S
from ..ops.gemm import gemm, gemm_skinny, gemm_skinny_acc, gemm_swiglu
from .kv_cache import KVCache


_FUSE_ACC = os.environ.get("DLTI_DECODE_FUSE_ACC", "1") == "1"    # A/B switch: consume the fp32 accumulators directly


# opt-in until measured on the serving shapes: prefill gate_up GEMM with SwiGLU in its epilogue (the training kernel)
_PREFILL_FUSE_SWIGLU = os.environ.get("DLTI_SERVE_FUSE_SWIGLU", "0") == "1"
_TP_LARGE = os.environ.get("DLTI_TP_LARGE_RSAG", "1") == "1"          # prefill-sized steps: reduce-scatter + all-gather
# how prefill-sized steps do it: "fused" = reduce-scatter in the row-parallel GEMM's epilogue (K10) and all-gather in
# the next column-parallel GEMM's producer (K11); "pull" = the round-1 stand-alone pull-reduce-push kernel
_TP_LARGE_MODE = os.environ.get("DLTI_TP_LARGE_MODE", "fused")


class ShardedAct:
    """Normalised activations of a prefill-sized TP step whose rows still live on their owner ranks: the next
    column-parallel GEMM gathers them itself (`handle`, K11); `rows()` is that GEMM's A operand."""

    def __init__(self, T: int, handle):
        self.T, self.handle = T, handle
        self.dtype, self.device = handle.full.dtype, handle.full.device

    def rows(self) -> torch.Tensor:
        return self.handle.full[: self.T]

    def gathered(self) -> torch.Tensor:
        self.handle.materialize()
        return self.handle.full[: self.T]
_FUSE_TP_BARRIER = os.environ.get("DLTI_TP_FUSE_BARRIER", "1") == "1"  # rank barrier inside the reduce+norm kernel


class TPContext:
    """Collectives used by the TP forward. world == 1 => no-ops."""

    def __init__(self, info=None, backend: str = "nccl", fused=None, cfg: LlamaConfig | None = None,
                 max_tokens: int = 8192, dtype=torch.bfloat16):
        self.rank = info.rank if info else 0
        self.world = info.world_size if info else 1
        self.backend, self.fused = backend, fused
        self.use_fused = fused is not None and self.world > 1
        self.fused_large = False
        if self.use_fused:
            d = cfg.hidden_size
            # steps of more than SMALL rows (prefill) reduce-scatter + all-gather through one partial and one
            # activation buffer per rank; smaller steps (decode) push their partial tile to every peer's
            # receive slots: two alternating buffers [world, rows, d] in the symmetric arena
            self.large = _TP_LARGE
            self.small_rows = 256 if self.large else max_tokens
            self.recv = [fused.alloc_symmetric((self.world, self.small_rows, d), dtype) for _ in range(2)]
            self.flip = 0
            self.max_tokens = max_tokens
            if self.large:
                self.partial = fused.alloc_symmetric((max_tokens, d), dtype)
                self.xn = fused.alloc_symmetric((max_tokens, d), dtype)
                self.fused_large = _TP_LARGE_MODE == "fused" and d % 64 == 0
                if self.fused_large:
                    self.per_max = -(-(-(-max_tokens // self.world)) // 64) * 64
                    self.slots = fused.alloc_symmetric((self.world * self.per_max, d), dtype)   # [src rank][row]
                    self.xn_own = fused.alloc_symmetric((self.per_max, d), dtype)
                    self.xn_full = torch.empty(self.world * self.per_max, d, dtype=dtype, device=info.device)
                    self.ctr_pool = torch.zeros(1 << 16, dtype=torch.int32, device=info.device)
                    self.ctr_off = 0

    def adopt(self, w: "TPWeights", cfg: LlamaConfig, max_rows: int = 1024) -> None:
        """Fused backend: move the pieces of the model that are addressed across ranks into the symmetric arena.
        (collective: every rank calls it in the same order)
          * K15: the embedding table becomes vocabulary-sharded (1/tp of it per rank) and is read in place on the
            owner ranks by the embedding kernel (parallel.fused.PeerRows) instead of being replicated;
          * C11: a [max_rows, vocab] logits buffer per rank that every rank's LM-head GEMM epilogue writes its
            vocabulary slice into (peer stores), replacing the NCCL all-gather of the logits."""
        if not self.use_fused or hasattr(w.embed, "lookup") or os.environ.get("DLTI_TP_ADOPT", "1") != "1":
            return
        import torch.distributed as dist
        from ..parallel.fused import PeerRows
        V, d = cfg.vocab_size, cfg.hidden_size
        Vr = V // self.world
        sh = self.fused.alloc_symmetric((Vr, d), w.embed.dtype)
        sh.copy_(w.embed[self.rank * Vr:(self.rank + 1) * Vr])
        self.logits_rows = max_rows
        self.logits_buf = self.fused.alloc_symmetric((max_rows, V), w.embed.dtype)
        torch.cuda.synchronize()
        dist.barrier()
        w.embed = PeerRows(self.fused, sh, V)

    def lm_head_logits(self, xl: torch.Tensor, w_head: torch.Tensor) -> torch.Tensor:
        """[B, vocab] logits on every rank from the vocabulary-parallel LM head."""
        B = xl.shape[0]
        if self.world == 1:
            return (gemm_skinny if B <= 256 else gemm)(xl, w_head)
        if self.use_fused and getattr(self, "logits_buf", None) is not None and B <= self.logits_rows:
            buf = self.logits_buf
            Vl, V = w_head.shape[0], buf.shape[1]
            # the epilogue stores this rank's [B, V/tp] slice into EVERY rank's buffer at column rank * V/tp
            gemm(xl, w_head, push=(self.fused.peer_array(buf), buf, self.rank, self.world, self.logits_rows, V, Vl))
            self.fused.barrier_graph_safe(7)
            return buf[:B]
        return self.all_gather_cols((gemm_skinny if B <= 256 else gemm)(xl, w_head))

    def reduce_add_norm(self, partial_fn, T: int, residual, w, eps):
        """partial_fn(out=None|tensor, push=None|spec) runs the row-parallel GEMM. Returns the normalised
        activations; `residual` (fp [T, d]) is updated in place with residual += sum_r partial_r."""
        if self.world == 1:
            part = partial_fn(None, None)
            y, _, _ = E.rmsnorm_fwd(part, w, eps, residual=residual)
            return y
        if self.use_fused and T > self.small_rows and self.fused_large:
            # K10: the GEMM's epilogue scatters each output row to its owner's receive slot | barrier | the owner
            # sums its slots, adds the residual, normalises (local memory only) | barrier | K11: the NEXT
            # column-parallel GEMM pulls the normalised rows from their owners while it computes
            from ..parallel.fused import ActGather, tp_own_reduce_norm
            assert T <= self.max_tokens
            self.fused_large_calls = getattr(self, "fused_large_calls", 0) + 1
            per = -(-(-(-T // self.world)) // 64) * 64
            partial_fn(self.slots[:T], None, scatter=(self.fused.peer_array(self.slots), self.world, self.rank, per))
            self.fused.barrier_graph_safe(6)
            tp_own_reduce_norm(self.fused, self.slots, per, T, residual, w, eps, self.xn_own)
            self.fused.barrier_graph_safe(6)
            nu = self.world * per // 64
            if self.ctr_off + nu > self.ctr_pool.numel():
                self.ctr_pool.zero_()
                self.ctr_off = 0
            ctr = self.ctr_pool[self.ctr_off:self.ctr_off + nu]
            self.ctr_off += nu
            return ShardedAct(T, ActGather(self.fused, self.xn_own, self.xn_full, ctr, per))
        if self.use_fused and T > self.small_rows:
            # reduce-scatter + all-gather over peer memory: partial via the plain (fast) GEMM into this rank's
            # symmetric buffer | barrier | pull my rows from all ranks, add residual, norm, push to all | barrier
            from ..parallel.fused import tp_pull_reduce_norm_push
            assert T <= self.max_tokens
            partial_fn(self.partial[:T], None)
            self.fused.barrier_graph_safe(6)
            tp_pull_reduce_norm_push(self.fused, self.partial, self.xn, T, residual, w, eps)
            self.fused.barrier_graph_safe(6)
            return self.xn[:T]
        if self.use_fused:
            from ..parallel.fused import tp_reduce_add_norm
            buf = self.recv[self.flip]
            self.flip ^= 1
            partial_fn(None, (self.fused.peer_array(buf), buf, self.rank, self.world, self.small_rows))
            if _FUSE_TP_BARRIER:
                return tp_reduce_add_norm(buf, T, residual, w, eps, self.world, self.small_rows, barrier=(self.fused, 5))
            self.fused.barrier_graph_safe(5)
            return tp_reduce_add_norm(buf, T, residual, w, eps, self.world, self.small_rows)
        import torch.distributed as dist
        part = partial_fn(None, None)
        dist.all_reduce(part)
        y, _, _ = E.rmsnorm_fwd(part, w, eps, residual=residual)
        return y

    def all_gather_cols(self, local: torch.Tensor) -> torch.Tensor:
        if self.world == 1:
            return local
        import torch.distributed as dist
        rows, cols = local.shape
        out = torch.empty(self.world * rows, cols, dtype=local.dtype, device=local.device)   # rank-major concatenation
        dist.all_gather_into_tensor(out, local.contiguous())
        return out.view(self.world, rows, cols).permute(1, 0, 2).reshape(rows, -1)


class TPWeights:
    """This rank's slice of every weight, built from a ResidentWeights-like provider or random init."""

    def __init__(self, cfg: LlamaConfig, rank: int, world: int, device, dtype):
        self.cfg, self.rank, self.world = cfg, rank, world
        assert cfg.num_attention_heads % world == 0 and cfg.num_key_value_heads % world == 0, \
            "tp size must divide the number of q and kv heads"
        assert cfg.intermediate_size % world == 0 and cfg.vocab_size % world == 0
        self.H, self.KV = cfg.num_attention_heads // world, cfg.num_key_value_heads // world
        self.ff = cfg.intermediate_size // world
        self.V = cfg.vocab_size // world
        d, hd = cfg.hidden_size, cfg.head_dim
        mk = lambda *s: torch.empty(*s, dtype=dtype, device=device)
        self.embed = mk(cfg.vocab_size, d)
        self.final_norm = mk(d)
        self.lm_head = mk(self.V, d)
        self.lora = None          # unmerged PEFT adapter (attach_adapter)
        self.layers = [dict(w_qkv=mk((self.H + 2 * self.KV) * hd, d), w_o=mk(d, self.H * hd),
                            w_gu=mk(2 * self.ff, d), w_down=mk(d, self.ff), ln1=mk(d), ln2=mk(d))
                       for _ in range(cfg.num_hidden_layers)]

    def load_from_full(self, w) -> "TPWeights":
        """Slice a full ResidentWeights (any device) into this rank's shards."""
        c, r, hd = self.cfg, self.rank, self.cfg.head_dim
        q, kv, ff = c.q_size, c.kv_size, c.intermediate_size
        Hl, KVl, ffl = self.H * hd, self.KV * hd, self.ff
        self.embed.copy_(w.embed)
        self.final_norm.copy_(w.final_norm)
        self.lm_head.copy_(w.lm_head[r * self.V:(r + 1) * self.V])
        for dst, src in zip(self.layers, w.layers):
            dst["w_qkv"][:Hl].copy_(src.w_qkv[r * Hl:(r + 1) * Hl])
            dst["w_qkv"][Hl:Hl + KVl].copy_(src.w_qkv[q + r * KVl:q + (r + 1) * KVl])
            dst["w_qkv"][Hl + KVl:].copy_(src.w_qkv[q + kv + r * KVl:q + kv + (r + 1) * KVl])
            dst["w_o"].copy_(src.w_o[:, r * Hl:(r + 1) * Hl])
            dst["w_gu"][:ffl].copy_(src.w_gu[r * ffl:(r + 1) * ffl])
            dst["w_gu"][ffl:].copy_(src.w_gu[ff + r * ffl:ff + (r + 1) * ffl])
            dst["w_down"].copy_(src.w_down[:, r * ffl:(r + 1) * ffl])
            dst["ln1"].copy_(src.ln1)
            dst["ln2"].copy_(src.ln2)
        return self

    def attach_adapter(self, adapter_state: dict, lcfg) -> "TPWeights":
        """Serve a PEFT LoRA adapter UNMERGED on top of the base weights (what vLLM's LoRA loader does):
        per layer, packed 16-bit operands for the LoRA branch of q/k/v (a K-extension of the QKV GEMM, as in
        training) and of o_proj (extra K columns of the o_proj GEMM: its input becomes [attention | t_o] and its
        weight [W_o | (alpha/r) B_o], so every o_proj kernel — decode, push, scatter — serves the adapter without
        knowing about it). TP slicing: B_q/k/v rows and A_o columns follow this rank's heads; A_q/k/v and B_o are
        whole. The base weights are left untouched, so adapters can be swapped without reloading the model."""
        cfg, hd, r = self.cfg, self.cfg.head_dim, lcfg.r
        dev, dt = self.final_norm.device, self.final_norm.dtype
        d = cfg.hidden_size
        Hl, KVl = self.H * hd, self.KV * hd
        q, kv = cfg.q_size, cfg.kv_size
        Rq = max(64, (3 * r + 63) // 64 * 64)
        Ro = max(64, (r + 63) // 64 * 64)
        sc = lcfg.scaling
        rk = self.rank

        def find(li, mod, which):
            for k, v in adapter_state.items():
                if f"layers.{li}.self_attn.{mod}.lora_{which}" in k:
                    return v.to(torch.float32)
            return None

        rows_full = {"q_proj": (rk * Hl, (rk + 1) * Hl), "k_proj": (rk * KVl, (rk + 1) * KVl),
                     "v_proj": (rk * KVl, (rk + 1) * KVl)}
        rows_local = {"q_proj": (0, Hl), "k_proj": (Hl, Hl + KVl), "v_proj": (Hl + KVl, Hl + 2 * KVl)}
        self.lora = []
        for li, l in enumerate(self.layers):
            ent = {"A_qkv": None, "B_qkv": None, "A_o": None}
            if any(m in lcfg.target_modules for m in ("q_proj", "k_proj", "v_proj")):
                A_c = torch.zeros(Rq, d, dtype=torch.float32)
                B_c = torch.zeros(Hl + 2 * KVl, Rq, dtype=torch.float32)
                for j, mod in enumerate(("q_proj", "k_proj", "v_proj")):
                    A, B = find(li, mod, "A"), find(li, mod, "B")
                    if A is None or B is None:
                        continue
                    A_c[j * r:(j + 1) * r] = A
                    f0, f1 = rows_full[mod]
                    l0, l1 = rows_local[mod]
                    B_c[l0:l1, j * r:(j + 1) * r] = B[f0:f1] * sc
                ent["A_qkv"], ent["B_qkv"] = A_c.to(dev, dt), B_c.to(dev, dt)
            A, B = find(li, "o_proj", "A"), find(li, "o_proj", "B")
            if "o_proj" in lcfg.target_modules and A is not None and B is not None:
                A_o = torch.zeros(Ro, Hl, dtype=torch.float32)
                A_o[:r] = A[:, rk * Hl:(rk + 1) * Hl]
                ent["A_o"] = A_o.to(dev, dt)
                w_ext = torch.zeros(d, Hl + Ro, dtype=dt, device=dev)
                w_ext[:, :Hl].copy_(l["w_o"])
                w_ext[:, Hl:Hl + r].copy_((B * sc).to(dev, dt))
                l["w_o_ext"] = w_ext
                ent["Ro"] = Ro
            self.lora.append(ent)
        return self

    def init_random(self, seed: int = 0) -> "TPWeights":
        """Random init identical to slicing ResidentWeights.random_init-style full matrices: every full
        matrix is generated from (seed, matrix id) on the device and sliced (so TP=1 and TP=N serve the
        same model)."""
        c, hd = self.cfg, self.cfg.head_dim
        dev = self.embed.device
        gen = torch.Generator(device=dev)

        def full(rows, cols, key):
            gen.manual_seed((seed * 1000003 + key) & 0x7FFFFFFFFFFF)
            t = torch.empty(rows, cols, dtype=self.embed.dtype, device=dev)
            t.normal_(0.0, 0.02, generator=gen)
            return t

        class _L:
            pass

        class _W:
            pass

        w = _W()
        w.embed = full(c.vocab_size, c.hidden_size, 1)
        w.lm_head = w.embed if c.tie_word_embeddings else full(c.vocab_size, c.hidden_size, 2)
        w.final_norm = torch.ones(c.hidden_size, dtype=self.embed.dtype, device=dev)
        self.embed.copy_(w.embed)
        self.final_norm.copy_(w.final_norm)
        self.lm_head.copy_(w.lm_head[self.rank * self.V:(self.rank + 1) * self.V])
        del w.embed, w.lm_head
        one = torch.ones(c.hidden_size, dtype=self.embed.dtype, device=dev)
        r = self.rank
        q, kv, ff = c.q_size, c.kv_size, c.intermediate_size
        Hl, KVl, ffl = self.H * hd, self.KV * hd, self.ff
        for li, dst in enumerate(self.layers):
            wq = full(c.qkv_size, c.hidden_size, li * 8 + 16)
            dst["w_qkv"][:Hl].copy_(wq[r * Hl:(r + 1) * Hl])
            dst["w_qkv"][Hl:Hl + KVl].copy_(wq[q + r * KVl:q + (r + 1) * KVl])
            dst["w_qkv"][Hl + KVl:].copy_(wq[q + kv + r * KVl:q + kv + (r + 1) * KVl])
            del wq
            dst["w_o"].copy_(full(c.hidden_size, c.q_size, li * 8 + 17)[:, r * Hl:(r + 1) * Hl])
            wg = full(2 * ff, c.hidden_size, li * 8 + 18)
            dst["w_gu"][:ffl].copy_(wg[r * ffl:(r + 1) * ffl])
            dst["w_gu"][ffl:].copy_(wg[ff + r * ffl:ff + (r + 1) * ffl])
            del wg
            dst["w_down"].copy_(full(c.hidden_size, ff, li * 8 + 19)[:, r * ffl:(r + 1) * ffl])
            dst["ln1"].copy_(one)
            dst["ln2"].copy_(one)
        return self


class ModelRunner:
    def __init__(self, cfg: LlamaConfig, weights: TPWeights, kv: KVCache, tp: TPContext):
        self.cfg, self.w, self.kv, self.tp = cfg, weights, kv, tp
        self.scale = 1.0 / math.sqrt(cfg.head_dim)
        self.cos, self.sin = E.rope_tables(cfg.max_position_embeddings, cfg.head_dim, cfg.rope_theta,
                                           weights.embed.device)
        self.device = weights.embed.device
        tp.adopt(weights, cfg)
        if self.device.type == "cuda" and "DLTI_PDL" not in os.environ:
            # serving weights are static: the decode GEMMs may fetch weight tiles while the previous kernel drains
            from ..ops import _lib
            _lib.gpu_lib().dlti_set_pdl(2)

    @torch.no_grad()
    def forward(self, input_ids, positions, slots, *, num_decode: int = 0, block_tables=None, ctx_lens=None,
                cu_seqlens=None, max_seqlen: int = 0, last_idx=None, max_ctx=None, prefix=None) -> torch.Tensor:
        """One pass over T token rows: the first `num_decode` rows are single new tokens of running
        sequences (attention over the paged cache via block_tables / ctx_lens), the remaining rows are whole
        prompts packed back to back (cu_seqlens, relative to the first prompt row; causal attention inside
        each prompt). Either part may be empty. Every row's rotated K and V are written to the cache at
        `slots`. Returns logits [rows selected by last_idx (all rows if None), vocab].

        prefix = (pslots, pdst, csrc, ecu, max_ext): chunked prefill. Some prompt rows continue prompts whose
        earlier tokens are already in the paged cache. Their attention runs in an extended layout: per prompt,
        [cached prefix | this step's rows] is laid out as one causal sequence (ecu = boundaries), the prefix
        K/V rows are gathered from the cache (pslots -> rows pdst; their q is zero and their outputs are
        dropped), this step's rows sit at csrc, and the same varlen kernel is used. The extra score work on
        the prefix rows is attention only (a few percent of a prefill step) and needs no second kernel."""
        cfg, w, tp = self.cfg, self.w, self.tp
        H, KV, hd, eps = w.H, w.KV, cfg.head_dim, cfg.rms_norm_eps
        T = input_ids.numel()
        nd = int(num_decode)
        residual = E.embedding(input_ids.view(-1), w.embed)                  # [T, d] residual stream
        xn, _, _ = E.rmsnorm_fwd(residual, w.layers[0]["ln1"], eps)
        L = cfg.num_hidden_layers
        # decode-sized steps stream the weights through the swap-AB stream-K kernel; its fp32 accumulator is
        # consumed directly by the next elementwise op (no stand-alone convert pass)
        skinny = T <= 256
        acc_path = skinny and _FUSE_ACC and residual.is_cuda and hd == 128 and residual.dtype != torch.float32
        mm = gemm_skinny if skinny else (lambda a, b_: gemm(a, b_))

        def rowpar(act, wgt):
            return lambda out, push, scatter=None: (gemm(act, wgt, out=out, push=push, scatter=scatter)
                                                    if (push is not None or out is not None) else mm(act, wgt))

        def colpar(x, wgt):                        # column-parallel GEMM; gathers its rows itself when they are sharded
            if isinstance(x, ShardedAct):
                return gemm(x.rows(), wgt, gather=x.handle)
            return mm(x, wgt)

        def row_reduce(act, wgt, norm_w):          # row-parallel GEMM -> (all-reduce) -> residual add -> norm
            if acc_path and tp.world == 1:
                return E.acc_add_rmsnorm(gemm_skinny_acc(act, wgt), residual, norm_w, eps)
            return tp.reduce_add_norm(rowpar(act, wgt), T, residual, norm_w, eps)

        for li, lw in enumerate(w.layers):
            ad = w.lora[li] if w.lora is not None else None          # unmerged adapter operands of this layer
            if ad is not None and ad["A_qkv"] is not None:
                # LoRA branch of q/k/v: t = x A^T (rank-sized), then (alpha/r) t B^T joins the base projection in the
                # same accumulator (second stream-K pass into the fp32 workspace / K-extension of the GEMM)
                if isinstance(xn, ShardedAct):
                    t = gemm(xn.rows(), ad["A_qkv"], gather=xn.handle)
                    x_in = xn.rows()
                else:
                    t = (gemm_skinny if (skinny and xn.is_cuda) else gemm)(xn, ad["A_qkv"])
                    x_in = xn
                if acc_path:
                    acc = gemm_skinny_acc(x_in, lw["w_qkv"])
                    gemm_skinny_acc(t, ad["B_qkv"])
                    qkv = S.acc_rope_cache(acc, x_in.dtype, positions, self.cos, self.sin, slots, self.kv.k[li],
                                           self.kv.v[li], H, KV, hd)
                else:
                    qkv = gemm(x_in, lw["w_qkv"], a2=t, b2=ad["B_qkv"])
                    S.rope_cache_(qkv, positions, self.cos, self.sin, slots, self.kv.k[li], self.kv.v[li], H, KV, hd)
            elif acc_path:
                qkv = S.acc_rope_cache(gemm_skinny_acc(xn, lw["w_qkv"]), xn.dtype, positions, self.cos, self.sin,
                                       slots, self.kv.k[li], self.kv.v[li], H, KV, hd)
            else:
                qkv = colpar(xn, lw["w_qkv"])
                S.rope_cache_(qkv, positions, self.cos, self.sin, slots, self.kv.k[li], self.kv.v[li], H, KV, hd)
            if prefix is not None:
                attn = torch.empty((T, H * hd), dtype=qkv.dtype, device=qkv.device)
                if nd:
                    S.paged_decode(qkv[:nd], self.kv.k[li], self.kv.v[li], block_tables, ctx_lens, H, KV, hd,
                                   self.scale, out=attn[:nd], max_ctx=max_ctx)
                pslots, pdst, csrc, ecu, max_ext = prefix
                ext = torch.zeros((int(pdst.numel() + csrc.numel()), qkv.shape[1]), dtype=qkv.dtype, device=qkv.device)
                ext[csrc] = qkv[nd:]
                ext[pdst, H * hd:(H + KV) * hd] = self.kv.k[li].view(-1, KV * hd)[pslots]
                ext[pdst, (H + KV) * hd:] = self.kv.v[li].view(-1, KV * hd)[pslots]
                out_ext, _ = A.attention_fwd(ext, ecu, max_ext, H, KV, hd, self.scale, want_lse=False)
                attn[nd:] = out_ext[csrc]
            elif nd == 0:
                attn, _ = A.attention_fwd(qkv, cu_seqlens, max_seqlen, H, KV, hd, self.scale, want_lse=False)
            elif nd == T:
                attn = S.paged_decode(qkv, self.kv.k[li], self.kv.v[li], block_tables, ctx_lens, H, KV, hd,
                                      self.scale, max_ctx=max_ctx)
            else:
                attn = torch.empty((T, H * hd), dtype=qkv.dtype, device=qkv.device)
                S.paged_decode(qkv[:nd], self.kv.k[li], self.kv.v[li], block_tables, ctx_lens, H, KV, hd,
                               self.scale, out=attn[:nd], max_ctx=max_ctx)
                A.attention_fwd(qkv[nd:], cu_seqlens, max_seqlen, H, KV, hd, self.scale, want_lse=False,
                                out=attn[nd:])
            if ad is not None and ad["A_o"] is not None:
                # LoRA branch of o_proj as extra K columns: [attention | t_o] x [W_o | (alpha/r) B_o]^T
                Hl = H * hd
                act = torch.empty((T, Hl + ad["Ro"]), dtype=attn.dtype, device=attn.device)
                act[:, :Hl].copy_(attn)
                (gemm_skinny if (skinny and attn.is_cuda) else gemm)(attn, ad["A_o"], out=act[:, Hl:])
                xn = row_reduce(act, lw["w_o_ext"], lw["ln2"])
            else:
                xn = row_reduce(attn, lw["w_o"], lw["ln2"])
            if acc_path:
                h = E.acc_swiglu(gemm_skinny_acc(xn, lw["w_gu"]), xn.dtype)
            elif _PREFILL_FUSE_SWIGLU and not skinny:
                # SwiGLU in the CTA-pair GEMM epilogue (falls back by itself)
                h = (gemm_swiglu(xn.rows(), lw["w_gu"], gather=xn.handle) if isinstance(xn, ShardedAct)
                     else gemm_swiglu(xn, lw["w_gu"]))
            else:
                h = E.swiglu_fwd(colpar(xn, lw["w_gu"]))
            nxt = w.layers[li + 1]["ln1"] if li + 1 < L else w.final_norm
            xn = row_reduce(h, lw["w_down"], nxt)
        if last_idx is not None and last_idx.numel() == 0:      # a step of partial prompt chunks samples nothing
            return torch.empty((0, cfg.vocab_size), dtype=xn.dtype, device=xn.device)
        if isinstance(xn, ShardedAct):                          # the LM head needs the sampled rows on every rank
            xn = xn.gathered()
        xl = xn if last_idx is None else xn.index_select(0, last_idx)
        return tp.lm_head_logits(xl, w.lm_head)
