"""Plain-PyTorch autograd restatement of Llama + LoRA (the oracle for the hand-written
forward/backward in dlti_b200.models.llama). Mirrors HF LlamaForCausalLM + PEFT lora.Linear."""
import math

import torch
import torch.nn.functional as F

from dlti_b200.ops.rng import keep_mask


def ref_loss(cfg, w, lora, batch, p_drop, seed, micro_step):
    """Returns (sum of per-token CE, dict of leaf tensors name -> requires_grad fp32 copies)."""
    leaves = {}
    r = lora.lcfg.r
    s = lora.lcfg.scaling
    for n in lora.slots:
        leaves[n] = lora.view(n).detach().clone().double().requires_grad_(True)
    H, KV, hd = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
    T = batch.input_ids.numel()
    x = w.embed.double()[batch.input_ids]
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, hd, 2, dtype=torch.float64) / hd))
    ang = batch.positions.double()[:, None] * inv[None, :]
    cos, sin = ang.cos()[:, None, :], ang.sin()[:, None, :]

    def rms(x, wt):
        return wt.double() * (x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + cfg.rms_norm_eps))

    def rope(v):
        a, b = v[..., : hd // 2], v[..., hd // 2:]
        return torch.cat([a * cos - b * sin, b * cos + a * sin], -1)

    def drop(v, li, site):
        if p_drop <= 0:
            return v
        m = keep_mask(v.shape[0], v.shape[1], p_drop, seed, (micro_step << 20) | (li << 4) | site)
        return v * m.double() / (1 - p_drop)

    def lin(xin, xdrop, W, li, mod):
        y = xin @ W.double().t()
        na = f"layers.{li}.self_attn.{mod}.lora_A"
        if na in leaves:
            y = y + s * (xdrop @ leaves[na].t()) @ leaves[na.replace("lora_A", "lora_B")].t()
        return y

    q_, kv_ = cfg.q_size, cfg.kv_size
    cu = batch.cu_seqlens.tolist()
    for li, l in enumerate(w.layers):
        xn = rms(x, l.ln1)
        xd = drop(xn, li, 0)
        q = lin(xn, xd, l.w_qkv[:q_], li, "q_proj").view(T, H, hd)
        k = lin(xn, xd, l.w_qkv[q_:q_ + kv_], li, "k_proj").view(T, KV, hd)
        v = lin(xn, xd, l.w_qkv[q_ + kv_:], li, "v_proj").view(T, KV, hd)
        q, k = rope(q), rope(k)
        outs = []
        for b in range(len(cu) - 1):
            s0, e0 = cu[b], cu[b + 1]
            qq = q[s0:e0].transpose(0, 1)
            kk = k[s0:e0].transpose(0, 1).repeat_interleave(H // KV, 0)
            vv = v[s0:e0].transpose(0, 1).repeat_interleave(H // KV, 0)
            o = F.scaled_dot_product_attention(qq, kk, vv, is_causal=True)
            outs.append(o.transpose(0, 1).reshape(e0 - s0, H * hd))
        attn = torch.cat(outs)
        x = x + lin(attn, drop(attn, li, 1), l.w_o, li, "o_proj")
        xn2 = rms(x, l.ln2)
        ff = cfg.intermediate_size
        g, u = xn2 @ l.w_gu[:ff].double().t(), xn2 @ l.w_gu[ff:].double().t()
        x = x + (F.silu(g) * u) @ l.w_down.double().t()
    logits = rms(x, w.final_norm) @ w.lm_head.double().t()
    loss = F.cross_entropy(logits, batch.labels, ignore_index=-100, reduction="sum")
    return loss, leaves, logits
