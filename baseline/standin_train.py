"""STAND-IN baseline (not the reference, not the product): the same LoRA fine-tuning step written with
stock libraries only — transformers' LlamaForCausalLM (cuBLAS GEMMs, SDPA attention), a plain-PyTorch
LoRA wrapper (PEFT is not on the image), torch.utils.checkpoint, torch fused AdamW, and DDP/NCCL over
the trainable adapter parameters for N > 1 (the frozen base is replicated, which is *cheaper* than
ZeRO-3's parameter gathers). The reference's own stack (DeepSpeed + PEFT + HF Trainer) cannot be
installed here, so this is the "NCCL + cuBLAS" comparator BASELINE.md asks for.

Run through `python bench.py --impl standin --gpus N ...`; prints the same JSON line with
"impl": "standin".
"""
from __future__ import annotations

import json
import math
import os
import time

import torch
import torch.nn as nn
import torch.nn.functional as F


class LoRALinear(nn.Module):
    def __init__(self, base: nn.Linear, r: int, alpha: int, dropout: float):
        super().__init__()
        self.base = base
        self.scaling = alpha / r
        self.drop = nn.Dropout(dropout)
        self.lora_A = nn.Parameter(torch.empty(r, base.in_features, dtype=torch.float32, device=base.weight.device))
        self.lora_B = nn.Parameter(torch.zeros(base.out_features, r, dtype=torch.float32, device=base.weight.device))
        nn.init.kaiming_uniform_(self.lora_A, a=math.sqrt(5))

    def forward(self, x):
        y = self.base(x)
        z = F.linear(F.linear(self.drop(x).float(), self.lora_A), self.lora_B) * self.scaling
        return y + z.to(y.dtype)


def build_model(cfg, r, device):
    import transformers

    hf_cfg = transformers.LlamaConfig(**{k: v for k, v in cfg.to_hf_dict().items()
                                         if k not in ("architectures", "model_type", "torch_dtype")})
    hf_cfg._attn_implementation = "sdpa"
    with torch.device(device):
        model = transformers.LlamaForCausalLM(hf_cfg).to(torch.bfloat16)
    for p in model.parameters():
        p.requires_grad_(False)
    for layer in model.model.layers:
        for name in ("q_proj", "k_proj", "v_proj", "o_proj"):
            setattr(layer.self_attn, name, LoRALinear(getattr(layer.self_attn, name), r, 2 * r, 0.05))
    model.gradient_checkpointing_enable(gradient_checkpointing_kwargs={"use_reentrant": False})
    model.config.use_cache = False
    model.enable_input_require_grads()
    return model


class Zero3Base:
    """ZeRO-3 stand-in for the FROZEN base of the decoder layers, with stock PyTorch collectives: every layer's
    frozen parameters are flattened, only 1/world of the flat tensor is kept per rank, and a forward pre-hook
    all-gathers the layer into one of two full-size buffers (NCCL all_gather_into_tensor on a side stream, issued
    one layer ahead in the direction of travel — forward, and again in reverse for the checkpoint recompute) and
    points the layer's parameters at it. This is what /root/reference/configs/ds_config_zero3.json makes DeepSpeed
    do (stage 3 parameter partitioning with prefetch and overlap_comm), minus DeepSpeed's own bookkeeping. The
    embedding and the LM head stay replicated (which only favours the stand-in)."""

    def __init__(self, model, info):
        import torch.distributed as dist
        self.dist, self.info = dist, info
        self.world = info.world_size
        self.layers = list(model.model.layers)
        self.metas, self.shards = [], []
        maxn = 0
        for layer in self.layers:
            ps = [p for p in layer.parameters() if not p.requires_grad]
            n = sum(p.numel() for p in ps)
            npad = -(-n // self.world) * self.world
            maxn = max(maxn, npad)
            flat = torch.zeros(npad, dtype=ps[0].dtype, device=ps[0].device)
            torch.cat([p.data.reshape(-1) for p in ps], out=flat[:n])
            per = npad // self.world
            self.shards.append(flat[info.rank * per:(info.rank + 1) * per].clone())
            self.metas.append((ps, [p.shape for p in ps], npad))
            for p in ps:
                p.data = torch.empty(0, dtype=p.dtype, device=p.device)
            del flat
        torch.cuda.empty_cache()
        self.bufs = [torch.empty(maxn, dtype=self.shards[0].dtype, device=info.device) for _ in range(2)]
        self.stream = torch.cuda.Stream(device=info.device)
        self.ready = {}            # layer index -> event
        self.last = -1
        self.gathered_bytes = 0
        for i, layer in enumerate(self.layers):
            layer.register_forward_pre_hook(lambda mod, args, kwargs=None, i=i: self._pre(i), with_kwargs=False)

    def _launch(self, i: int) -> None:
        if i < 0 or i >= len(self.layers) or i in self.ready:
            return
        npad = self.metas[i][2]
        self.stream.wait_stream(torch.cuda.current_stream())      # the buffer's previous user is done
        with torch.cuda.stream(self.stream):
            self.dist.all_gather_into_tensor(self.bufs[i % 2][:npad], self.shards[i])
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.ready[i] = ev
        self.gathered_bytes += npad * 2 * (self.world - 1) // self.world

    def _pre(self, i: int) -> None:
        self._launch(i)
        torch.cuda.current_stream().wait_event(self.ready.pop(i))
        ps, shapes, _ = self.metas[i]
        buf, off = self.bufs[i % 2], 0
        for p, sh in zip(ps, shapes):
            n = math.prod(sh)
            p.data = buf[off:off + n].view(sh)
            off += n
        nxt = i + 1 if i >= self.last else i - 1                   # direction of travel
        self.last = i
        self.ready = {k: v for k, v in self.ready.items() if k == nxt}
        self._launch(nxt)


def bench(a) -> None:
    import torch.distributed as dist

    from dlti_b200.models.config import resolve_config
    from dlti_b200.parallel import comm

    info = comm.init_distributed("cuda")
    world = info.world_size
    cfg = resolve_config(a.model)
    torch.manual_seed(0)
    model = build_model(cfg, a.lora_r, info.device)
    model.train()
    params = [p for p in model.parameters() if p.requires_grad]
    net = model
    mode = os.environ.get("STANDIN_MODE", "ddp")     # "ddp": frozen base replicated; "zero3": sharded + gathered
    z3 = Zero3Base(model, info) if (mode == "zero3" and world > 1) else None
    if world > 1:
        net = nn.parallel.DistributedDataParallel(model, device_ids=[info.device.index], find_unused_parameters=False)
    opt = torch.optim.AdamW(params, lr=2e-4, weight_decay=0.0, fused=True)
    g = torch.Generator().manual_seed(1234 + info.rank)

    def make_step():
        return [torch.randint(3, cfg.vocab_size, (a.micro_batch, a.seq_len), generator=g).pin_memory()
                for _ in range(a.grad_accum)]

    steps = [make_step() for _ in range(a.warmup + a.steps)]

    def run_step(mbs):
        total = 0.0
        for k, ids in enumerate(mbs):
            ids = ids.to(info.device, non_blocking=True)
            ctx = net.no_sync() if (world > 1 and k < len(mbs) - 1) else torch.autocast("cuda", enabled=False)
            with ctx:
                out = net(input_ids=ids, labels=ids)
                (out.loss / len(mbs)).backward()
            total = total + out.loss.detach() / len(mbs)
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
        return total

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for s in range(a.warmup):
        run_step(steps[s])
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for s in range(a.warmup, a.warmup + a.steps):
        loss = float(run_step(steps[s]).item())        # e2e: H2D of the batch, D2H of the loss
    e1.record()
    torch.cuda.synchronize()
    ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / a.steps
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    tokens = a.micro_batch * a.seq_len * a.grad_accum * world
    if info.rank == 0:
        print(json.dumps({
            "metric": "Llama-2-7B LoRA train tok/s (stand-in: HF transformers + SDPA + cuBLAS + torch AdamW + "
                      + ("NCCL ZeRO-3 gathers of the frozen base + DDP on the adapters)" if z3 else "DDP)"),
            "impl": "standin", "value": tokens / ms * 1e3, "unit": "tokens/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "loss": loss,
            "config": {"model": a.model, "global_batch": a.micro_batch * a.grad_accum * world, "seq_len": a.seq_len,
                       "parallelism": (f"zero3-dp{world} (frozen decoder layers sharded, all_gather_into_tensor per layer, "
                                       "one layer of look-ahead)" if z3 else f"ddp{world} (frozen base replicated)"),
                       "gradient_checkpointing": True,
                       "lora_r": a.lora_r},
            "peak_memory_gb": torch.cuda.max_memory_allocated() / 1e9, "gpu_launches": 0}))
    comm.barrier(info)
    comm.shutdown()
