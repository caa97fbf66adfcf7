#!/usr/bin/env python
"""Headline benchmark: Llama-2-7B LoRA (r=16, q/k/v/o) ZeRO-3 bf16 training throughput, seq 2048,
synthetic data, random-init weights (BASELINE.json metric "Llama-2-7B LoRA ZeRO-3 train tok/s").

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W

One optimizer step = gradient_accumulation_steps (4) micro-batches of per_device_batch (2) x 2048
tokens per GPU (the reference's ZeRO-3 defaults: training/train_deepspeed_zero3.py:69-81) with
gradient checkpointing and LoRA dropout 0.05 on, then the ZeRO gradient reduce, global-norm clip,
AdamW and parameter refresh. Weak scaling: per-GPU work is fixed, `value` is total tokens/s.

Two timed regions of K steps each (W warm-up steps first; CUDA events; barrier + synchronize on
both sides; max over ranks):
  value : inputs already resident on the device (kernel-side number)
  e2e   : through the public API `Trainer.train_step(host_batches) -> float`: every step copies its
          micro-batches from pinned host memory and reads the loss back to the host.
Working set per step (13.5 GB of weights + activations) exceeds the 126 MB L2 by two orders of
magnitude, so no explicit L2 flush is needed between iterations ("inputs larger than L2").

`--impl reference` must run the unmodified reference; it is not installable here (no setup.py /
pyproject; it needs deepspeed + peft + accelerate, absent from the image and from /opt/wheelhouse),
so that arm prints {"impl": "reference", "unavailable": ...} and exits 0 (DESIGN.md §Reference arm).
`--impl standin` runs the stock-PyTorch (cuBLAS + SDPA + NCCL) restatement under baseline/.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", type=str, default="dlti", choices=["dlti", "reference", "standin"])
    ap.add_argument("--model", type=str, default="llama2-7b")
    ap.add_argument("--seq_len", type=int, default=2048)
    ap.add_argument("--micro_batch", type=int, default=2)
    ap.add_argument("--grad_accum", type=int, default=4)
    ap.add_argument("--zero_stage", type=int, default=3)
    ap.add_argument("--backend", type=str, default="fused", choices=["fused", "nccl"])
    ap.add_argument("--no_grad_ckpt", action="store_true")
    ap.add_argument("--lora_r", type=int, default=16)
    ap.add_argument("--offload", action="store_true",
                    help="run the ds_config as shipped (ZeRO-Offload of optimizer state and parameters to pinned host "
                         "memory) instead of the HBM-resident default; a secondary measurement, not the headline (689 ms vs 631 ms per step on one B200)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.p = None
        self.path = f"/tmp/dlti_clocks_{os.getpid()}.csv"
        try:
            self.f = open(self.path, "w")
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200", "-i", str(gpu_index)], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.close()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
            except ValueError:
                continue
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


def reference_arm(a) -> None:
    why = ("reference is not pip-installable (no setup.py/pyproject.toml: pip says 'not installable') and its "
           "scripts import deepspeed, peft and accelerate, none of which is in the image or /opt/wheelhouse")
    ref = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(ref) and os.listdir(ref):
        why = "baseline/_ref exists but deepspeed/peft/accelerate are not importable on this image"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))


def main() -> None:
    a = parse()
    if a.impl == "reference":
        reference_arm(a)
        return
    if a.impl == "standin":
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import standin_train
        standin_train.bench(a)
        return

    import torch
    import torch.distributed as dist

    from dlti_b200.config.args import TrainingArgs
    from dlti_b200.engine.trainer import Trainer
    from dlti_b200.models.config import resolve_config
    from dlti_b200.models.llama import pack_batch
    from dlti_b200.models.lora import LoraConfig
    from dlti_b200.ops import _lib
    from dlti_b200.parallel import comm

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200)")
    info = comm.init_distributed("cuda")
    world = info.world_size
    assert world == a.gpus, f"--gpus {a.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run)"
    cfg = resolve_config(a.model)
    ds_path = os.path.join(ROOT, "configs", f"ds_config_zero{a.zero_stage}.json") if a.zero_stage else None
    targs = TrainingArgs(output_dir="/tmp/dlti_bench", per_device_train_batch_size=a.micro_batch,
                         gradient_accumulation_steps=a.grad_accum, learning_rate=2e-4, precision="bf16",
                         gradient_checkpointing=not a.no_grad_ckpt, deepspeed=ds_path, no_offload=not a.offload,
                         device="cuda", backend=a.backend, max_steps=a.steps + a.warmup + 8)
    lcfg = LoraConfig(r=a.lora_r, lora_alpha=2 * a.lora_r, lora_dropout=0.05)
    tr = Trainer(cfg, lcfg, targs, None, None, info=info)

    # synthetic fixed-length batches (seq_len tokens each), distinct per step and per rank
    g = torch.Generator().manual_seed(1234 + info.rank)
    n_steps_total = a.warmup + a.steps
    def make_step():
        mbs = []
        for _ in range(a.grad_accum):
            ids = torch.randint(3, cfg.vocab_size, (a.micro_batch, a.seq_len), generator=g)
            ids[:, 0] = cfg.bos_token_id
            mbs.append(pack_batch(ids, ids.clone(), None, pin=True))
        return mbs
    host_steps = [make_step() for _ in range(n_steps_total)]
    tokens_per_step = a.micro_batch * a.seq_len * a.grad_accum * world
    h2d = sum(t.numel() * t.element_size() for mb in host_steps[0]
              for t in (mb.input_ids, mb.labels, mb.positions, mb.cu_seqlens))

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def min_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return float(t.item())

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- warm-up ----
    for s in range(a.warmup):
        tr.train_step(host_steps[s])
    # ---- region 1: device-resident inputs ----
    dev_steps = [[mb.to("cuda") for mb in st] for st in host_steps[a.warmup:]]
    sync_all()
    sampler = ClockSampler(torch.cuda.current_device()) if info.rank == 0 else None
    tr.ds.cfg["wall_clock_breakdown"] = True        # one extra event per step: forward+backward | reduce/clip/AdamW/refresh
    tr._phase_events = []
    _lib.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for st in dev_steps:
        loss_t, n_tot = tr._step_device(st)
    e1.record()
    sync_all()
    dev_ms = max_over_ranks(e0.elapsed_time(e1)) / a.steps
    launches = _lib.launch_count()
    # per-rank phase times: a rank that finishes its forward+backward early waits for the slowest one inside the
    # optimizer phase (first peer barrier), so min/max of the first phase over ranks shows the spread between GPUs
    fb = sum(x[0].elapsed_time(x[1]) for x in tr._phase_events) / max(1, len(tr._phase_events))
    tail = sum(x[1].elapsed_time(x[2]) for x in tr._phase_events) / max(1, len(tr._phase_events))
    phases = {"fwd_bwd_ms_min_rank": min_over_ranks(fb), "fwd_bwd_ms_max_rank": max_over_ranks(fb),
              "optimizer_phase_ms_min_rank": min_over_ranks(tail), "optimizer_phase_ms_max_rank": max_over_ranks(tail)}
    tr.ds.cfg["wall_clock_breakdown"] = False
    tr._phase_events = []
    loss_dev = float(loss_t.item()) / n_tot
    # ---- region 2: end to end through the public API ----
    sync_all()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e2.record()
    for st in host_steps[a.warmup:]:
        loss = tr.train_step(st)             # H2D from pinned memory ... D2H of the loss
    e3.record()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms = max_over_ranks(max(e2.elapsed_time(e3), wall_ms)) / a.steps
    clocks = sampler.stop() if sampler else None
    sync_all()
    arena_gb = tr.fused.arena.total_bytes() / 1e9 if tr.fused is not None else 0.0   # symmetric arena is
    peak = max_over_ranks(torch.cuda.max_memory_allocated() / 1e9 + arena_gb)         # outside torch's allocator

    # a step that produced a non-finite loss did not train anything: refuse to report a number for it
    import math
    if not (math.isfinite(loss_dev) and math.isfinite(float(loss))):
        raise SystemExit(f"bench: non-finite loss (device region {loss_dev}, e2e {loss}); no result reported")
    if info.rank == 0:
        value = tokens_per_step / dev_ms * 1e3
        e2e = tokens_per_step / e2e_ms * 1e3
        par = f"zero{a.zero_stage}-dp{world}" if a.zero_stage else f"dp{world}"
        out = {
            "metric": "Llama-2-7B LoRA ZeRO-3 train tok/s" if a.model == "llama2-7b" and a.zero_stage == 3
                      else f"{a.model} LoRA ZeRO-{a.zero_stage} train tok/s",
            "value": value, "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic (fixed-length random token ids; random-init weights)",
            "impl": "dlti_b200",
            "config": {"model": a.model, "global_batch": a.micro_batch * a.grad_accum * world,
                       "seq_len": a.seq_len, "parallelism": par, "micro_batch_per_gpu": a.micro_batch,
                       "grad_accum": a.grad_accum, "lora_r": a.lora_r, "lora_dropout": 0.05,
                       "lora_targets": "q,k,v,o", "gradient_checkpointing": not a.no_grad_ckpt,
                       "optimizer": "AdamW fp32 master, clip 1.0",
                       "offload": "optimizer + parameters in pinned host memory (ds_config as shipped)" if a.offload
                                  else "none (HBM-resident)",
                       "comm_backend": a.backend if world > 1 else "n/a (1 GPU)",
                       "l2": "inputs larger than L2 (13.5 GB weights streamed per pass)"},
            "e2e": {"value": e2e, "unit": "tokens/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 12},
            "gpu_launches": int(launches),
            "clocks": clocks, "phases": phases,
            "loss": loss, "loss_device_region": loss_dev, "peak_memory_gb_max_rank": peak,
            "model_tflops_per_gpu": (6 if not a.no_grad_ckpt else 4) * cfg.num_params() * tokens_per_step
                                    / world / dev_ms / 1e9,
        }
        # FLOPs the step actually executes (frozen base: no weight gradients; the recompute skips the down
        # projection and the LM head) and the recompute-free "model FLOPs" (4P: forward + input gradients)
        d_, ff_, L_ = cfg.hidden_size, cfg.intermediate_size, cfg.num_hidden_layers
        p_layer = d_ * cfg.qkv_size + cfg.q_size * d_ + 3 * ff_ * d_
        p_head = cfg.vocab_size * d_
        rec = 0 if a.no_grad_ckpt else 2 * L_ * (p_layer - ff_ * d_)
        lin = 2 * (L_ * p_layer + p_head) * 2 + rec
        attn_fwd = 2 * a.seq_len * cfg.q_size * L_                    # causal: half of 4*S*q_size per token
        attn = attn_fwd * (2.5 + (1 if a.no_grad_ckpt else 2))
        tok_gpu_s = tokens_per_step / world / dev_ms * 1e3
        out["executed_tflops_per_gpu"] = (lin + attn) * tok_gpu_s / 1e12
        out["mfu_4p_tflops_per_gpu"] = (4 * cfg.num_params() + 3.5 * attn_fwd) * tok_gpu_s / 1e12
        try:
            pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            out["executed_frac_of_measured_sustained_bf16"] = out["executed_tflops_per_gpu"] / pk["bf16_tflops_sustained"]
            out["mfu_4p_frac_of_measured_sustained_bf16"] = out["mfu_4p_tflops_per_gpu"] / pk["bf16_tflops_sustained"]
        except Exception:
            pass
        # same-box comparator measured with `bench.py --impl standin` (profiles/standin_bench_n*_r2.json)
        try:
            sb = [json.loads(l) for l in open(os.path.join(ROOT, "profiles", f"standin_bench_n{world}_r2.json"))
                  if l.startswith("{")][0]
            out["vs_standin_committed"] = {"standin_tok_s": sb["value"], "ratio": value / sb["value"],
                                           "file": f"profiles/standin_bench_n{world}_r2.json"}
        except Exception:
            pass
        print(json.dumps(out))
    comm.barrier(info)
    comm.shutdown()


if __name__ == "__main__":
    main()
