#!/usr/bin/env python
"""Developer probe: time one Llama-2-7B LoRA micro-step (forward+backward, layer recompute) on one GPU."""
import os, sys, time, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlti_b200.models.config import resolve_config
from dlti_b200.models.llama import LoraLlama, ResidentWeights, pack_batch
from dlti_b200.models.lora import LoraConfig, LoraParams
from dlti_b200.ops import _lib

name = os.environ.get("MODEL", "llama2-7b"); B = int(os.environ.get("B", 2)); S = int(os.environ.get("S", 2048))
ckpt = os.environ.get("CKPT", "1") == "1"
cfg = resolve_config(name)
t0 = time.time()
w = ResidentWeights.random_init(cfg, "cuda", torch.bfloat16, seed=0)
lora = LoraParams(cfg, LoraConfig(r=16, lora_alpha=32, lora_dropout=0.05), "cuda", torch.bfloat16)
model = LoraLlama(cfg, w, lora, gradient_checkpointing=ckpt)
torch.cuda.synchronize(); print("init s", time.time() - t0, "weights GB", w.nbytes() / 1e9, flush=True)
ids = torch.randint(3, cfg.vocab_size, (B, S)); b = pack_batch(ids, ids.clone(), None).to("cuda")
gs = torch.tensor([1.0 / b.n_valid], device="cuda")
for i in range(int(os.environ.get('WARM', 3))):
    loss = model.forward_backward(b, gs)
torch.cuda.synchronize()
_lib.reset_launch_count()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = int(os.environ.get('N', 5))
s.record()
t_cpu0 = time.perf_counter()
for i in range(n):
    loss = model.forward_backward(b, gs)
cpu_enqueue_ms = (time.perf_counter() - t_cpu0) * 1e3 / n      # host time to enqueue one micro-step
e.record(); torch.cuda.synchronize()
ms = s.elapsed_time(e) / n
toks = B * S
flops = 6 * cfg.num_params() * toks if ckpt else 4 * cfg.num_params() * toks
print(json.dumps({"model": name, "B": B, "S": S, "ckpt": ckpt, "ms_per_microstep": ms, "tok_per_s": toks / ms * 1e3,
                  "model_tflops": flops / ms / 1e9, "loss": loss.item() / b.n_valid, "launches_per_step": _lib.launch_count() / n,
                  "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9, "cpu_enqueue_ms": cpu_enqueue_ms}))
