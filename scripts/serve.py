#!/usr/bin/env python
"""Launch the OpenAI-compatible server, single GPU or tensor parallel.

    python scripts/serve.py --model_config llama2-7b --synthetic --port 8000
    python -m dlti_b200.launch --num_gpus=8 --no_local_rank scripts/serve.py --model <hf_dir> [--adapter <peft_dir>]

Rank 0 runs the HTTP server + scheduler; the other ranks run the TP worker loop. `--adapter` serves a PEFT
LoRA adapter unmerged on top of the base weights (LoRA branch inside the QKV / o_proj GEMMs); with
`--merge_adapter` it is folded into the base weights at load time instead (W' = W + alpha/r * B A). The reference declares
"vLLM serving with tensor parallelism" (README.md:10) without shipping a launch script."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build_engine(a):
    import torch

    from dlti_b200.models.config import resolve_config
    from dlti_b200.parallel import comm
    from dlti_b200.serve.engine import LLMEngine
    from dlti_b200.serve.model_runner import TPContext, TPWeights
    from dlti_b200.serve.tokenizer import load_tokenizer

    info = comm.init_distributed(a.device)
    dtype = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[a.dtype]
    model_dir = a.model if a.model and os.path.isdir(a.model) else None
    cfg = resolve_config(a.model_config or a.model)
    fused = None
    if info.distributed and a.backend == "fused" and info.device.type == "cuda":
        from dlti_b200.parallel.fused import FusedComm
        fused = FusedComm(info)
    tp = TPContext(info, a.backend, fused, cfg, max_tokens=a.max_num_batched_tokens, dtype=dtype)
    w = TPWeights(cfg, info.rank, info.world_size, info.device, dtype)
    if model_dir and not a.synthetic:
        from dlti_b200.models.hf_io import load_resident
        _, full = load_resident(model_dir, "cpu", dtype)
        if a.adapter and a.merge_adapter:
            from dlti_b200.models.merge import merge_into_weights
            from dlti_b200.models.peft_io import load_adapter_config, load_adapter_state
            merge_into_weights(full, load_adapter_state(a.adapter), load_adapter_config(a.adapter))
        w.load_from_full(full)
        del full
        if a.adapter and not a.merge_adapter:      # base + adapter served side by side (the adapter stays swappable)
            from dlti_b200.models.peft_io import load_adapter_config, load_adapter_state
            w.attach_adapter(load_adapter_state(a.adapter), load_adapter_config(a.adapter))
    else:
        w.init_random(a.seed)
    tok = load_tokenizer(model_dir, cfg.vocab_size)
    eng = LLMEngine(cfg, w, tp, tok, info=info, block_size=a.block_size, num_blocks=a.num_blocks,
                    gpu_memory_utilization=a.gpu_memory_utilization, max_num_seqs=a.max_num_seqs,
                    max_num_batched_tokens=a.max_num_batched_tokens, max_model_len=a.max_model_len,
                    use_cuda_graphs=not a.enforce_eager, mixed_batches=not a.no_mixed_batches,
                    async_scheduling=not a.no_async_scheduling, prefill_chunk=a.prefill_chunk,
                    enable_prefix_caching=a.enable_prefix_caching)
    if info.rank == 0 and not a.no_warmup:
        eng.warmup()                      # workers follow from worker_loop()
    return info, cfg, eng


def add_args(ap):
    ap.add_argument("--model", type=str, default=None, help="HF checkpoint directory (or preset name)")
    ap.add_argument("--model_config", type=str, default=None, help="tiny | llama2-7b | llama2-70b")
    ap.add_argument("--adapter", type=str, default=None, help="PEFT LoRA adapter directory served on top of the base")
    ap.add_argument("--merge_adapter", action="store_true",
                    help="fold the adapter into the base weights at load time (W' = W + alpha/r B A) instead of "
                         "serving it unmerged")
    ap.add_argument("--synthetic", action="store_true", help="random-init weights (no checkpoint needed)")
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp16", "fp32"])
    ap.add_argument("--device", default="cuda")
    ap.add_argument("--backend", default="fused", choices=["fused", "nccl"])
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8000)
    ap.add_argument("--block_size", type=int, default=16)
    ap.add_argument("--num_blocks", type=int, default=None)
    ap.add_argument("--gpu_memory_utilization", type=float, default=0.85)
    ap.add_argument("--max_num_seqs", type=int, default=256)
    ap.add_argument("--max_num_batched_tokens", type=int, default=8192)
    ap.add_argument("--max_model_len", type=int, default=None)
    ap.add_argument("--enforce_eager", action="store_true", help="disable CUDA-graph decode")
    ap.add_argument("--no_mixed_batches", action="store_true",
                    help="prefill-first steps: never put decode rows and prompt rows in one forward pass")
    ap.add_argument("--prefill_chunk", type=int, default=0,
                    help="> 0: chunked prefill, at most this many tokens of one prompt per step (experimental)")
    ap.add_argument("--enable_prefix_caching", action="store_true",
                    help="share the KV blocks of a common prompt prefix between requests (experimental)")
    ap.add_argument("--no_warmup", action="store_true", help="capture the decode CUDA graphs lazily")
    ap.add_argument("--no_async_scheduling", action="store_true",
                    help="read each step's tokens on the host before launching the next step")
    ap.add_argument("--inprocess_frontend", action="store_true",
                    help="run the HTTP server as a thread of the engine process instead of a process of its own")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--served_model_name", type=str, default=None)


def main():
    ap = argparse.ArgumentParser()
    add_args(ap)
    a = ap.parse_args()
    info, cfg, eng = build_engine(a)
    if info.rank != 0:
        eng.worker_loop()
        return
    name = a.served_model_name or a.model or a.model_config or cfg.name
    print(f"serving {name} (tp={info.world_size}) on http://{a.host}:{a.port}", flush=True)
    if not a.inprocess_frontend:
        # HTTP server in its own process; this process (which owns the GPU) only runs the engine loop
        from dlti_b200.serve.api_server import engine_process_loop, start_frontend_process
        proc, conn = start_frontend_process(eng, a.host, a.port, name)
        try:
            engine_process_loop(eng, conn, should_stop=lambda: not proc.is_alive())
        except KeyboardInterrupt:
            eng.shutdown_workers()
        finally:
            proc.terminate()
        return
    import uvicorn

    from dlti_b200.serve.api_server import AsyncEngine, build_app
    aeng = AsyncEngine(eng)
    aeng.start()
    try:
        uvicorn.run(build_app(aeng, name), host=a.host, port=a.port, log_level="warning")
    finally:
        aeng.stop = True
        aeng.thread.join(timeout=10)


if __name__ == "__main__":
    main()
