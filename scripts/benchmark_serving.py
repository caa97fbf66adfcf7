#!/usr/bin/env python
"""Serving benchmark (BASELINE.json config 4): start the OpenAI-compatible server in this process
group and drive it with the closed-loop load generator: 512-token prompts, 128 output tokens, a sweep
over concurrent users; reports output tokens/s and p50 TTFT per level.

    python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 1,8,32,64
    python -m dlti_b200.launch --num_gpus=8 --no_local_rank scripts/benchmark_serving.py --model_config llama2-7b --synthetic
"""
import argparse
import asyncio
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scripts.serve import add_args, build_engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    add_args(ap)
    ap.add_argument("--concurrency", type=str, default="1,8,32,64")
    ap.add_argument("--input_len", type=int, default=512)
    ap.add_argument("--output_len", type=int, default=128)
    ap.add_argument("--requests_per_level", type=int, default=None)
    ap.add_argument("--out", type=str, default="gpurun_out/serving_bench.json")
    a = ap.parse_args()
    info, cfg, eng = build_engine(a)
    if info.rank != 0:
        eng.worker_loop()
        return
    import urllib.request

    import uvicorn

    from dlti_b200.ops import _lib
    from dlti_b200.serve.api_server import AsyncEngine, build_app
    from dlti_b200.serve.loadgen import sweep

    aeng = AsyncEngine(eng)
    aeng.start()
    name = a.served_model_name or a.model_config or cfg.name
    server = uvicorn.Server(uvicorn.Config(build_app(aeng, name), host=a.host, port=a.port, log_level="warning"))
    th = threading.Thread(target=server.run, daemon=True)
    th.start()
    url = f"http://{a.host}:{a.port}"
    for _ in range(200):
        try:
            urllib.request.urlopen(url + "/health", timeout=1)
            break
        except Exception:
            time.sleep(0.1)
    levels = [int(x) for x in a.concurrency.split(",")]
    # warm-up: compile-free, but capture the CUDA graphs for the buckets the sweep will hit
    asyncio.run(sweep(url, name, [max(levels)], requests_per_level=max(levels), input_len=a.input_len,
                      output_len=8, vocab=cfg.vocab_size, seed=12345))
    _lib.reset_launch_count()
    res = asyncio.run(sweep(url, name, levels, a.requests_per_level, input_len=a.input_len,
                            output_len=a.output_len, vocab=cfg.vocab_size))
    out = {"model": name, "tp": info.world_size, "backend": a.backend if info.world_size > 1 else "n/a",
           "dtype": a.dtype, "num_kv_blocks": eng.blocks.num_blocks, "cuda_graphs": eng.use_graphs,
           "native_kernel_launches": _lib.launch_count(), "engine_stats": eng.stats,
           "preemptions": eng.sched.num_preemptions, "levels": res}
    print(json.dumps(out))
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(out, open(a.out, "w"), indent=1)
    server.should_exit = True
    aeng.stop = True
    aeng.thread.join(timeout=20)


if __name__ == "__main__":
    main()
