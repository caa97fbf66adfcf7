#!/usr/bin/env python
"""Serving benchmark (BASELINE.json config 4): start the OpenAI-compatible server in this process
group and drive it with the closed-loop load generator: 512-token prompts, 128 output tokens, a sweep
over concurrent users; reports output tokens/s and p50 TTFT per level.

    python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 1,8,32,64
    python -m dlti_b200.launch --num_gpus=8 --no_local_rank scripts/benchmark_serving.py --model_config llama2-7b --synthetic
"""
import argparse
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

    from dlti_b200.ops import _lib
    from dlti_b200.serve.api_server import AsyncEngine, build_app, engine_process_loop, start_frontend_process

    name = a.served_model_name or a.model_config or cfg.name
    url = f"http://{a.host}:{a.port}"
    done = threading.Event()
    if a.inprocess_frontend:
        import uvicorn
        aeng = AsyncEngine(eng)
        aeng.start()
        server = uvicorn.Server(uvicorn.Config(build_app(aeng, name), host=a.host, port=a.port, log_level="warning"))
        threading.Thread(target=server.run, daemon=True).start()
    else:
        proc, conn = start_frontend_process(eng, a.host, a.port, name)

    def wait_ready():
        for _ in range(600):
            try:
                urllib.request.urlopen(url + "/health", timeout=1)
                return
            except Exception:
                time.sleep(0.1)
        raise RuntimeError("server did not come up")

    levels = [int(x) for x in a.concurrency.split(",")]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    def client(conc, out_len, seed, per_level=None):
        """The load generator runs as a separate process (its own interpreter: no GIL shared with the engine)."""
        import subprocess
        import tempfile
        with tempfile.NamedTemporaryFile(suffix=".json") as tf:
            cmd = [sys.executable, "-m", "dlti_b200.serve.loadgen", "--url", url, "--model", name, "--concurrency", conc,
                   "--input_len", str(a.input_len), "--output_len", str(out_len), "--vocab", str(cfg.vocab_size),
                   "--seed", str(seed), "--out", tf.name]
            if per_level:
                cmd += ["--requests_per_level", str(per_level)]
            subprocess.run(cmd, check=True, cwd=root, env=dict(os.environ, PYTHONPATH=root, CUDA_VISIBLE_DEVICES=""))
            return json.load(open(tf.name))

    result = {}

    def drive():
        try:
            wait_ready()
            # warm-up: compile-free, but capture the CUDA graphs for the buckets the sweep will hit
            client(str(max(levels)), 8, 12345, max(levels))
            _lib.reset_launch_count()
            for k in eng.stats:
                eng.stats[k] = type(eng.stats[k])(0)
            result["levels"] = client(a.concurrency, a.output_len, 0, a.requests_per_level)
            result["launches"] = _lib.launch_count()
            result["stats"] = dict(eng.stats)
        finally:
            done.set()

    th = threading.Thread(target=drive, daemon=True)
    th.start()
    if a.inprocess_frontend:
        done.wait()
        server.should_exit = True
        aeng.stop = True
        aeng.thread.join(timeout=20)
    else:
        engine_process_loop(eng, conn, should_stop=done.is_set)
        proc.terminate()
    th.join(timeout=5)
    if "levels" not in result:
        raise SystemExit("load generation failed")
    out = {"model": name, "tp": info.world_size, "backend": a.backend if info.world_size > 1 else "n/a",
           "dtype": a.dtype, "num_kv_blocks": eng.blocks.num_blocks, "cuda_graphs": eng.use_graphs,
           "frontend": "thread" if a.inprocess_frontend else "process", "mixed_batches": not a.no_mixed_batches, "async_scheduling": not a.no_async_scheduling,
           "native_kernel_launches": result["launches"], "engine_stats": result["stats"],
           "preemptions": eng.sched.num_preemptions, "levels": result["levels"]}
    print(json.dumps(out))
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(out, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
