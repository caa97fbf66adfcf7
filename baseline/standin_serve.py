#!/usr/bin/env python
"""STAND-IN serving baseline: vLLM (installed on the image, sm_100 kernels) serving a random-init
Llama-2-7B (`--load-format dummy`, no tokenizer) driven by the SAME closed-loop load generator as
scripts/benchmark_serving.py. Not the reference (which pins vllm==0.6.0 and ships no serving code).

    python baseline/standin_serve.py --tp 1 --concurrency 1,16,64
"""
import argparse, asyncio, json, os, subprocess, sys, tempfile, time, urllib.request

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dlti_b200.models.config import resolve_config  # noqa: E402
from dlti_b200.serve.loadgen import sweep  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tp", type=int, default=1)
ap.add_argument("--concurrency", default="1,16,64")
ap.add_argument("--port", type=int, default=8030)
ap.add_argument("--startup_timeout", type=int, default=420)
ap.add_argument("--compile_mode", type=int, default=None, help="vLLM compilation mode (None = vLLM's default)")
a = ap.parse_args()
cfg = resolve_config("llama2-7b")
d = tempfile.mkdtemp(prefix="llama7b_cfg_")
cfg.save(d)
cmd = [sys.executable, "-m", "vllm.entrypoints.openai.api_server", "--model", d, "--served-model-name", "llama2-7b",
       "--load-format", "dummy", "--skip-tokenizer-init", "--tensor-parallel-size", str(a.tp), "--dtype", "bfloat16",
       "--max-model-len", "4096", "--port", str(a.port), "--host", "127.0.0.1", "--no-enable-log-requests"]
if a.compile_mode is not None:
    cmd += ["--compilation-config", json.dumps({"mode": a.compile_mode})]
env = dict(os.environ, VLLM_NO_USAGE_STATS="1", HF_HUB_OFFLINE="1", TRANSFORMERS_OFFLINE="1")
log = open("gpurun_out/vllm_server.log", "w") if os.path.isdir("gpurun_out") else subprocess.DEVNULL
srv = subprocess.Popen(cmd, env=env, stdout=log, stderr=subprocess.STDOUT, start_new_session=True)
url = f"http://127.0.0.1:{a.port}"
ok = False
t0 = time.time()
while time.time() - t0 < a.startup_timeout and srv.poll() is None:
    try:
        urllib.request.urlopen(url + "/health", timeout=2)
        ok = True
        break
    except Exception:
        time.sleep(2)
startup_s = time.time() - t0
try:
    if not ok:
        print(json.dumps({"impl": "standin-vllm", "unavailable": f"vLLM server did not come up in {a.startup_timeout}s (exit {srv.poll()})"}))
    else:
        levels = [int(x) for x in a.concurrency.split(",")]
        extra = {"detokenize": False}
        asyncio.run(sweep(url, "llama2-7b", [max(levels)], requests_per_level=max(levels), input_len=512, output_len=8,
                          vocab=cfg.vocab_size, extra=extra, seed=12345))
        res = asyncio.run(sweep(url, "llama2-7b", levels, input_len=512, output_len=128, vocab=cfg.vocab_size, extra=extra))
        out = {"impl": "standin-vllm", "tp": a.tp, "startup_s": startup_s, "compile_mode": a.compile_mode, "levels": res}
        print(json.dumps(out))
        if os.path.isdir("gpurun_out"):
            json.dump(out, open(f"gpurun_out/vllm_standin_tp{a.tp}.json", "w"), indent=1)
finally:
    import signal
    try:
        os.killpg(srv.pid, signal.SIGTERM)
    except ProcessLookupError:
        pass
