#!/bin/bash
# First GPU call of a new round (1 GPU): everything that was left unmeasured, in one box acquisition.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_first_call.sh'
# Outputs land in gpurun_out/ (copy what matters to profiles/).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t_all.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/t_all.log
grep -n "XFAIL\|xfail\|XPASS" gpurun_out/t_all.log | head -5
timeout 200 python -m pytest tests/test_serve_gpu.py -q -rxX -k chunked > gpurun_out/t_chunked.log 2>&1; tail -15 gpurun_out/t_chunked.log
python bench.py --steps 4 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench_err.log; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json
OPTIN=1 bash scripts/gpu_serving_check.sh
timeout 600 python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 16,64 --enable_prefix_caching --out gpurun_out/serving_bench_tp1_apc_unique.json > gpurun_out/serve_apc.log 2>&1
echo "serve apc rc=$?"; grep '"concurrency"' gpurun_out/serve_apc.log | cut -c1-300 | tail -2
