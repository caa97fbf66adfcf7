#!/bin/bash
# One GPU-box pass over the serving path: kernel/engine tests, decode-step microbench, HTTP serving sweep
# (two-process front-end; async vs synchronous scheduling).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_kernels_gpu.py tests/test_serve_gpu.py -x -q > gpurun_out/t_serve.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/t_serve.log
timeout 300 python scripts/bench_decode.py --model_config llama2-7b --synthetic --batches ${DEC_BATCHES:-1,16,64,128} > gpurun_out/dec.log 2>&1
echo "dec rc=$?"; grep ms_per_step gpurun_out/dec.log | cut -c1-100
timeout 600 python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 1,16,64,128 --out gpurun_out/serving_bench_tp1_v6.json > gpurun_out/serve_v6.log 2>&1
echo "serve rc=$?"; grep '"concurrency"' gpurun_out/serve_v6.log | cut -c1-420 | tail -5
timeout 600 python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 16,64 --no_async_scheduling --out gpurun_out/serving_bench_tp1_v6_sync.json > gpurun_out/serve_v6_sync.log 2>&1
echo "serve sync rc=$?"; grep '"concurrency"' gpurun_out/serve_v6_sync.log | cut -c1-420 | tail -3
# opt-in features awaiting their first GPU measurement (round 2): SwiGLU-epilogue GEMM in prefill, chunked prefill
if [ -n "$OPTIN" ]; then
  DLTI_SERVE_FUSE_SWIGLU=1 timeout 600 python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 16,64 --out gpurun_out/serving_bench_tp1_fuse_swiglu.json > gpurun_out/serve_fuse_swiglu.log 2>&1
  echo "serve fuse_swiglu rc=$?"; grep '"concurrency"' gpurun_out/serve_fuse_swiglu.log | cut -c1-420 | tail -2
  timeout 600 python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 16,64 --prefill_chunk 2048 --out gpurun_out/serving_bench_tp1_chunk2048.json > gpurun_out/serve_chunk.log 2>&1
  echo "serve chunk rc=$?"; grep '"concurrency"' gpurun_out/serve_chunk.log | cut -c1-420 | tail -2
fi
