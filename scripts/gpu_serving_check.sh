#!/bin/bash
# One GPU-box pass over the serving path: kernel/engine tests, decode-step microbench (A/B over the PDL levels),
# HTTP serving sweep (two-process front-end), launch list of one eager decode step.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_kernels_gpu.py tests/test_serve_gpu.py -x -q > gpurun_out/t_serve.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/t_serve.log
for lvl in ${PDL_LEVELS:-0 1 2}; do
  DLTI_PDL=$lvl timeout 300 python scripts/bench_decode.py --model_config llama2-7b --synthetic --batches ${DEC_BATCHES:-1,16,64,128} > gpurun_out/dec_pdl$lvl.log 2>&1
  echo "dec pdl=$lvl rc=$?"; grep ms_per_step gpurun_out/dec_pdl$lvl.log | cut -c1-100; cp gpurun_out/decode_bench_tp1.json gpurun_out/decode_bench_tp1_pdl$lvl.json
done
timeout 600 python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 1,16,64,128 --out gpurun_out/serving_bench_tp1_v5.json > gpurun_out/serve_v5.log 2>&1
echo "serve rc=$?"; grep '"concurrency"' gpurun_out/serve_v5.log | cut -c1-420 | tail -5
