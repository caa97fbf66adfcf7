#!/bin/bash
# Multi-GPU serving pass (run with gpurun --gpus N): TP tests, decode-step A/B (barrier fused into the reduce
# kernel, PDL), HTTP sweep.
N=${1:-2}
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then timeout 900 python -m pytest tests/test_multigpu_serve.py -x -q > gpurun_out/t_mg.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/t_mg.log; fi
for cfg in $(echo ${AB:-1:1,0:1,1:0} | tr ',' ' '); do
  fb=${cfg%%:*}; pd=${cfg##*:}
  DLTI_TP_FUSE_BARRIER=$fb DLTI_PDL=$pd timeout 300 python -m dlti_b200.launch --num_gpus=$N --no_local_rank scripts/bench_decode.py --model_config llama2-7b --synthetic --batches 1,16,64,128 > gpurun_out/dec_tp${N}_b${fb}_p${pd}.log 2>&1
  echo "dec fuse_barrier=$fb pdl=$pd rc=$?"; grep ms_per_step gpurun_out/dec_tp${N}_b${fb}_p${pd}.log | cut -c1-90
done
timeout 500 python -m dlti_b200.launch --num_gpus=$N --no_local_rank scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency ${CONC:-1,16,64} --out gpurun_out/serving_bench_tp${N}_v7.json > gpurun_out/serve_tp${N}.log 2>&1
echo "serve rc=$?"; grep "\"concurrency\"" gpurun_out/serve_tp${N}.log | cut -c1-400 | tail -4
