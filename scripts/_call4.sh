#!/bin/bash
NG=${NG:-2}; mkdir -p gpurun_out
for mode in fused pull; do
DLTI_TP_LARGE_MODE=$mode timeout 400 python -m dlti_b200.launch --num_gpus=$NG --no_local_rank scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency ${CONC:-16,64} --out gpurun_out/serving_tp${NG}_$mode.json > gpurun_out/serve_tp${NG}_$mode.log 2>&1
echo "serve $mode rc=$?"; grep "\"concurrency\"" gpurun_out/serve_tp${NG}_$mode.log | cut -c1-330 | tail -4
done
