#!/bin/bash
# TP serving checks on NG GPUs: token agreement with TP=1, decode-step bench
NG=${NG:-2}; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu_serve.py -x -q -s > gpurun_out/r2_mgs_w$NG.log 2>&1; echo "serve mg rc=$?"; grep -E "agreement|PASSED|passed|failed|Error" gpurun_out/r2_mgs_w$NG.log | tail -12
timeout 300 python -m dlti_b200.launch --num_gpus=$NG --no_local_rank scripts/bench_decode.py --model_config llama2-7b --synthetic --batches 1,16,64,128 > gpurun_out/dec_tp$NG.log 2>&1; echo "dec rc=$?"; grep ms_per_step gpurun_out/dec_tp$NG.log | cut -c1-150; tail -3 gpurun_out/dec_tp$NG.log | cut -c1-300
