#!/bin/bash
NG=${NG:-8}; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_multigpu_serve.py -q -s > gpurun_out/r2_mgs_final_w$NG.log 2>&1; echo "serve mg rc=$?"; grep -E "agreement|passed|failed|PASSED" gpurun_out/r2_mgs_final_w$NG.log | tail -6 | cut -c1-220
timeout 200 python -m dlti_b200.launch --num_gpus=$NG --no_local_rank scripts/bench_gather2.py 2>&1 | grep '^{' | cut -c1-330
