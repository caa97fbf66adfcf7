#!/bin/bash
NG=${NG:-2}; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py tests/test_multigpu_serve.py -q -s > gpurun_out/r2_mg_final_w$NG.log 2>&1; echo "mg rc=$?"; grep -E "\[ok\]|\[FAIL\]|agreement|passed|failed|PASSED" gpurun_out/r2_mg_final_w$NG.log | tail -45 | cut -c1-200
timeout 300 python -m dlti_b200.launch --num_gpus=$NG --no_local_rank scripts/bench_gather2.py 2>&1 | grep '^{' | cut -c1-330
