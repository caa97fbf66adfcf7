#!/bin/bash
# 2-GPU validation: peer-memory kernels, K1 on the CTA-pair GEMM, fused vs nccl training; then the 2-GPU bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -x -q -s > gpurun_out/r2_mg_w${NG:-2}.log 2>&1; echo "mg rc=$?"; grep -E "\[ok\]|\[FAIL\]|stage|passed|failed|Error|error" gpurun_out/r2_mg_w${NG:-2}.log | tail -40
python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NG:-2} --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus ${NG:-2} --steps 6 --warmup 3 2>gpurun_out/r2_bench2_err.log > gpurun_out/r2_bench_n${NG:-2}.json; echo "bench2 rc=$?"; cut -c1-250 gpurun_out/r2_bench_n${NG:-2}.json; tail -5 gpurun_out/r2_bench2_err.log
