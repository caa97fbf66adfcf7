#!/bin/bash
# Round-2 baseline call (1 GPU): HEAD's GPU tests, bench, stand-in bench, sustained GEMM microbench.
mkdir -p gpurun_out
timeout 1000 python -m pytest tests -m gpu -x -q > gpurun_out/r2_t_all.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_t_all.log
python bench.py --steps 6 --warmup 3 > gpurun_out/r2_bench_head.json 2> gpurun_out/r2_bench_err.log; echo "bench rc=$?"; cut -c1-300 gpurun_out/r2_bench_head.json
python bench.py --impl standin --steps 6 --warmup 3 > gpurun_out/r2_standin_n1.json 2> gpurun_out/r2_standin_err.log; echo "standin rc=$?"; cut -c1-300 gpurun_out/r2_standin_n1.json
OUT=gpurun_out/r2_gemm_sustained_head.json python scripts/bench_gemm_sustained.py > gpurun_out/r2_gemm_sustained_head.log 2>&1; echo "gemm rc=$?"
