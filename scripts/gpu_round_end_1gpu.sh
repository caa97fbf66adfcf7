#!/bin/bash
# 1-GPU round-end pass: full GPU tests, final bench + secondary benches, GEMM microbenches, ncu capture of the K1 kernel
mkdir -p gpurun_out
timeout 1000 python -m pytest tests -m gpu -x -q > gpurun_out/r2_t_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_t_final.log
python bench.py --steps 8 --warmup 3 2>/dev/null > gpurun_out/r2_bench_final_n1.json; cut -c1-160 gpurun_out/r2_bench_final_n1.json
python bench.py --steps 5 --warmup 3 --no_grad_ckpt 2>/dev/null > gpurun_out/r2_bench_nockpt_n1.json; cut -c1-160 gpurun_out/r2_bench_nockpt_n1.json
python bench.py --steps 3 --warmup 3 --offload 2>gpurun_out/offload_err.log > gpurun_out/r2_bench_offload_n1.json; cut -c1-160 gpurun_out/r2_bench_offload_n1.json; tail -2 gpurun_out/offload_err.log | cut -c1-200
for z in 1 2; do python bench.py --steps 4 --warmup 3 --zero_stage $z 2>/dev/null > gpurun_out/r2_bench_zero${z}_n1.json; cut -c1-140 gpurun_out/r2_bench_zero${z}_n1.json; done
OUT=gpurun_out/r2_gemm_sustained_final.json SECS=1.0 python scripts/bench_gemm_sustained.py > /dev/null 2>&1; echo "sust rc=$?"
python scripts/bench_gemm.py > /dev/null 2>&1; cp gpurun_out/bench_gemm.json gpurun_out/r2_gemm_microbench.json
python scripts/profile_k1_local.py > gpurun_out/k1_local_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemm2_kernel -s 2 -c 3 -o gpurun_out/prof_k1_local python scripts/profile_k1_local.py > gpurun_out/k1_local_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/k1_local_plain.log | tail -2
