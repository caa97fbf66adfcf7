timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_model_gpu.py -x -q 2>&1 | tail -4
OUT=gpurun_out/r2_gemm_sust_tailsplit2.json SECS=1.0 python scripts/bench_gemm_sustained.py > gpurun_out/sust2.log 2>&1; echo rc=$?
python scripts/bench_gemm.py > gpurun_out/burst2.log 2>&1; cp gpurun_out/bench_gemm.json gpurun_out/r2_gemm_burst_tailsplit2.json
python bench.py --steps 6 --warmup 3 2>/dev/null > gpurun_out/r2_bench_tailsplit2.json; cut -c1-200 gpurun_out/r2_bench_tailsplit2.json
