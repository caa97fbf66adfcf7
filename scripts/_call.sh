timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_model_gpu.py -x -q 2>&1 | tail -3
python bench.py --steps 6 --warmup 3 2>/dev/null > gpurun_out/r2_bench_2prod.json; cut -c1-200 gpurun_out/r2_bench_2prod.json
DLTI_GPU_LIB=$PWD/dlti_b200/lib/variants/libdlti_b200_b420dad.so python bench.py --steps 6 --warmup 3 2>/dev/null > gpurun_out/r2_bench_headlib.json; cut -c1-200 gpurun_out/r2_bench_headlib.json
python bench.py --steps 6 --warmup 3 2>/dev/null > gpurun_out/r2_bench_2prod_b.json; cut -c1-200 gpurun_out/r2_bench_2prod_b.json
