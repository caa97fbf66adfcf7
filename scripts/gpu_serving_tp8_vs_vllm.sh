#!/bin/bash
# TP=8 serving evaluation on one box: correctness at 8 ranks, decode-step bench, HTTP sweep, then vLLM 0.22 TP=8
NG=${NG:-8}; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu_serve.py -x -q -s > gpurun_out/r2_mgs_w$NG.log 2>&1; echo "serve mg rc=$?"; grep -E "agreement|PASSED|passed|failed" gpurun_out/r2_mgs_w$NG.log | tail -6
timeout 200 python -m dlti_b200.launch --num_gpus=$NG --no_local_rank scripts/bench_decode.py --model_config llama2-7b --synthetic --batches 1,16,64,128 > gpurun_out/dec_tp$NG.log 2>&1; echo "dec rc=$?"; grep ms_per_step gpurun_out/dec_tp$NG.log | cut -c1-110
timeout 400 python -m dlti_b200.launch --num_gpus=$NG --no_local_rank scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 1,16,64,128 --out gpurun_out/serving_tp${NG}_r2.json > gpurun_out/serve_tp${NG}.log 2>&1
echo "serve rc=$?"; grep "\"concurrency\"" gpurun_out/serve_tp${NG}.log | cut -c1-40,250-420 | tail -4
timeout 900 python baseline/standin_serve.py --tp $NG --concurrency 1,16,64,128 --startup_timeout 480 > gpurun_out/vllm_tp$NG.log 2>&1
echo "vllm rc=$?"; python - <<PY
import json
try:
    d=json.load(open('gpurun_out/vllm_standin_tp$NG.json')); print('startup_s', d.get('startup_s'))
    for l in d['levels']: print({k:(round(v,1) if isinstance(v,float) else v) for k,v in l.items() if k in ('concurrency','output_tokens_per_s','ttft_p50_ms','itl_mean_ms','errors')})
except Exception as e: print('no vllm result', e)
PY
tail -2 gpurun_out/vllm_tp$NG.log | cut -c1-300
