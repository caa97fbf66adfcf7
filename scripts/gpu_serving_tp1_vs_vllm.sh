#!/bin/bash
# TP=1 serving: this framework vs the vLLM stand-in on the same box, same load generator
mkdir -p gpurun_out
timeout 500 python scripts/benchmark_serving.py --model_config llama2-7b --synthetic --concurrency 1,16,64,128 --out gpurun_out/serving_tp1_r2.json > gpurun_out/serve_tp1.log 2>&1
echo "dlti rc=$?"; grep "\"concurrency\"" gpurun_out/serve_tp1.log | cut -c1-120,250-420 | tail -4
timeout 900 python baseline/standin_serve.py --tp 1 --concurrency 1,16,64,128 --startup_timeout 600 > gpurun_out/vllm_tp1.log 2>&1
echo "vllm rc=$?"; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/vllm_standin_tp1.json')); print('startup_s', d.get('startup_s'))
    for l in d['levels']: print({k:(round(v,1) if isinstance(v,float) else v) for k,v in l.items() if k in ('concurrency','output_tokens_per_s','ttft_p50_ms','itl_mean_ms','errors')})
except Exception as e: print('no vllm result', e)
PY
tail -3 gpurun_out/vllm_tp1.log | cut -c1-300
