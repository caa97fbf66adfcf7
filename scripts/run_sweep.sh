#!/bin/bash
# ZeRO-1/2/3 comparison sweep through the real entrypoints (BASELINE.json config 3), synthetic Llama-2-7B,
# bf16, seq 2048, micro 2 x accum 4. Usage: scripts/run_sweep.sh <num_gpus> [stages...]; appends to
# results/training_metrics.csv and prints scripts/compare_training.py.
set -uo pipefail
N=${1:-1}; shift || true
STAGES=${@:-"1 2 3"}
COMMON="--synthetic --model_config llama2-7b --bf16 --max_length 2048 --fixed_length --num_samples 4096 --max_steps 6 --per_device_train_batch_size 2 --gradient_accumulation_steps 4 --logging_steps 2"
run() { if [ "$N" = "1" ]; then python "$@"; else python -m dlti_b200.launch --num_gpus=$N "$@"; fi; }
if [ "$N" = "1" ]; then python training/train_baseline.py $COMMON --output_dir /tmp/ckpt/baseline 2>&1 | grep -E "^\{'loss|Throughput|Error|error" | tail -4; fi
for S in $STAGES; do
  FLAG=--deepspeed_config; [ "$S" = "3" ] && FLAG=--deepspeed
  run training/train_deepspeed_zero$S.py $FLAG configs/ds_config_zero$S.json $COMMON --no_offload --output_dir /tmp/ckpt/z$S 2>&1 | grep -E "^\{'loss|Throughput|Tokens/sec|Error|error|Traceback" | tail -4
done
mkdir -p gpurun_out; cp results/training_metrics.csv gpurun_out/training_metrics_${N}gpu.csv
python scripts/compare_training.py 2>&1 | tail -30
