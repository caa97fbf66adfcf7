#!/bin/bash
# Same-box comparators for the training headline at NG GPUs:
#   dlti fused backend | dlti NCCL backend | stand-in DDP (frozen base replicated) | stand-in ZeRO-3 (NCCL gathers)
NG=${NG:-2}; STEPS=${STEPS:-6}
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $NG --steps $STEPS --warmup 3 "$@"; }
run                     2>gpurun_out/cmp_err_fused.log  | grep '^{' > gpurun_out/cmp_dlti_fused_n$NG.json;  echo "fused rc=$?";  cut -c1-200 gpurun_out/cmp_dlti_fused_n$NG.json
run --backend nccl      2>gpurun_out/cmp_err_nccl.log   | grep '^{' > gpurun_out/cmp_dlti_nccl_n$NG.json;   echo "nccl rc=$?";   cut -c1-200 gpurun_out/cmp_dlti_nccl_n$NG.json
STANDIN_MODE=zero3 run --impl standin 2>gpurun_out/cmp_err_sz3.log | grep '^{' > gpurun_out/cmp_standin_zero3_n$NG.json; echo "standin zero3 rc=$?"; cut -c1-330 gpurun_out/cmp_standin_zero3_n$NG.json
[ -n "$SKIP_DDP" ] || { run --impl standin 2>gpurun_out/cmp_err_sddp.log | grep '^{' > gpurun_out/cmp_standin_ddp_n$NG.json; echo "standin ddp rc=$?"; cut -c1-330 gpurun_out/cmp_standin_ddp_n$NG.json; }
tail -3 gpurun_out/cmp_err_sz3.log
