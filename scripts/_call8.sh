#!/bin/bash
mkdir -p gpurun_out
run() { echo "== $*"; env "$@" timeout 200 python -m dlti_b200.launch --num_gpus=2 --master_port=29741 --no_local_rank tests/mp_serve_worker.py 2>&1 | grep -E "agreement|PASSED|Error|assert" | cut -c1-230; }
run X=1
run MP_SERVE_PRESET=tiny-2h
