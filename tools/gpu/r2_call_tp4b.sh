#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_tp.py -q -k "expert_parallel" > gpurun_out/ep2_parity.log 2>&1; echo "ep parity rc=$?"; grep -E "PASS|FAIL|passed|failed" gpurun_out/ep2_parity.log | head -8
PARITY_PATHS="" bash tools/gpu/r2_call_tp.sh 4 mistral-7b-fp16-4k mixtral-8x7b-fp16-4k
