#!/bin/bash
# Round-2 GPU call F (1 GPU): where the token's time goes — the whole-token kernel without GEMV math (1), without attention
# math (4), without copies (8), with fewer ring stages; on the flagged-hb path (YALM_MEGA_W2_PIPE=0) and the pipelined one.
cd /root/repo
mkdir -p gpurun_out
timeout 1200 python tools/mega_ab.py "$@" > gpurun_out/f_ab.log 2>&1; echo "ab rc=$?"; grep -v "loading model" gpurun_out/f_ab.log | tail -20
