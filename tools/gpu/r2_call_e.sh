#!/bin/bash
# Round-2 GPU call E (1 GPU): pipelined hb hand-over — parity of the whole-token kernel, then A/B against the flagged
# hand-over for hb (tools/mega_ab.py: the model is re-planned per setting), then the phase profile.
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mega.py tests/test_gpu_model.py -x -q > gpurun_out/e_mega_tests.log 2>&1; echo "mega+model tests rc=$?"; tail -6 gpurun_out/e_mega_tests.log
timeout 600 python tools/mega_ab.py "YALM_MEGA_W2_PIPE=0" ${AB_EXTRA} > gpurun_out/e_ab.log 2>&1; echo "ab rc=$?"; tail -8 gpurun_out/e_ab.log
timeout 600 python tools/mega_profile.py --out gpurun_out/e_mega_profile.txt > gpurun_out/e_profile.log 2>&1; echo "profile rc=$?"; tail -14 gpurun_out/e_profile.log
