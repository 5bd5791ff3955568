#!/bin/bash
# Round-2 GPU call H (1 GPU): pipelined-hb tests, then the whole GPU suite.
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mega.py -x -q -k "pipelined" > gpurun_out/h_mega_tests.log 2>&1; echo "pipe tests rc=$?"; tail -8 gpurun_out/h_mega_tests.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/h_gpu_tests.log 2>&1; echo "gpu suite rc=$?"; tail -8 gpurun_out/h_gpu_tests.log
