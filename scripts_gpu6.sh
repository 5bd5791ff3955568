#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_mega.py -x -q 2>&1 | tail -2
for ah in 0 16 32 64; do
  echo "=== L2_AHEAD=$ah"
  YALM_MEGA_L2_AHEAD=$ah timeout 600 python tools/mega_profile.py --layers 8 2>&1 | tail -10
done
