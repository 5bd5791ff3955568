#!/bin/bash
# L2 look-ahead sweep
cd /root/repo
for d in 0 8 24 48 96; do
  YALM_MEGA_L2_AHEAD=$d timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('ahead $d', round(d['value'],1), round(d['ms_per_step'],3))"
done
YALM_MEGA_L2_AHEAD=32 timeout 900 python -m pytest tests/test_gpu_mega.py -x -q 2>&1 | tail -3
YALM_MEGA_L2_AHEAD=32 timeout 600 python tools/mega_profile.py --out gpurun_out/mega_profile_a32.txt 2>&1 | tail -12
