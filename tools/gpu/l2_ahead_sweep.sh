#!/bin/bash
# L2 look-ahead sweep
cd /root/repo
for d in 0 -32 -64 32; do
  YALM_MEGA_L2_AHEAD=$d timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('ahead $d', round(d['value'],1), round(d['ms_per_step'],3))"
done

YALM_MEGA_L2_AHEAD=-48 timeout 600 python tools/mega_profile.py --out gpurun_out/mega_profile_pre48.txt 2>&1 | tail -12
