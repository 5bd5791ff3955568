#!/bin/bash
cd /root/repo
for d in 4 12 9; do
  echo "=== YALM_MEGA_DEBUG=$d"
  YALM_MEGA_DEBUG=$d timeout 300 python tools/mega_profile.py --layers 8 --out gpurun_out/mega_profile_d$d.txt 2>&1 | tail -9
done
