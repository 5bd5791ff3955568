#!/bin/bash
cd /root/repo
for d in 0 1 8 5 9 13; do
  echo "=== YALM_MEGA_DEBUG=$d"
  YALM_MEGA_DEBUG=$d timeout 300 python tools/mega_profile.py --layers 8 --out gpurun_out/mega_profile_d$d.txt 2>&1 | tail -9
done
for st in 2 3; do
  echo "=== stages=$st"
  YALM_MEGA_STAGES=$st timeout 300 python tools/mega_profile.py --layers 8 --out gpurun_out/mega_profile_s$st.txt 2>&1 | tail -9
done
