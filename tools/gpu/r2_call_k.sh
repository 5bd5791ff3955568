#!/bin/bash
cd /root/repo
mkdir -p gpurun_out
YALM_MEGA_MMA=1 timeout 600 python tools/mega_profile.py --out gpurun_out/k_profile_mma.txt > gpurun_out/k_profile_mma.log 2>&1; echo "profile rc=$?"; tail -14 gpurun_out/k_profile_mma.log
YALM_MEGA_MMA=1 YALM_MEGA_DEBUG=9 timeout 600 python tools/mega_profile.py --out gpurun_out/k_profile_mma9.txt > gpurun_out/k_profile_mma9.log 2>&1; echo "profile9 rc=$?"; tail -14 gpurun_out/k_profile_mma9.log
