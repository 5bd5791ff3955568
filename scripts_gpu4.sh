#!/bin/bash
cd /root/repo
timeout 300 python tools/mega_profile.py --layers 4 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:decode_mega -s 2 -c 1 -o gpurun_out/mega_r1b -f python tools/mega_profile.py --layers 4 > gpurun_out/ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu.log
