#!/bin/bash
# one --set full capture of the whole-token kernel on the bench workload
cd /root/repo
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:decode_mega -s 3 -c 1 -o gpurun_out/mega_full -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_full.log
