#!/bin/bash
# ncu launch list of the bench command (per-launch gpu time; cold-cache, serialised)
cd /root/repo
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:decode_mega -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_launches.log; wc -l gpurun_out/launches.csv
