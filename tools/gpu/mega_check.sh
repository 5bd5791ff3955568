#!/bin/bash
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_mega.py -x -q > gpurun_out/mega_tests.log 2>&1; echo "mega tests rc=$?"; tail -5 gpurun_out/mega_tests.log
timeout 600 python tools/mega_profile.py --out gpurun_out/mega_profile.txt 2>&1 | tail -14
timeout 900 python bench.py --steps 64 --warmup 4 --no-cpu-baseline > gpurun_out/bench_mega.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_mega.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'])"
