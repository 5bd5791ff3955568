#!/bin/bash
# A/B of two library builds on the same box: build/lib_base.so vs the in-tree library; then tests + phase profile
cd /root/repo
one() { # label lib
  YALM_B200_LIB=$2 timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('$1', round(d['value'],1), round(d['ms_per_step'],3))"
}
timeout 900 python -m pytest tests/test_gpu_mega.py -x -q 2>&1 | tail -3
one base /root/repo/build/lib_base.so
one new /root/repo/yalm_b200/libyalm_b200.so
timeout 600 python tools/mega_profile.py --out gpurun_out/mega_profile.txt 2>&1 | tail -12
python tools/_probe.py
