#!/bin/bash
# A/B/C of library builds on the same box: parity tests per build, then the bench (64 steps) per build, then the
# phase profile of the in-tree build.  usage: ab3.sh label=path ...   (default: base, in-tree, llattn)
cd /root/repo
LIBS=${@:-"base=/root/repo/build/lib_base.so new=/root/repo/yalm_b200/libyalm_b200.so llattn=/root/repo/build/lib_llattn.so"}
for kv in $LIBS; do
  l=${kv%%=*}; p=${kv#*=}
  [ -f "$p" ] || { echo "$l: missing $p"; continue; }
  YALM_B200_LIB=$p timeout 600 python -m pytest tests/test_gpu_mega.py -x -q > gpurun_out/mega_tests_$l.log 2>&1; echo "$l mega tests rc=$? $(tail -1 gpurun_out/mega_tests_$l.log)"
done
for rep in 1 2; do
for kv in $LIBS; do
  l=${kv%%=*}; p=${kv#*=}
  [ -f "$p" ] || continue
  YALM_B200_LIB=$p timeout 600 python bench.py --steps 64 --warmup 4 --no-cpu-baseline 2>gpurun_out/bench_$l.err | tail -1 > gpurun_out/bench_$l.json
  python -c "import sys,json; d=json.loads(open('gpurun_out/bench_$l.json').readline()); print('$l', round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1))" || tail -3 gpurun_out/bench_$l.err
done
done
timeout 600 python tools/mega_profile.py --out gpurun_out/mega_profile.txt 2>&1 | tail -12
