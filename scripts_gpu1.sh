#!/bin/bash
# first GPU check of the megakernel: probe, mega tests, full gpu suite, short bench
cd /root/repo
python - <<'PY' > gpurun_out/probe.log 2>&1
from yalm_b200 import api
us, by = api.bench_op("stream", iters=5, warmup=0)
print("stream probe GB/s", by/us/1e3)
PY
cat gpurun_out/probe.log
timeout 900 python -m pytest tests/test_gpu_mega.py -x -q > gpurun_out/mega_tests.log 2>&1; echo "mega tests rc=$?"; tail -30 gpurun_out/mega_tests.log
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -15 gpurun_out/gpu_tests.log
timeout 900 python bench.py --steps 64 --warmup 4 > gpurun_out/bench_mega.log 2>&1; echo "bench rc=$?"; tail -3 gpurun_out/bench_mega.log
