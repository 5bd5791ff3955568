#!/bin/bash
cd /root/repo
for cb in 16384 8192 4096 2048 1024; do
YALM_PROBE_COPY=$cb python - <<PY
from yalm_b200 import api
us, by = api.bench_op("stream", iters=5, warmup=0)
print("copy_bytes $cb: stream probe GB/s", by/us/1e3)
PY
done
