#!/bin/bash
# streaming probe with an L2 prefetch running `ahead` tiles in front of the ring
cd /root/repo
for ah in 0 8 32 64; do
YALM_PROBE_AHEAD=$ah python - <<PY
from yalm_b200 import api
us, by = api.bench_op("stream", iters=5, warmup=0)
print("ahead $ah: stream probe GB/s", round(by/us/1e3, 1))
PY
done
