#!/bin/bash
# tensor parallel at N ranks (default 2): parity of the sharded tiny models, then the bench with the in-kernel
# all-reduce (megakernel) and, for comparison, on the per-op kernels + one-shot all-reduce.
N=${1:-2}
cd /root/repo
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 tests/tp_worker.py 2>&1 | grep -E "PASS|FAIL|Error|error|all-reduce over|Traceback" | head -40
for mega in ${2:-1 0}; do
YALM_TP_MEGA=$mega timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 64 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tp${N}_mega$mega.json 2> gpurun_out/bench_tp${N}_mega$mega.err
echo "mega=$mega rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_tp${N}_mega$mega.json').read().strip().splitlines()[-1]); print(d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],3), d['e2e']['value'], d['gpu_launches'])"
grep -E "rror|FATAL" gpurun_out/bench_tp${N}_mega$mega.err | head -3
done
