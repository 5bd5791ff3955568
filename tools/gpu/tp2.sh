#!/bin/bash
# tensor parallel at 2 ranks: parity of the sharded tiny models, then the bench (per-op path + all-reduce)
cd /root/repo
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/tp_check.py 2>&1 | grep -E "PASS|FAIL|Error|error|all-reduce over" | head -20
for peer in 1 0; do
YALM_TP_PEER=$peer timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 32 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tp2_peer$peer.json 2> gpurun_out/bench_tp2_peer$peer.err
echo "peer=$peer rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_tp2_peer$peer.json').read().strip().splitlines()[-1]); print(d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],3), d['e2e']['value'])"
grep -E "all-reduce over|rror" gpurun_out/bench_tp2_peer$peer.err | head -3
done
