#!/bin/bash
# Round-2 2-GPU call after the kernel refactor: tensor-parallel parity at world 2 (both paths, expert parallelism) and the
# Mistral-7B TP2 bench line.
cd /root/repo
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_tp.py -x -q > gpurun_out/r2b_tp2_tests.log 2>&1; echo "tp tests rc=$?"; tail -3 gpurun_out/r2b_tp2_tests.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus 2 --steps 32 --warmup 4 > gpurun_out/r2b_bench_mistral_tp2.json 2> gpurun_out/r2b_bench_mistral_tp2.err
echo "tp2 bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r2b_bench_mistral_tp2.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['parity'])"
