#!/bin/bash
# tensor-parallel smoke of the bench at 2 ranks (per-op path + NCCL all-reduce)
cd /root/repo
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 32 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tp8.json 2> gpurun_out/bench_tp8.err
echo "rc=$?"; tail -1 gpurun_out/bench_tp8.json | cut -c1-900; tail -5 gpurun_out/bench_tp8.err
