#!/bin/bash
# Round-2 8-GPU call: the Mixtral-8x7B TP8 bench line (BASELINE config 4's widest split).
cd /root/repo
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 8 --workload mixtral-8x7b-fp16-4k --steps 32 --warmup 4 > gpurun_out/r2_bench_mixtral_tp8.json 2> gpurun_out/r2_bench_mixtral_tp8.err
echo "mixtral tp8 rc=$?"; tail -c 700 gpurun_out/r2_bench_mixtral_tp8.json; echo; grep -iE "error|Traceback|PARITY|timed out" gpurun_out/r2_bench_mixtral_tp8.err | head -5
