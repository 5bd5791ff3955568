#!/bin/bash
# Round-2 GPU call I (1 GPU): the Llama-3-70B-arch single-GPU line again, now with the pipelined hb hand-over (20 ring stages
# instead of 14).
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python bench.py --workload llama-3-70b-fp16-4k --steps 16 --warmup 3 --no-cpu-baseline --no-gpu-reference > gpurun_out/r2b_bench_llama70b_tp1.json 2> gpurun_out/r2b_bench_llama70b_tp1.err; echo "70b rc=$?"; tail -c 1500 gpurun_out/r2b_bench_llama70b_tp1.json; tail -3 gpurun_out/r2b_bench_llama70b_tp1.err
