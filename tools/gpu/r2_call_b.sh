#!/bin/bash
# Round-2 GPU call B (1 GPU): full GPU suite (MoE on the whole-token kernel, CLI, sampler, loader), then the single-GPU
# lines of the sharding workloads (both models fit one 180 GB B200): Mixtral-8x7B and Llama-3-70B-arch.
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/gpu_tests_b.log 2>&1; echo "gpu tests rc=$?"; tail -8 gpurun_out/gpu_tests_b.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_b.log 2>&1; tail -3 gpurun_out/smoke_b.log
timeout 900 python bench.py --workload mixtral-8x7b-fp16-4k --steps 32 --warmup 4 --no-cpu-baseline > gpurun_out/r2_bench_mixtral_tp1.json 2> gpurun_out/r2_bench_mixtral_tp1.err; echo "mixtral rc=$?"; tail -c 1200 gpurun_out/r2_bench_mixtral_tp1.json; tail -3 gpurun_out/r2_bench_mixtral_tp1.err
timeout 1500 python bench.py --workload llama-3-70b-fp16-4k --steps 16 --warmup 3 --no-cpu-baseline --no-gpu-reference > gpurun_out/r2_bench_llama70b_tp1.json 2> gpurun_out/r2_bench_llama70b_tp1.err; echo "70b rc=$?"; tail -c 1200 gpurun_out/r2_bench_llama70b_tp1.json; tail -3 gpurun_out/r2_bench_llama70b_tp1.err
