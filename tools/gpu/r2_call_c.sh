#!/bin/bash
# Round-2 GPU call C (1 GPU): prefill tests first (new code), then the rest of the GPU suite, then the prefill bench line.
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prefill.py -q -x > gpurun_out/gpu_tests_prefill.log 2>&1; echo "prefill tests rc=$?"; tail -25 gpurun_out/gpu_tests_prefill.log
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gpu_prefill.py > gpurun_out/gpu_tests_c.log 2>&1; echo "gpu tests rc=$?"; tail -6 gpurun_out/gpu_tests_c.log
timeout 900 python bench.py --workload mistral-7b-fp16-prefill --steps 4 > gpurun_out/r2_bench_prefill.json 2> gpurun_out/r2_bench_prefill.err; echo "prefill bench rc=$?"; tail -c 1500 gpurun_out/r2_bench_prefill.json; tail -3 gpurun_out/r2_bench_prefill.err
