#!/bin/bash
# Round-2 GPU call D (1 GPU): the no-allocation entry point's test alone in a fresh process, then the whole GPU suite
# (which collects it last), then smoke().
cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/gpu_tests_no_alloc_alone.log 2>&1; echo "no-alloc alone rc=$?"; tail -15 gpurun_out/gpu_tests_no_alloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_d.log 2>&1; echo "gpu suite rc=$?"; tail -12 gpurun_out/gpu_tests_d.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_d.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke_d.log
