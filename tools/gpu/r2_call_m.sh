#!/bin/bash
# Round-2 last GPU call (1 GPU): the whole GPU suite and smoke() on the final tree.
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2c_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/r2c_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2c_smoke.log 2>&1; tail -3 gpurun_out/r2c_smoke.log
