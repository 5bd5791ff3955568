#!/bin/bash
# Round-2 GPU call J (1 GPU): A/B on the headline model (tools/mega_ab.py settings as arguments).
cd /root/repo
mkdir -p gpurun_out
timeout 900 python tools/mega_ab.py "$@" > gpurun_out/j_ab2.log 2>&1; echo "ab rc=$?"; grep -v "loading model" gpurun_out/j_ab2.log | tail -12
