#!/bin/bash
cd /root/repo
N=${1:-4}
timeout 400 python tools/tp_debug.py --world $N --layers 2 --timeout-s 12 > gpurun_out/tp_debug_$N.log 2>&1; echo "tp_debug rc=$?"; grep -v "^$\|loading" gpurun_out/tp_debug_$N.log | tail -25
