#!/bin/bash
# Round-2 GPU call G (1 GPU): per-tile event trace of CTA 0 (library built with -DYALM_MEGA_TRACE) on the pipelined-hb path and
# on the flagged-hb path; analysed offline (tools/mega_trace.py).
cd /root/repo
mkdir -p gpurun_out
export YALM_B200_LIB=/root/repo/build/lib_trace.so
YALM_MEGA_TRACE=gpurun_out/g_trace_pipe.bin timeout 600 python tools/mega_profile.py --layers 8 --out gpurun_out/g_profile_pipe.txt > gpurun_out/g_pipe.log 2>&1; echo "pipe rc=$?"; tail -3 gpurun_out/g_pipe.log
YALM_MEGA_W2_PIPE=0 YALM_MEGA_TRACE=gpurun_out/g_trace_flag.bin timeout 600 python tools/mega_profile.py --layers 8 --out gpurun_out/g_profile_flag.txt > gpurun_out/g_flag.log 2>&1; echo "flag rc=$?"; tail -3 gpurun_out/g_flag.log
ls -la gpurun_out/*.bin
