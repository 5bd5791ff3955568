#!/bin/bash
cd /root/repo
timeout 900 python tools/mega_profile.py --workload mistral-7b-fp8-32k --pos 32700 --out gpurun_out/mega_profile_fp8_32k.txt 2>&1 | tail -12
