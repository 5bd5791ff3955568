#!/bin/bash
cd /root/repo
timeout 600 python tools/mega_profile.py --out gpurun_out/mega_profile.txt 2>&1 | tail -3
python tools/_probe.py
