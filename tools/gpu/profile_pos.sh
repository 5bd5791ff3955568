#!/bin/bash
cd /root/repo
for pos in 2800 4000; do
timeout 600 python tools/mega_profile.py --pos $pos 2>&1 | grep -E "attn |attn detail|wo  "
done
