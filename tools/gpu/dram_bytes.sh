#!/bin/bash
cd /root/repo
timeout 300 python tools/mega_profile.py --layers 4 > gpurun_out/plain.log 2>&1 || exit 1
for ah in 0 32; do
YALM_MEGA_L2_AHEAD=$ah timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:decode_mega -s 2 -c 1 python tools/mega_profile.py --layers 4 2>&1 | grep -E "dram__bytes|gpu__time|lts__" 
done
