#!/bin/bash
# A/B of two library builds on the same box: build/lib_base.so (another commit) vs the in-tree library, interleaved twice.
cd /root/repo
mkdir -p gpurun_out
one() { # label lib
  YALM_B200_LIB=$2 timeout 600 python tools/mega_ab.py 2>/dev/null | grep -v "loading model" | sed "s/^default/$1/"
}
for rep in 1 2; do
one base /root/repo/build/lib_base.so
one new /root/repo/yalm_b200/libyalm_b200.so
done | tee gpurun_out/ab_libs.log
