#!/bin/bash
set -u
for T in 1,3,32,0 2,3,32,0 4,3,32,0 8,3,32,0 16,3,32,0 4,1,32,0 8,1,32,0; do
  RT_CHAIN_TUNE=$T timeout 120 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8 | tail -2 | head -1 | tr '\n' ' '; echo " tune=$T"
done
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_stats.so
RT_CHAIN_TUNE=1,3,32,0 timeout 120 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 8 2>&1 | tail -7
