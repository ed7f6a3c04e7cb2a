#!/bin/bash
set -u
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_stats.so
for B in 1 8; do
  python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 $B 2>&1
  RT_CHAIN_SLOTS=3 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 $B 2>&1
  RT_CHAIN_TUNE=0,16,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 $B 2>&1
done
