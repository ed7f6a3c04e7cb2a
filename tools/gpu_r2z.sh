#!/bin/bash
# timelines of the chain kernel (stats build): evaluations started / samples consumed / pixels finished per 0.25 ms
set -u
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_stats.so
for T in 16,1,32,0 16,3,32,0; do
  for B in 1 8; do
    echo "== tune=$T bands=$B"
    RT_CHAIN_TUNE=$T timeout 120 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 $B 2>&1 | tail -8
  done
done
