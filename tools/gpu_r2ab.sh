#!/bin/bash
set -u
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_fine.so
for K in 100 106 112 118 125; do
  for B in 8 4 2; do
    RT_SPEC_K0=$K RT_SPEC_MARGIN=1030,10 python tools/prof_case.py cfg4_open 3840 2160 64 samples 3 $B | tail -2 | head -1 | tr '\n' ' '; echo " K0=$K"
  done
done
