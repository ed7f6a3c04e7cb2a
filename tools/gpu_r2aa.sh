#!/bin/bash
set -u
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_fine.so
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "samples" 2>&1 | tail -2
for L in librt_b200.so librt_b200_fine.so; do
  export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/$L
  for M in 1125,20 1060,15 1030,10 1000,5; do
    for B in 8 2; do
      RT_SPEC_MARGIN=$M python tools/prof_case.py cfg4_open 3840 2160 64 samples 3 $B | tail -2 | head -1 | tr '\n' ' '; echo " $L margin=$M"
    done
  done
done
