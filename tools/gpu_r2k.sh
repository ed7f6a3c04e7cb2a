#!/bin/bash
set -u
for B in 6 8; do
  export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_burst$B.so
  echo "burst $B"
  python tools/prof_case.py cfg4_open 3840 2160 64 persistent 3 1
  RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1
  python tools/prof_case.py cfg4_box 3840 2160 8 persistent 3 1
done
unset RT_B200_LIB
echo "thresholds"
for T in 6,10 8,12 8,10 10,12 6,12 10,16; do
  RT_VOTE_TUNE=$T python tools/prof_case.py cfg4_open 3840 2160 64 persistent 3 1 | tail -1
  RT_VOTE_TUNE=$T,0 RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1 | tail -1
  RT_VOTE_TUNE=$T python tools/prof_case.py cfg4_box 3840 2160 8 persistent 3 1 | tail -1
done
