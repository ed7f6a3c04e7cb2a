#!/bin/bash
set -u
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_stats.so
python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 1 2>&1
python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 8 2>&1
RT_CHAIN_TUNE=16,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 8 2>&1
