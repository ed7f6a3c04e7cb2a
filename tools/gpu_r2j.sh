#!/bin/bash
set -u
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_stats.so
RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 1 2>&1
RT_CHAIN_TUNE=32,1,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 1 2>&1
RT_CHAIN_TUNE=64,1,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 1 2>&1
RT_CHAIN_TUNE=16,0,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 1 2>&1
unset RT_B200_LIB
RT_CHAIN_TUNE=32,1,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1
RT_CHAIN_TUNE=64,1,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1
RT_CHAIN_TUNE=16,0,32,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1
RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg4_box 3840 2160 8 chain 3 1
RT_CHAIN_TUNE=16,0,32,0 python tools/prof_case.py cfg4_box 3840 2160 8 chain 3 1
RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg4_box 3840 2160 64 chain 2 8
python tools/prof_case.py cfg4_box 3840 2160 64 persistent 2 8
RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg3_open 1920 1080 16 chain 3 1
RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg2_open 1280 720 8 chain 3 1
python tools/prof_case.py cfg2_open 1280 720 8 persistent 3 1
RT_CHAIN_TUNE=16,1,32,0 python tools/prof_case.py cfg3_box 1920 1080 16 chain 3 1
python tools/prof_case.py cfg3_box 1920 1080 16 persistent 3 1
