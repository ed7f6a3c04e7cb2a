#!/bin/bash
# round-2 call F: chain v2 (pixel per lane + in-warp sample stealing)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "chain" > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2f_pytest.log
for B in 1 2 4 8; do
  python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
done
RT_CHAIN_TUNE=16,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1
RT_CHAIN_TUNE=16,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8
RT_CHAIN_TUNE=8,1 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1
RT_CHAIN_TUNE=8,1 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8
RT_CHAIN_TUNE=12,1 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8
RT_CHAIN_TUNE=24,1 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8
python tools/prof_case.py cfg4_box 3840 2160 8 chain 3 1
python tools/prof_case.py cfg4_box 3840 2160 64 chain 2 8
python tools/prof_case.py cfg3_open 1920 1080 16 chain 3 1
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_stats.so
python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 1 2>&1
python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 8 2>&1
