#!/bin/bash
# round-2 call H: chain v2 with global rings and a wide window
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "chain" > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2h_pytest.log
for B in 1 2 4 8; do python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B; done
for W in 4 8 16; do RT_CHAIN_TUNE=16,1,$W python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8; done
RT_CHAIN_TUNE=8,1,32 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8
RT_CHAIN_TUNE=4,1,32 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8
RT_CHAIN_TUNE=8,1,32 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1
export RT_B200_LIB=$PWD/opengl_raytracer_b200/csrc/librt_b200_stats.so
python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 8 2>&1
RT_CHAIN_TUNE=4,1,32 python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 8 2>&1
