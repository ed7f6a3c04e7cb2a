#!/bin/bash
# round-2 call C: chain variant — parity, then timings over slots / margins / shares
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "chain" > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2c_pytest.log
for B in 1 8; do
  python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
  RT_CHAIN_SLOTS=3 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
  RT_CHAIN_SLOTS=4 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
  RT_CHAIN_TUNE=0,16,1 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
  RT_CHAIN_TUNE=2,16,1 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
  RT_CHAIN_TUNE=4,16,1 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
  RT_CHAIN_TUNE=1,16,0 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B
done
python tools/prof_case.py cfg4_open 3840 2160 64 auto 3 1
python tools/prof_case.py cfg4_open 3840 2160 64 auto 3 2
python tools/prof_case.py cfg4_open 3840 2160 64 auto 3 4
python tools/prof_case.py cfg4_box 3840 2160 64 auto 2 8
python tools/prof_case.py cfg3_open 1920 1080 16 auto 3 1
python tools/prof_case.py cfg3_open 1920 1080 16 persistent 3 1
python tools/prof_case.py cfg3_open 1920 1080 16 chain 3 1
