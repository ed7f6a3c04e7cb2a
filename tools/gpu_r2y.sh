#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "chain" 2>&1 | tail -3
for T in 16,1,32,0 16,3,32,0; do
  for B in 1 2 4 8; do
    RT_CHAIN_TUNE=$T timeout 120 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 $B | tail -2 | tr '\n' ' '; echo " tune=$T"
  done
done
RT_CHAIN_TUNE=16,1,32,0 timeout 300 python tools/prof_case.py cfg4_box 3840 2160 8 chain 3 1 | tail -2 | tr '\n' ' '; echo
RT_CHAIN_TUNE=16,3,32,0 timeout 300 python tools/prof_case.py cfg4_box 3840 2160 8 chain 3 1 | tail -2 | tr '\n' ' '; echo
