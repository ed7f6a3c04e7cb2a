#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "chain or samples" 2>&1 | tail -3
for H in 1 4 6 8 12; do
  echo "hitLanes $H"
  RT_VOTE_TUNE=0,0,0,$H python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 1 | tail -1
done
RT_VOTE_TUNE=0,0,0,6 python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 8 | tail -1
RT_VOTE_TUNE=0,0,0,6 python tools/prof_case.py cfg4_box 3840 2160 8 chain 3 1 | tail -1
RT_VOTE_TUNE=0,0,0,6 python tools/prof_case.py cfg3_open 1920 1080 64 chain 3 1 | tail -1
