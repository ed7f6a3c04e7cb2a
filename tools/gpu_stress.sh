#!/bin/bash
# Perturb the scheduling of the speculating variants and re-run their parity tests: occupancy, quota, window, scatter,
# thresholds.  Any dependence of the result on timing (ring protocol, claims, drops) would show as a pixel difference.
set -u
K="chain or samples or random_scenes"
run() { echo "== $*"; env "$@" timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "$K" 2>&1 | tail -1; }
run RT_VOTE_CTAS=1
run RT_VOTE_CTAS=3 RT_CHAIN_TUNE=1,1,32,0
run RT_CHAIN_TUNE=4,1,4,1
run RT_CHAIN_TUNE=16,1,2,0
run RT_CHAIN_TUNE=64,1,32,1
run RT_CHAIN_TUNE=16,0,32,0
run RT_SPEC_K0=100 RT_SPEC_MARGIN=1000,0
run RT_SPEC_K0=30 RT_SPEC_MARGIN=1000,0
