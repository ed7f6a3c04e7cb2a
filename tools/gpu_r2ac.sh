#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "samples or auto or random or launch" 2>&1 | tail -2
for B in 1 2 4 8; do python tools/prof_case.py cfg4_open 3840 2160 64 samples 3 $B | tail -2 | head -1; done
python tools/prof_case.py cfg4_open 3840 2160 64 auto 3 1 | tail -2 | head -1
