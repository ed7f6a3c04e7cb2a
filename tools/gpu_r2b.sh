#!/bin/bash
# round-2 call B: parity suite on the LDG.256 build + variant timings
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2b_pytest.log
python tools/prof_case.py cfg4_open 3840 2160 64 persistent 3
python tools/prof_case.py cfg4_box 3840 2160 8 persistent 3
python tools/prof_case.py cfg4_box 3840 2160 8 pixel 3
python tools/prof_case.py cfg4_open 3840 2160 64 samples 3
python tools/prof_case.py cfg4_open 3840 2160 64 samples 3 8
python tools/prof_case.py cfg4_open 3840 2160 64 persistent 3 8
