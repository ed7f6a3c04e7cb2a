#!/bin/bash
# quick GPU check: parity subset + timings of the variants on the two 4K workloads
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -k "${PYTEST_K:-persistent}" > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/quick_pytest.log
for V in ${VARIANTS:-pixel persistent}; do
  python tools/prof_case.py cfg4_box 3840 2160 8 $V 2
  python tools/prof_case.py cfg4_open 3840 2160 64 $V 2
done
