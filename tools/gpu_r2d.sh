#!/bin/bash
# round-2 call D: ncu of the chain kernel (whole frame)
set -u
mkdir -p gpurun_out
python tools/prof_case.py cfg4_open 3840 2160 64 chain 3 > gpurun_out/r2d_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'rtChainKernel' -s 1 -c 1 -f -o gpurun_out/r2d_chain_open \
   python tools/prof_case.py cfg4_open 3840 2160 64 chain 2 > gpurun_out/r2d_ncu.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2d_plain.log
