#!/bin/bash
# ncu --set full of the two kernels of one bench pass (AUTO -> CHAIN) at HEAD
set -u
mkdir -p gpurun_out
python tools/prof_case.py cfg4_open 3840 2160 64 auto 2 > gpurun_out/r2ak_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'rt(SpecPrimary|Chain)Kernel' -s 2 -c 2 -f -o gpurun_out/r2_chain_open_dfs \
   python tools/prof_case.py cfg4_open 3840 2160 64 auto 2 > gpurun_out/r2ak_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r2ak_plain.log
