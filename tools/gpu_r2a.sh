#!/bin/bash
# round-2 call A: LDG.256 probe, baseline per-line capture of the vote kernel
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 120 tools/native/ldg256_probe > gpurun_out/r2a_ldg256.txt 2>&1; echo "probe rc=$?"; cat gpurun_out/r2a_ldg256.txt
python tools/prof_case.py cfg4_open 3840 2160 64 persistent 3 > gpurun_out/r2a_plain_open.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'rtVoteKernel' -s 1 -c 1 -f -o gpurun_out/r2a_vote_open \
   python tools/prof_case.py cfg4_open 3840 2160 64 persistent 2 > gpurun_out/r2a_ncu_open.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/r2a_plain_open.log
python tools/prof_case.py cfg4_box 3840 2160 8 persistent 3
python tools/prof_case.py cfg4_box 3840 2160 8 pixel 3
python tools/prof_case.py cfg4_open 3840 2160 64 samples 3
python tools/prof_case.py cfg4_open 3840 2160 64 samples 3 8
python tools/prof_case.py cfg4_open 3840 2160 64 persistent 3 8
