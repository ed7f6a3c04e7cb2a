#!/bin/bash
# ncu --set full of ONE whole SAMPLES pass (stage 1 + all rounds) at the 1/N share: usage bash tools/gpu_r2u.sh N
set -u
N=${1:-8}
mkdir -p gpurun_out
python tools/prof_case.py cfg4_open 3840 2160 64 samples 2 $N > gpurun_out/plain_samples_share$N.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rtSpec -s 12 -c 12 -f -o gpurun_out/r2_samples_share$N \
    python tools/prof_case.py cfg4_open 3840 2160 64 samples 2 $N > gpurun_out/ncu_samples_share$N.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/plain_samples_share$N.log
