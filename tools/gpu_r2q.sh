#!/bin/bash
# launch list of the bench command (per-kernel durations; serialised and cold, so shares, not absolutes)
set -u
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu --no-box > gpurun_out/plain_bench_r2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_r2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-box > gpurun_out/ncu_bench_r2.log 2>&1
echo "launch list rc=$?"
