#!/bin/bash
# One GPU-box session: parity tests, probe, bench, ncu launch list, ncu full captures.  Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python tools/gpu_probe.py > gpurun_out/probe.log 2>&1; echo "probe rc=$?"; tail -12 gpurun_out/probe.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.log
python bench.py --steps 3 --warmup 3 --variant persistent --no-cpu > gpurun_out/bench_persistent.log 2>gpurun_out/bench_persistent.err; echo "bench-p rc=$?"; cat gpurun_out/bench_persistent.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?"; cat gpurun_out/bench_ref.log
