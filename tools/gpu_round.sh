#!/bin/bash
# One GPU-box session: parity tests, bench (both arms), ncu launch list + full captures of the bench kernels.
set -u
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.log
python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.log; tail -3 gpurun_out/bench.err
if [ "${NCU:-1}" = "1" ]; then
  LIST=1 VARIANTS="persistent samples" CASE="cfg4_open 3840 2160 64" bash tools/gpu_ncu.sh ${TAG:-r1}
  VARIANTS=pixel CASE="cfg4_box 3840 2160 8" bash tools/gpu_ncu.sh ${TAG:-r1}
fi
