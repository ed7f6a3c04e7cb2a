#!/bin/bash
# ncu session: launch list of the bench command + full captures of the traversal kernels (box scene, 1080p, 2 spp)
set -u
mkdir -p gpurun_out
TAG=${1:-a}
python bench.py --steps 2 --warmup 3 --no-cpu --no-box > gpurun_out/plain_bench_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-box > gpurun_out/ncu_bench_$TAG.log 2>&1
echo "launch list rc=$?"
for V in ${VARIANTS:-pixel persistent}; do
  python tools/prof_case.py cfg4_box 1920 1080 2 $V 2 > gpurun_out/plain_${V}_$TAG.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'rt(Pixel|Vote|Pool)Kernel' -s 1 -c 1 -f -o gpurun_out/prof_box_${V}_$TAG \
      python tools/prof_case.py cfg4_box 1920 1080 2 $V 2 > gpurun_out/ncu_${V}_$TAG.log 2>&1
  echo "ncu $V rc=$?"; tail -2 gpurun_out/ncu_${V}_$TAG.log
done
