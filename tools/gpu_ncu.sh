#!/bin/bash
# ncu session.  usage: [VARIANTS="pixel pool"] [CASE="cfg4_box 1920 1080 2"] [LIST=1] bash tools/gpu_ncu.sh <tag>
# LIST=1 also records the launch list of the bench command.  Full captures: one launch of the traversal kernel per variant.
set -u
mkdir -p gpurun_out
TAG=${1:-a}
CASE=${CASE:-cfg4_box 1920 1080 2}
NAME=$(echo $CASE | tr ' ' '_')
if [ "${LIST:-0}" = "1" ]; then
  python bench.py --steps 2 --warmup 3 --no-cpu --no-box > gpurun_out/plain_bench_$TAG.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_$TAG.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu --no-box > gpurun_out/ncu_bench_$TAG.log 2>&1
  echo "launch list rc=$?"
fi
for V in ${VARIANTS:-pixel persistent pool}; do
  python tools/prof_case.py $CASE $V 2 > gpurun_out/plain_${V}_$TAG.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'rt(Pixel|Vote|Pool|SpecEval)Kernel' -s 1 -c 1 -f -o gpurun_out/prof_${NAME}_${V}_$TAG \
      python tools/prof_case.py $CASE $V 2 > gpurun_out/ncu_${V}_$TAG.log 2>&1
  echo "ncu $V rc=$?"; tail -1 gpurun_out/ncu_${V}_$TAG.log
done
