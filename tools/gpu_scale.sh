#!/bin/bash
# Strong-scaling bench on N GPUs of one box: usage  bash tools/gpu_scale.sh N   (run under `gpurun --gpus N`)
set -u
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err
echo "bench N=$N rc=$?"; tail -1 gpurun_out/bench_n$N.log | cut -c1-300
