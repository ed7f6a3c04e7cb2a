#!/bin/bash
# 2..N GPUs: NCCL leg of the peer test, then the bench with both exchanges
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_peer.py -m gpu -x -q -k nccl 2>&1 | tail -5
for X in peer gather; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 10 --no-box --exchange $X > gpurun_out/r2o_n${N}_$X.json 2> gpurun_out/r2o_n${N}_$X.err
  echo "bench N=$N $X rc=$?"; tail -3 gpurun_out/r2o_n${N}_$X.err
  python - <<P
import json
d=json.loads(open("gpurun_out/r2o_n${N}_$X.json").read().strip().splitlines()[-1])
print(d["n_gpus"], d["exchange"], d["variant"], d["ms_per_step"], d["value"], d["phases_ms"], d["e2e"]["ms_per_step"], d["gpu_launches"])
P
done
