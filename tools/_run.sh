N=${1:-8}
for S in cfg5_box cfg5_open; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/bench_frames.py --scene $S --frames 1024 --repeats 2 2>gpurun_out/frames_$S_n$N.err | tail -1 | tee gpurun_out/frames_${S}_n$N.json
done
