"""Frame mode: progressive accumulation of `frames` 1-spp frames at 3840x2160, the frames dealt
round-robin to the ranks (rt_render_frames_sum), NCCL reduce of the sums, rt_finish_sum on rank 0.

  python tools/bench_frames.py [--scene cfg5_box] [--frames 256] [--check]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_frames.py ...

Prints one JSON line on rank 0: ms for the whole job (CUDA events, max over ranks, inputs resident), Mrays/s over the
distinct rays of all frames, and with --check (N=1 only, small frame counts) the largest difference between the summed
result and the reference's sequential running mean (float re-association only)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from opengl_raytracer_b200 import configs, distributed, rtlib  # noqa: E402
from opengl_raytracer_b200.scene import Scene  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="cfg5_box")
    ap.add_argument("--frames", type=int, default=256)
    ap.add_argument("--repeats", type=int, default=3)
    ap.add_argument("--warmups", type=int, default=1)
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = configs.scene_configs()[args.scene]
    tris, boxes, idx = Scene(configs.build_objects(cfg.objects)).pack()
    params = cfg.params()
    ctx = rtlib.Context(local)
    ctx.use_torch_stream()
    ctx.scene_upload(tris, boxes, idx)
    ctx.set_params(params)
    ctx.target_create(cfg.width, cfg.height)

    # distinct rays of this rank's frames (one counting pass per distinct frame would be the exact figure; frames differ
    # only by their seeds, so one frame's count x frames is exact to < 0.1 %): counted on frame 0
    ctx.reset_counters()
    ctx.target_clear()
    ctx.render_pass(0, 0, 0, flags=rtlib.PASS_COUNTERS)
    rays_per_frame = ctx.read_counters()["rays_traced"]

    def job():
        ctx.target_clear()
        distributed.render_frame_mode(ctx, dist, params, rank, world, 0, args.frames, lambda c: c.accum_tensor())

    for _ in range(args.warmups):
        job()
    times = []
    for _ in range(args.repeats):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        job()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    ms = float(np.median(times))
    line = {"workload": f"{args.scene}: {cfg.width}x{cfg.height}, 1 spp, {cfg.bounces} bounces, frames 0..{args.frames - 1} "
                        "accumulated (frame mode: frames round-robin over the ranks, NCCL reduce of the sums)",
            "n_gpus": world, "frames": args.frames, "ms_total": ms, "ms_per_frame": ms / args.frames,
            "mrays_s": rays_per_frame * args.frames / (ms * 1e-3) / 1e6, "rays_per_frame": rays_per_frame,
            "launches_per_frame": ctx.last_pass_launches(), "times_ms": times}
    if args.check and world == 1:
        summed = ctx.read_accum_f32()
        ctx.target_clear()
        ctx.render_frames(0, args.frames)                      # the reference's sequential running mean
        mean = ctx.read_accum_f32()
        line["max_abs_diff_vs_running_mean"] = float(np.abs(summed - mean).max())
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
