#!/usr/bin/env python
"""bench.py — BASELINE.json's metric (Mrays/s and ms/frame, dragon 3840x2160, 64 spp, 4 bounces) on N B200s.

    python bench.py --gpus 1 --steps K --warmup W             # one process per GPU (torchrun for N > 1)
    python bench.py --impl reference ...                      # the reference's shader math on the host cores

A *step* is one whole frame of the workload: 3840 x 2160 pixels x 64 samples per pixel x up to 5 rays per
path (bounces = 4 -> nBounces = 5), scene models/dragon under open sky as BASELINE.json names it.  A *ray*
is one raycast() (fragment.glsl:246) actually traversed; the primary ray of a pixel-frame is traversed once
and reused by its 64 samples (they share it, fragment.glsl:361-362), so traced rays < reference-equivalent
rays — both are reported, the headline uses traced rays.

N > 1: tile mode — rank g owns the interleaved 8-row bands (y // 8) % N == g of the SAME frame (strong
scaling).  The bands reach rank 0 inside the timed region: by default the render kernels of the other ranks store
their pixels straight into rank 0's target over NVLink (CUDA IPC mapping, RT_PASS_TO_PEER; two 4-byte all-reduces as
fences, "exchange": "peer"); --exchange gather packs them and uses an NCCL gather instead.

JSON keys: "roofline" is the on-SM ceiling that binds the pass (L1 data-pipe wavefronts or warp issue, whichever
fraction is larger; both kept as sub-objects; peaks micro-benchmarked on the same GPU right after the timed region,
"measured_peaks"; counts per launch from the committed ncu capture, time live) — the scene is L2-resident by design, so
the contract's algorithmic-bytes-per-second figure is reported as "algorithmic" without a fraction of the HBM
peak, and "traffic" (DRAM bytes per launch) is the framebuffer.  "phases_ms": CUDA-event split of a step (flush+clear,
pass, and for N > 1 the fence after the pass, or pack / NCCL gather / scatter); "cpu_baseline"; "box_variant" (same frame size with the dragon inside
the reference's closed room: the realistic-throughput figure); "variant": what the library ran under AUTO.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from opengl_raytracer_b200 import configs  # noqa: E402
from opengl_raytracer_b200.scene import Scene  # noqa: E402

WORKLOAD = "cfg4_open"
BOX_WORKLOAD = "cfg4_box"
METRIC = "Mrays/s"
BAND_ROWS = 8


def bench_config(world):
    """The `config` object of the JSON line: what is measured.  The SAME object in both arms (the driver compares them)."""
    cfg = configs.scene_configs()[WORKLOAD]
    return {"workload": f"{WORKLOAD}: models/dragon (37,986 triangles) under open sky, {cfg.width}x{cfg.height}, "
                        f"{cfg.rays_per_pixel} spp, {cfg.bounces} bounces, 1 frame",
            "parallelism": f"b200 arm: tile mode, {world} x interleaved {BAND_ROWS}-row bands, assembled on rank 0 (the render "
                           "kernels store their pixels into rank 0's target over NVLink; 4-byte all-reduces as fences); "
                           "reference arm: rank 0's host cores",
            "l2": "b200 arm: 256 MiB memset between steps, inside the timed region; reference arm: a bounded sample of the "
                  "frame per step (cpu_baseline.sample)",
            "ray_unit": "distinct rays traced: raycast() calls with the primary ray of a pixel-frame counted once, not once per sample"}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def algorithmic_bytes(cnt):
    """64 B per expanded internal node + 16 B per leaf entered + 40 B per triangle test + 68 B per
    hit + 24 B (root bounds) per ray, over the work actually executed (``cnt``: counters of the traced rays)."""
    return 64 * cnt["internal"] + 16 * cnt["leaves"] + 40 * cnt["tri_tests"] + 68 * cnt["hits"] + 24 * cnt["rays"]


class ClockSampler:
    """SM clock / throttle reasons of this rank's GPU, sampled during the timed region.  NVML in-process (a thread,
    started before the warm-up so that NVML's start-up cost stays outside the timed region; `mark()` drops what was
    sampled before it) — forking nvidia-smi next to the timed loop stalled the launching thread for tens of ms."""
    PERIOD = 0.01

    def __init__(self, index):
        self.index, self.rows, self.handle, self.nvml, self.running = index, [], None, None, False

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = str(torch_uuid(self.index))
            try:
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.nvml, self.running = pynvml, True
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception as e:                                     # no NVML: clocks are reported as unknown
            sys.stderr.write(f"clock sampling unavailable: {e}\n")
            self.nvml = None

    def _pump(self):
        n = self.nvml
        while self.running:
            try:
                self.rows.append((n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM),
                                  n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)))
            except Exception:
                pass
            time.sleep(self.PERIOD)

    def mark(self):
        self.rows = []

    def stop(self):
        if self.nvml is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self.running = False
        self.thread.join()
        n = self.nvml
        bits = {"hw_slowdown": n.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": n.nvmlClocksEventReasonHwThermalSlowdown,
                "sw_thermal_slowdown": n.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": n.nvmlClocksEventReasonSwPowerCap}
        sm = [r[0] for r in self.rows]
        reasons = sorted(name for name, bit in bits.items() if any(r[1] & bit for r in self.rows))
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(self.max_mhz),
                "reasons": reasons, "samples": len(sm)}


def torch_uuid(index):
    import torch
    return torch.cuda.get_device_properties(index).uuid


def build_scene(name):
    cfg = configs.scene_configs()[name]
    tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
    return cfg, tris, boxes, indices


# ------------------------------------------------------------------------------------ reference arm
def cpu_sample_setup(stride):
    """The bounded CPU sample both CPU legs time: the SAME frame as the GPU arm (frameNumber 0, 64 samples per
    pixel, 4 bounces), restricted by the shader's own tile mask (fragment.glsl:382) to every `stride`-th
    8-row band.  Returns (scene buffers, uniform block, distinct rays of the sample, description)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    if not os.path.isfile(oracle.PORT_LIB):
        oracle.build_port()
    cfg, tris, boxes, indices = build_scene(WORKLOAD)
    params = cfg.params().replace(tileSizeX=cfg.width, numTilesX=1, tileSizeY=BAND_ROWS, numTilesY=stride, tileY=0,
                                  frameNumber=0)
    # untimed counting pass of the hand-written port: rays = raycast() calls the shader makes; the primary ray of a
    # pixel-frame is the same ray for all its samples, so distinct rays = rays - (spp - 1) per pixel that casts
    _, _, cnt = oracle.render_pass(tris, boxes, indices, params, impl="port", want_counters=True)
    pixels = cnt["paths"] // max(cfg.rays_per_pixel, 1)
    distinct = cnt["rays"] - (cfg.rays_per_pixel - 1) * pixels
    desc = (f"{WORKLOAD}: frame 0 at {cfg.rays_per_pixel} spp on every {stride}th {BAND_ROWS}-row band of the "
            f"{cfg.width}x{cfg.height} frame = {pixels} pixels, {cnt['rays']} raycast() calls, {distinct} distinct rays")
    return oracle, (tris, boxes, indices), params, distinct, cnt["rays"], desc


def run_reference(args):
    """The reference's own shader math on the host cores: oracle/_ref (fragment.glsl compiled as C++) when it
    was built, else the hand-written port; all host threads.  A step is one pass over the bounded sample above.
    Rays are counted like the GPU arm counts them (distinct rays), so value ratios are frame-time ratios."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    oracle, scene, params, distinct, calls, desc = cpu_sample_setup(args.cpu_band_stride)
    kind = "reference" if oracle.have_ref() else "port"
    impl = "ref" if kind == "reference" else "port"
    secs = []
    for step in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        oracle.render_pass(*scene, params, impl=impl)
        dt = time.perf_counter() - t0
        if step >= args.warmup:
            secs.append(dt)
    value = distinct * len(secs) / sum(secs) / 1e6
    cores = oracle.num_threads(impl)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args.gpus),
            "raycast_calls_per_s_M": calls * len(secs) / sum(secs) / 1e6,
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": desc},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print_line(line)


# ------------------------------------------------------------------------------------------ GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist

    from opengl_raytracer_b200 import distributed, rtlib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(vec):
        t = torch.tensor(vec, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(x) for x in t.tolist()]

    ctx = rtlib.Context(local)
    ctx.use_torch_stream()
    ctx.set_variant(args.variant)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    peak_gbs, peak_src, sm_max = load_peaks()

    def measure(name, steps, warmup, with_e2e):
        cfg, tris, boxes, indices = build_scene(name)
        if args.spp:
            cfg = cfg.scaled(cfg.width, cfg.height, args.spp)
        params = cfg.params()
        # pinned host staging for the end-to-end leg
        host = {}
        for key, arr in (("tris", tris), ("boxes", boxes), ("indices", indices)):
            t = torch.empty(arr.nbytes, dtype=torch.uint8).pin_memory()
            t.numpy()[:] = arr.view(np.uint8).reshape(-1)
            host[key] = t
        out_host = torch.empty((cfg.height, cfg.width, 4), dtype=torch.float32).pin_memory()
        n_tris, n_boxes = len(tris), len(boxes)

        def upload():
            ctx._check(ctx.lib.rt_scene_upload(ctx.handle, host["tris"].data_ptr(), n_tris, host["boxes"].data_ptr(),
                                               n_boxes, host["indices"].data_ptr(), indices.size))

        upload()
        ctx.set_params(params)
        ctx.target_create(cfg.width, cfg.height)
        dense = None
        # exchange step of tile mode: peer stores from inside the render kernels (default), or pack / NCCL gather / unpack
        peer = None
        if world > 1 and args.exchange == "peer":
            peer = distributed.open_peer_targets(ctx, dist, rank, world)
            if peer is None and rank == 0:
                sys.stderr.write("bench: peer mapping refused, falling back to the NCCL gather exchange\n")

        def step(frame=0, clear=True):
            nonlocal dense
            if clear:
                ctx.target_clear()
            dense = distributed.render_tile_mode(ctx, dist, params, rank, world, frame, BAND_ROWS, dense, peer=peer)

        # work counters of one step (untimed; reference-equivalent counts, rays_traced = actual traversals)
        ctx.reset_counters()
        ctx.set_params(distributed.band_params(params, rank, world, BAND_ROWS))
        ctx.target_clear()
        ctx.render_pass(0, 0, rank, flags=rtlib.PASS_COUNTERS | rtlib.PASS_SKIP_MASKED)
        mine = ctx.read_counters()
        total = dict(zip(rtlib.COUNTER_NAMES, sum_over_ranks([mine[k] for k in rtlib.COUNTER_NAMES])))
        # work actually executed: the counters are reference-equivalent (the shader re-traces the primary ray for
        # each of its spp samples); a primary-only counting pass gives that ray's work, traced exactly once
        ctx.reset_counters()
        ctx.set_params(distributed.band_params(params.replace(rays_per_pixel=1, nBounces=1), rank, world, BAND_ROWS))
        ctx.target_clear()
        ctx.render_pass(0, 0, rank, flags=rtlib.PASS_COUNTERS | rtlib.PASS_SKIP_MASKED)
        prim = ctx.read_counters()
        extra = max(cfg.rays_per_pixel - 1, 0)
        traced = {k: mine[k] - extra * prim[k] for k in ("rays", "internal", "leaves", "tri_tests", "hits")}
        assert traced["rays"] == mine["rays_traced"], (traced["rays"], mine["rays_traced"])
        local_alg_bytes = algorithmic_bytes(traced)
        ctx.set_params(params)

        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        kernel_events = []

        def timed_step():
            """One step exactly as timed (the warm-up runs the same body, so that lazily loaded kernels, NCCL
            channels and allocator blocks are all in place before the clock starts)."""
            nonlocal dense
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(6 if world > 1 else 3)]
            ev[0].record()
            flush_buf.zero_()                                   # L2 flush between steps (inside the timed region)
            ctx.target_clear()
            ctx.set_params(distributed.band_params(params, rank, world, BAND_ROWS))
            if peer is not None:
                if rank == 0:
                    ctx.target_wait_reads()
                dist.all_reduce(peer.token)                     # fence: rank 0 has cleared the target the peers write
                ev[1].record()
                ctx.render_pass(0, 0, rank, flags=rtlib.PASS_SKIP_MASKED | (rtlib.PASS_TO_PEER if rank else 0))
                ev[2].record()
                dist.all_reduce(peer.token)                     # fence: every rank's kernels and their stores are done
                ev[3].record()
                kernel_events.append(ev[:4])
                return
            ev[1].record()
            ctx.render_pass(0, 0, rank, flags=rtlib.PASS_SKIP_MASKED)
            ev[2].record()
            if world > 1:
                rows = distributed.max_band_rows(cfg.height, BAND_ROWS, world)
                if dense is None:
                    dense = torch.empty((rows, cfg.width, 4), dtype=torch.float32, device=dev)
                ctx.band_gather(rank, dense.data_ptr())
                ev[3].record()
                bufs = distributed.gather_bands(dist, dense, rank, world)
                ev[4].record()
                if rank == 0:
                    for g in range(1, world):
                        ctx.band_scatter(g, bufs[g].data_ptr())
                ev[5].record()
            kernel_events.append(ev)

        for _ in range(warmup):
            timed_step()
        kernel_events.clear()
        barrier()
        sampler.mark()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            timed_step()
        e1.record()
        barrier()
        clocks = sampler.stop() if rank == 0 else None
        ms_total = max_over_ranks(e0.elapsed_time(e1))
        def phase(i):
            return float(np.mean([ev[i].elapsed_time(ev[i + 1]) for ev in kernel_events]))

        kernel_ms = phase(1)
        phases = {"flush_and_clear": phase(0), "pass": kernel_ms}
        if os.environ.get("RT_BENCH_TRACE"):
            for ev in kernel_events:
                sys.stderr.write(f"rank {rank} step phases ms: " + " ".join(f"{a.elapsed_time(b):.3f}" for a, b in zip(ev, ev[1:])) + "\n")
        if peer is not None:
            # phase 0 includes the fence before the pass, "fence_after" the wait for the slowest rank; rank 0's view
            phases.update({"pass_max_over_ranks": max_over_ranks(kernel_ms), "fence_after": phase(2)})
        elif world > 1:
            # where a multi-GPU step goes: slowest rank's pass, then pack / NCCL gather (includes waiting for the
            # slowest rank) / scatter; rank 0's view
            phases.update({"pass_max_over_ranks": max_over_ranks(kernel_ms), "pack": phase(2), "nccl_gather": phase(3),
                           "scatter": phase(4)})
        ms_step = ms_total / steps
        res = {"cfg": cfg, "counters": total, "ms_per_step": ms_step, "kernel_ms": kernel_ms,
               "mrays_s": total["rays_traced"] / (ms_step * 1e-3) / 1e6,
               "ref_equiv_mrays_s": total["rays"] / (ms_step * 1e-3) / 1e6, "clocks": clocks,
               "local_alg_bytes": local_alg_bytes, "launches": steps, "phases_ms": phases,
               "pass_launches": ctx.last_pass_launches(), "variant": ctx.last_pass_variant(),
               "exchange": "none" if world == 1 else "peer" if peer is not None else "gather"}   # traversal-kernel passes; the caller multiplies by kernels per pass
        if with_e2e:
            # end to end through the C ABI with HOST buffers: every step uploads the scene (H2D from pinned memory),
            # sets the uniforms, draws, gathers, and reads the RGBA32F frame back on rank 0 (D2H into pinned memory).
            # The read-back is asynchronous (rt_read_accum_f32_async, two host buffers) and the upload fills the
            # library's second set of scene buffers on its own stream, so both overlap the neighbouring steps' render;
            # the clock stops after the last read-back has landed.  serial = the same loop with a blocking read-back.
            out_hosts = [out_host, torch.empty_like(out_host).pin_memory()] if rank == 0 else [out_host]

            def e2e_loop(n, blocking):
                for i in range(n):
                    upload()
                    step(clear=False)       # frame 0 overwrites every pixel; the targets alternate, so the pass never
                    if rank == 0:           # writes the one whose read-back is still in flight
                        if blocking:
                            ctx._check(ctx.lib.rt_read_accum_f32(ctx.handle, out_hosts[i & 1].data_ptr()))
                        else:
                            ctx.read_accum_f32_async(out_hosts[i & 1].data_ptr())
                ctx.read_wait()

            def timed(blocking):
                e2e_loop(max(1, warmup // 2), blocking)
                barrier()
                t0 = time.perf_counter()
                e2e_loop(steps, blocking)
                barrier()
                return max_over_ranks((time.perf_counter() - t0) * 1e3) / steps

            serial_ms = timed(True)
            e2e_ms = timed(False)
            res["e2e"] = {"value": total["rays_traced"] / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                          "ms_per_step": e2e_ms, "serial_ms_per_step": serial_ms,
                          "h2d_bytes_per_step": int(tris.nbytes + boxes.nbytes + indices.nbytes),
                          "d2h_bytes_per_step": int(out_host.numel() * 4),
                          "note": "copies overlap the neighbouring steps' render (second scene-buffer set + upload "
                                  "stream, asynchronous read-back); serial_ms_per_step = blocking read-back every step"}
            res["e2e_launches"] = steps
        if peer is not None:
            peer.close()            # before rank 0 re-creates (frees) the targets the others have mapped
            barrier()
        res["scene"] = (tris, boxes, indices)
        return res

    main = measure(WORKLOAD, args.steps, args.warmup, True)
    cfg = main["cfg"]
    cnt = main["counters"]
    # + 16 B written per pixel (RGBA32F target; frame 0 reads no previous texel)
    fb_bytes = 16.0 * cfg.width * cfg.height / world
    alg_bytes = main["local_alg_bytes"] + fb_bytes
    achieved = alg_bytes / (main["kernel_ms"] * 1e-3) / 1e9
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
    variant = main["variant"]          # what the library ran (under AUTO the device may have made the choice)
    kernel = {"pixel": "rtPixelKernel", "persistent": "vote::rtVoteKernel", "pool": "pool::rtPoolKernel",
              "wavefront": "wave::rtPrimaryKernel + wave::rtPathKernel",
              "samples": "spec::rtSpecPrimaryKernel + rounds of spec::rtSpecEvalKernel / spec::rtSpecChainKernel",
              "chain": "spec::rtSpecPrimaryKernel + chain::rtChainKernel"}[variant]
    prof = {}
    ppath = os.path.join(ROOT, "profiles", "bench_kernel_counters.json")
    if os.path.isfile(ppath):
        with open(ppath) as f:
            table = json.load(f)
        # a rank's 1/N share runs its own mix of kernels (SAMPLES rounds): its capture is keyed by the share
        prof = table.get(f"{WORKLOAD}/{variant}/share{world}" if world > 1 else f"{WORKLOAD}/{variant}", {})
    per_rank = 1        # the captures are of exactly what rank 0 runs (its 1/N share for N > 1): nothing to divide
    clock_mhz = (main["clocks"] or {}).get("sm_mhz") or sm_max
    # What bounds this path.  The scene (8 MB) is L2/L1-resident by design, DRAM traffic is the framebuffer only, so the
    # HBM roofline does not bind; the ceilings that do are on the SM: the L1 data pipe (one wavefront per 128-byte line
    # per load instruction of a divergent warp) and warp issue.  Both are micro-benchmarked on this GPU right after the
    # timed region (csrc/rt_peaks.cu); the kernel's counts per launch come from the committed ncu capture of this workload
    # (profiles/bench_kernel_counters.json), the time is this run's.  `roofline` reports the one with the larger
    # fraction; the contract's algorithmic-bytes figure is kept as `algorithmic` — a statement of how much scene data the
    # kernels consume per second (it can exceed the HBM peak: the caches serve it), not a bound.
    measured = ctx.measure_peaks() if rank == 0 else {}
    subs = {}
    if prof.get("inst_executed_per_launch"):
        peak_issue = measured.get("issue_gwarpinst_s", 0.0) * 1e9 or sm_count * 4 * clock_mhz * 1e6
        ach = prof["inst_executed_per_launch"] / per_rank / (main["kernel_ms"] * 1e-3)
        subs["issue"] = {"achieved": ach / 1e9, "peak": peak_issue / 1e9, "unit": "G warp-inst/s", "frac": ach / peak_issue,
                         "lanes_per_inst": prof.get("lanes_per_inst"), "peak_nominal": sm_count * 4 * clock_mhz * 1e-3,
                         "source": "inst per launch from the committed ncu capture of this workload, time live; "
                                   "peak = FFMA issue micro-benchmark on this GPU (nominal: 4 schedulers x SMs x clock)"}
    if prof.get("l1_wavefronts_per_launch"):
        peak_wf = measured.get("l1_gwavefronts_s", 0.0) * 1e9 or sm_count * clock_mhz * 1e6
        ach = prof["l1_wavefronts_per_launch"] / per_rank / (main["kernel_ms"] * 1e-3)
        subs["l1_wavefronts"] = {"achieved": ach / 1e9, "peak": peak_wf / 1e9, "unit": "G wavefronts/s", "frac": ach / peak_wf,
                                 "peak_nominal": sm_count * clock_mhz * 1e-3,
                                 "source": "l1tex data-pipe wavefronts per launch from the committed ncu capture, time live; "
                                           "peak = divergent-load micro-benchmark on this GPU (nominal: 1 wavefront / cycle / SM)"}
    algorithmic = {"achieved": achieved, "unit": "GB/s", "hbm_peak": peak_gbs, "hbm_peak_source": peak_src,
                   "ratio_to_hbm_peak": achieved / peak_gbs, "bytes_per_launch": alg_bytes,
                   "note": "SURVEY 8d: 64 B/internal node, 16 B/leaf, 40 B/triangle test, 68 B/hit, 24 B/ray of the rays "
                           "actually traced + 16 B/pixel of framebuffer, per second of this rank's pass; cache-resident data, "
                           "so not a fraction of a bound"}
    if measured.get("l2_gbs"):
        # the scene data comes from L1/L2: L2 -> SM bandwidth is the memory ceiling the algorithmic bytes can be held against
        algorithmic["l2_peak"] = measured["l2_gbs"]
        algorithmic["ratio_to_l2_peak"] = achieved / measured["l2_gbs"]
    traffic = (prof["dram_bytes_per_launch"] / per_rank) if prof.get("dram_bytes_per_launch") else None
    if subs:
        name = max(subs, key=lambda k: subs[k]["frac"])
        roofline = {"bound": name, "achieved": subs[name]["achieved"], "peak": subs[name]["peak"], "unit": subs[name]["unit"],
                    "frac": subs[name]["frac"], "peak_source": "micro-benchmarked on this GPU in this run (csrc/rt_peaks.cu)"}
    else:
        # no committed capture of this variant: only the algorithmic figure is known
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                    "peak_source": peak_src, "cache_resident": True}
    roofline.update({"traffic": traffic, "kernel": kernel, "kernel_ms": main["kernel_ms"], "algorithmic": algorithmic,
                     "measured_peaks": measured, **subs})
    line = {"metric": METRIC, "value": main["mrays_s"], "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(world), "variant": variant,
            "ms_per_frame": main["ms_per_step"], "rays_per_step": cnt["rays_traced"],
            "reference_equivalent": {"rays_per_step": cnt["rays"], "mrays_s": main["ref_equiv_mrays_s"]},
            "e2e": main["e2e"], "roofline": roofline, "clocks": main["clocks"], "phases_ms": main["phases_ms"],
            # this repo's kernels launched by rank 0 inside the timed region: the pass (1 kernel; the samples variant
            # launches its primary stage + rounds of sort/evaluate/chain kernels: the library reports the count) + band
            # gather (1) and scatters (N - 1)
            "exchange": main["exchange"],
            "gpu_launches": main["launches"] * (main["pass_launches"] + (world if main["exchange"] == "gather" else 0))}

    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args)
    if not args.no_box:
        box = measure(BOX_WORKLOAD, max(1, min(5, args.steps)), 3, False)
        bc = box["counters"]
        line["box_variant"] = {"workload": f"{BOX_WORKLOAD}: dragon inside the main.py room, same size/spp/bounces",
                               "variant": box["variant"],
                               "value": box["mrays_s"], "unit": "Mrays/s", "ms_per_frame": box["ms_per_step"],
                               "rays_per_step": bc["rays_traced"], "kernel_ms": box["kernel_ms"],
                               "algorithmic_gbs": (box["local_alg_bytes"] + fb_bytes) / (box["kernel_ms"] * 1e-3) / 1e9,
                               "per_ray": {"internal": bc["internal"] / bc["rays"], "tri_tests": bc["tri_tests"] / bc["rays"]}}
    if rank == 0:
        print_line(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline(args):
    """The reference shader (oracle/_ref, else the port) on the host cores, on the same bounded sample the
    --impl reference arm times, repeated until about `cpu_seconds` of CPU work."""
    oracle, scene, params, distinct, calls, desc = cpu_sample_setup(args.cpu_band_stride)
    kind = "reference" if oracle.have_ref() else "port"
    impl = "ref" if kind == "reference" else "port"
    secs, reps = 0.0, 0
    while secs < args.cpu_seconds and reps < 16:
        t0 = time.perf_counter()
        oracle.render_pass(*scene, params, impl=impl)
        secs += time.perf_counter() - t0
        reps += 1
    return {"value": distinct * reps / secs / 1e6, "unit": "Mrays/s", "cores": oracle.num_threads(impl), "kind": kind,
            "raycast_calls_per_s_M": calls * reps / secs / 1e6,
            "sample": f"{reps} x [{desc}], {secs:.1f} s"}


def main():
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner) go to stderr
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    global print_line
    def print_line(obj):
        real_stdout.write(json.dumps(obj) + "\n")
        real_stdout.flush()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--variant", default="auto", choices=["auto", "pixel", "persistent", "wavefront", "pool", "samples", "chain"])
    ap.add_argument("--exchange", default="peer", choices=["peer", "gather"],
                    help="N > 1: how the bands reach rank 0 (peer = stores from inside the render kernels over NVLink)")
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel (not a valid bench number)")
    ap.add_argument("--no-box", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--cpu-band-stride", type=int, default=8)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
