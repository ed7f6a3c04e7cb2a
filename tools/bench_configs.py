"""Every single-frame configuration of SURVEY §8(d) (cfg1-cfg4, open and box variants) on one GPU:

  python tools/bench_configs.py [--cpu] > gpurun_out/configs.jsonl

One JSON line per configuration: ms per frame (frame 0, CUDA events, mean of the timed passes after warm-up, scene
resident, variant AUTO), distinct rays traced, Mrays/s, work per ray from the library's exact counters; with --cpu also
the oracle's time for the same frame on the host cores when the frame is small enough (cfg1, cfg2), as SURVEY §8(d) asks."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from opengl_raytracer_b200 import configs, rtlib  # noqa: E402
from opengl_raytracer_b200.scene import Scene  # noqa: E402

NAMES = ["cfg1_open", "cfg2_open", "cfg2_box", "cfg3_open", "cfg3_box", "cfg4_open", "cfg4_box"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--passes", type=int, default=5)
    args = ap.parse_args()
    ctx = rtlib.Context(0)
    for name in NAMES:
        cfg = configs.scene_configs()[name]
        tris, boxes, idx = Scene(configs.build_objects(cfg.objects)).pack()
        params = cfg.params()
        ctx.set_variant("auto")
        ctx.scene_upload(tris, boxes, idx)
        ctx.set_params(params)
        ctx.target_create(cfg.width, cfg.height)
        ctx.reset_counters()
        ctx.target_clear()
        ctx.render_pass(0, 0, 0, flags=rtlib.PASS_COUNTERS)
        cnt = ctx.read_counters()
        passes = args.passes if cfg.width * cfg.height * cfg.rays_per_pixel < 3e8 or "open" in name else 2
        times = []
        for i in range(passes + 2):
            ctx.target_clear()
            ctx.timer_start()
            ctx.render_pass(0, 0, 0)
            ms = ctx.timer_stop()
            if i >= 2:
                times.append(ms)
        ms = sum(times) / len(times)
        rays = cnt["rays_traced"]
        line = {"config": name, "note": cfg.note, "size": [cfg.width, cfg.height], "spp": cfg.rays_per_pixel,
                "bounces": cfg.bounces, "triangles": int(len(tris)), "nodes": int(len(boxes)),
                "kernels_per_pass": ctx.last_pass_launches(), "ms_per_frame": ms, "rays_traced": rays,
                "mrays_s": rays / (ms * 1e-3) / 1e6, "reference_equivalent_rays": cnt["rays"],
                "per_ray": {"internal": cnt["internal"] / max(cnt["rays"], 1), "tri_tests": cnt["tri_tests"] / max(cnt["rays"], 1),
                            "hit_fraction": cnt["hits"] / max(cnt["rays"], 1)}}
        if args.cpu and cfg.width * cfg.height * cfg.rays_per_pixel <= 1280 * 720 * 8:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import oracle
            t0 = time.perf_counter()
            oracle.render_pass(tris, boxes, idx, params.replace(frameNumber=0), impl="ref" if os.path.isfile(oracle.REF_LIB) else "port")
            sec = time.perf_counter() - t0
            line["cpu"] = {"seconds": sec, "mrays_s": rays / sec / 1e6, "cores": os.cpu_count(),
                           "impl": "oracle/_ref" if os.path.isfile(oracle.REF_LIB) else "oracle port"}
        print(json.dumps(line), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
