"""Per-rank pass times of an N-GPU tile-mode frame for several band heights, measured on ONE GPU rank after rank
(each rank's share is rendered alone, as it would be on its own GPU):
python tools/band_rows_probe.py <scene> <w> <h> <spp> <world> <rows,rows,...>"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opengl_raytracer_b200 import configs, distributed, rtlib
from opengl_raytracer_b200.scene import Scene
name, w, h, spp, world = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
rows_list = [int(x) for x in sys.argv[6].split(",")]
cfg = configs.scene_configs()[name].scaled(w, h, spp)
tris, boxes, idx = Scene(configs.build_objects(cfg.objects)).pack()
with rtlib.Context(0) as ctx:
    ctx.set_variant("auto"); ctx.scene_upload(tris, boxes, idx); ctx.target_create(w, h)
    for rows in rows_list:
        times = []
        for rank in range(world):
            ctx.set_params(distributed.band_params(cfg.params(), rank, world, rows))
            best = 1e9
            for i in range(4):
                ctx.target_clear(); ctx.timer_start(); ctx.render_pass(0, 0, rank, flags=4)
                t = ctx.timer_stop()
                if i: best = min(best, t)
            times.append(best)
        print(f"{name} world={world} band_rows={rows}: max {max(times):.3f} mean {sum(times)/len(times):.3f} min {min(times):.3f} ms  ran {ctx.last_pass_variant()}", flush=True)
