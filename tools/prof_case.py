"""One render of a named scene for profiling / timing:
python tools/prof_case.py <scene> <w> <h> <spp> [variant] [passes] [bands]
bands = N emulates rank 0 of an N-GPU tile-mode run (interleaved 8-row bands, RT_PASS_SKIP_MASKED)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opengl_raytracer_b200 import configs, distributed, rtlib
from opengl_raytracer_b200.scene import Scene
name, w, h, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
variant = sys.argv[5] if len(sys.argv) > 5 else "auto"
passes = int(sys.argv[6]) if len(sys.argv) > 6 else 2
bands = int(sys.argv[7]) if len(sys.argv) > 7 else 1
cfg = configs.scene_configs()[name].scaled(w, h, spp)
tris, boxes, idx = Scene(configs.build_objects(cfg.objects)).pack()
params = cfg.params()
if bands > 1:
    params = distributed.band_params(params, 0, bands, 8)
with rtlib.Context(0) as ctx:
    ctx.set_variant(variant); ctx.scene_upload(tris, boxes, idx); ctx.set_params(params); ctx.target_create(w, h)
    for i in range(passes):
        ctx.target_clear(); ctx.timer_start(); ctx.render_pass(0, 0, 0, flags=4 if bands > 1 else 0)
        print(name, variant, f"bands={bands}", "tune=" + os.environ.get("RT_VOTE_TUNE", "-"), "ms", round(ctx.timer_stop(), 3), flush=True)
    print("ran:", ctx.last_pass_variant(), "launches per pass:", ctx.last_pass_launches(), flush=True)
