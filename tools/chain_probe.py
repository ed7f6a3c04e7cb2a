"""Latency probe: one 8x4-pixel tile of the 4K/64-spp open-dragon frame rendered alone (empty machine) —
the serial sample chain of a pixel-frame; bounds how far one frame can be strong-scaled."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from opengl_raytracer_b200 import configs, rtlib
from opengl_raytracer_b200.scene import Scene
cfg = configs.scene_configs()["cfg4_open"]
tris, boxes, idx = Scene(configs.build_objects(cfg.objects)).pack()
w, h = 3840, 2160
with rtlib.Context(0) as ctx:
    ctx.scene_upload(tris, boxes, idx); ctx.target_create(w, h)
    # find tiles on the dragon: primary ids at 1 spp
    ctx.set_params(cfg.params(rays_per_pixel=1)); ctx.set_variant("pixel")
    ctx.render_pass(0, 0, 0, flags=rtlib.PASS_PRIMARY_IDS | rtlib.PASS_COUNTERS)
    ids = ctx.read_primary_ids()
    hit = (ids >= 0)
    print("hit pixels", int(hit.sum()), "of", w * h)
    tiles = hit.reshape(h // 4, 4, w // 8, 8).all(axis=(1, 3))
    ty, tx = np.nonzero(tiles)
    print("full tiles on the mesh:", len(ty))
    rng = np.random.default_rng(1)
    pick = rng.choice(len(ty), 12, replace=False)
    for variant in ("pixel", "persistent"):
        ctx.set_variant(variant)
        out = []
        for k in pick:
            p = cfg.params().replace(tileSizeX=8, tileSizeY=4, numTilesX=w // 8, numTilesY=h // 4, tileX=int(tx[k]), tileY=int(ty[k]))
            ctx.set_params(p)
            ctx.reset_counters(); ctx.render_pass(0, int(tx[k]), int(ty[k]), flags=rtlib.PASS_SKIP_MASKED | rtlib.PASS_COUNTERS)
            rays = ctx.read_counters()["rays_traced"]
            ctx.timer_start(); ctx.render_pass(0, int(tx[k]), int(ty[k]), flags=rtlib.PASS_SKIP_MASKED); ms = ctx.timer_stop()
            out.append((round(ms, 3), rays))
        print(variant, "ms per lone 32-pixel tile, (ms, rays):", out)
