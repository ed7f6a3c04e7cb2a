"""First-light GPU probe: smoke(), then parity + timing on a few scenes.  Scratch tool, not a test."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import __graft_entry__ as g
g.smoke()
import oracle
from opengl_raytracer_b200 import configs, rtlib
from opengl_raytracer_b200.scene import Scene
cfgs = configs.scene_configs()
out = {}
with rtlib.Context(0) as ctx:
    for name, (w, h, spp, check) in {"cfg1_open": (640, 360, 1, True), "cfg2_box": (320, 180, 2, True),
                                     "cfg4_open": (3840, 2160, 8, False), "cfg4_box": (3840, 2160, 8, False)}.items():
        c = cfgs[name].scaled(w, h, spp)
        tris, boxes, idx = Scene(configs.build_objects(c.objects)).pack()
        p = c.params()
        ctx.scene_upload(tris, boxes, idx); ctx.set_params(p); ctx.target_create(w, h)
        for variant in ("pixel", "persistent"):
            ctx.set_variant(variant); ctx.target_clear(); ctx.reset_counters()
            ctx.render_pass(0, 0, 0, flags=rtlib.PASS_COUNTERS | rtlib.PASS_PRIMARY_IDS)
            cnt = ctx.read_counters(); got = ctx.read_accum_f32(); ids = ctx.read_primary_ids()
            rec = {"counters": cnt}
            if check:
                ref, rid, rc = oracle.render_pass(tris, boxes, idx, p, want_ids=True, want_counters=True)
                rec["bit_exact"] = got.tobytes() == ref.tobytes(); rec["ids_equal"] = bool(np.array_equal(ids, rid))
                rec["px_diff"] = int((got != ref).any(-1).sum())
                rec["counters_equal"] = all(cnt[k] == rc[k] for k in cnt if k != "rays_traced")
            ts = []
            for rep in range(3):
                ctx.target_clear(); ctx.timer_start(); ctx.render_pass(0, 0, 0); ts.append(ctx.timer_stop())
            rec["ms"] = ts; rec["mrays_s"] = cnt["rays_traced"] / (min(ts) * 1e-3) / 1e6
            out[f"{name}/{variant}"] = rec
            print(name, variant, {k: v for k, v in rec.items() if k != "counters"}, "rays_traced", cnt["rays_traced"], flush=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w"), indent=1)
