"""CPU-side checks of the two drivers above the C ABI: bench.py's reference arm (the JSON contract of its one output
line) and the headless App's loop sequencing (main.py:376-430) against a recording stand-in for the ctypes context."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    run = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--cpu-band-stride", "64"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert run.returncode == 0, run.stderr[-2000:]
    lines = [ln for ln in run.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, run.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Mrays/s" and d["unit"] == "Mrays/s"
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["higher_is_better"] is True and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] and "64 spp" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "cfg4_open" in d["config"]["workload"]
    # the driver compares the two arms' `config` objects: both come from one function
    sys.path.insert(0, ROOT)
    import bench
    assert d["config"] == bench.bench_config(1) and set(d["config"]) == {"workload", "parallelism", "l2", "ray_unit"}


class _RecordingContext:
    """Stands in for rtlib.Context: records the calls the App makes (no GPU, no library)."""
    def __init__(self, device=0):
        self.calls, self.width, self.height = [], 0, 0

    def __getattr__(self, name):
        def call(*a, **k):
            self.calls.append((name, a))
            if name == "target_create":
                self.width, self.height = a[0], a[1]
            if name == "timer_stop":
                return 1.0
            if name == "read_accum_f32":
                return np.zeros((self.height, self.width, 4), np.float32)
            return None
        return call


def test_app_loop_sequencing(monkeypatch):
    from opengl_raytracer_b200 import app as app_mod, configs, rtlib
    monkeypatch.setattr(rtlib, "Context", _RecordingContext)
    objects = configs.build_objects(configs.scene_configs()["rects_box"].objects)
    a = app_mod.App((90, 60), bounces=2, rays_per_pixel=1, tileSize=3, objects=objects)
    tiles = a.params.numTilesX * a.params.numTilesY
    assert tiles > 1 and a.screen.frame_count == 0
    a.main(frames=2)
    passes = [c[1] for c in a.context.calls if c[0] == "render_pass"]
    assert len(passes) == 2 * tiles and a.screen.frame_count == 2 and a.total_frames == 2 * tiles
    # one tile per iteration, x fastest, the frame number steps after the last tile (main.py:409-430)
    want = [(f, tx, ty) for f in range(2) for ty in range(a.params.numTilesY) for tx in range(a.params.numTilesX)]
    assert [tuple(p[:3]) for p in passes] == want
    # a camera change re-derives the uniforms and resets the accumulation (main.py:326-337, 252-271)
    before = len([c for c in a.context.calls if c[0] == "set_params"])
    a.step()
    a.set_camera(cam_pos=(0.0, 1.0, -5.0))
    assert a.screen.frame_count == 0 and (a.tileX, a.tileY) == (0, 0)
    assert len([c for c in a.context.calls if c[0] == "set_params"]) == before + 1
    assert a.context.calls[-1][0] in ("target_clear", "set_accum_index")
    a.set_lambertian(False)
    assert a.params.lambertian == 0
