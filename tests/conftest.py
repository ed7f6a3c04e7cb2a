import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

REFERENCE_ROOT = "/root/reference"
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA GPU (run on the B200 box with -m gpu)")


def have_reference():
    return os.path.isdir(REFERENCE_ROOT)


def have_models():
    return os.path.isfile(os.path.join(ROOT, "models", "sphere", "sphere.obj"))


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Oracle (always), staged assets and the reference-shader build (when /root/reference is mounted)."""
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "-s"], check=True)
    if have_reference():
        import stage_models
        stage_models.stage(quiet=True)
        import build_ref
        build_ref.build(quiet=True)


@pytest.fixture(scope="session")
def oracle_mod(_built):
    import oracle
    return oracle


@pytest.fixture(scope="session")
def scene_buffers():
    """name -> (tris, boxes, indices, SceneConfig), built once per session with this package's host code."""
    from opengl_raytracer_b200 import configs
    from opengl_raytracer_b200.scene import Scene
    cache = {}
    cfgs = configs.scene_configs()

    def get(name):
        if name not in cache:
            cfg = cfgs[name]
            if any(o["kind"] == "Mesh" for o in cfg.objects) and not have_models():
                pytest.skip("staged models/ not present (run tools/stage_models.py where /root/reference is mounted)")
            cache[name] = Scene(configs.build_objects(cfg.objects)).pack() + (cfg,)
        return cache[name]
    return get


def load_golden(name):
    from opengl_raytracer_b200.configs import RenderParams
    blob = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    kw = {}
    for key in blob.files:
        if key.startswith("p_"):
            v = blob[key]
            kw[key[2:]] = tuple(float(x) for x in v) if v.ndim == 1 else v.item()
    params = RenderParams(**kw)
    return dict(tris=blob["tris"], boxes=blob["boxes"], indices=blob["indices"], frames=int(blob["frames"]),
                expected=blob["expected"], params=params)


GOLDEN_CASES = ["rects_box_2f", "rects_box_nolambert", "rects_box_tiles", "cfg1_open", "knight_box"]


def random_scene(seed):
    """A seeded random scene of Rect boxes (no OBJ assets): 4-24 boxes of random size / pose / material — thin slabs,
    mirrors, emitters — under open sky or inside the main.py room, with random uniforms (bounces 0-6, 1-8 samples,
    lambertian on/off, sky brightness, jitter, ragged frame size, tiling).  Returns (tris, boxes, indices, params)."""
    from opengl_raytracer_b200 import configs
    from opengl_raytracer_b200.scene import Scene
    rng = np.random.default_rng(1000 + seed)
    specs = []
    for _ in range(int(rng.integers(4, 25))):
        size = rng.uniform(0.5, 6.0, 3)
        if rng.random() < 0.3:
            size[int(rng.integers(0, 3))] = 0.05                      # a slab
        emissive = rng.random() < 0.25
        specs.append(configs.rect(
            [float(x) for x in size], [float(x) for x in rng.uniform(-20, 20, 3)],
            [float(x) for x in rng.uniform(0, 360, 3)], [float(x) for x in rng.uniform(0, 1, 3)],
            emission_color=[float(x) for x in rng.uniform(0, 1, 3)] if emissive else (0, 0, 0),
            emission=float(rng.uniform(0.5, 4.0)) if emissive else 0,
            roughness=float(rng.choice([0.0, 0.15, 0.5, 1.0, rng.random()])), scale=int(rng.integers(1, 4))))
    if rng.random() < 0.5:
        specs += configs.box_room()
    tris, boxes, indices = Scene(configs.build_objects(specs)).pack()
    params = configs.make_params(
        int(rng.integers(37, 121)), int(rng.integers(23, 71)), int(rng.integers(0, 7)), int(rng.integers(1, 9)),
        lambertian=bool(rng.random() < 0.5), sky_brightness=float(rng.choice([0.0, 0.5, 1.0, 2.0])),
        jitter_amount=float(rng.choice([0.0, 0.001, 0.01])), tile_size=int(rng.choice([1, 1, 2, 3])),
        **configs.BOX_CAMERA)
    return tris, boxes, indices, params
