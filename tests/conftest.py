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
