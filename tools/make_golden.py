"""Generate tests/golden/*.npz: small scenes (the packed SSBO payloads themselves, so the fixtures do not
depend on BLAS bit patterns of the machine that replays them), the uniform block, and the RGBA32F frames
produced by the REFERENCE's own fragment.glsl compiled as C++ (oracle/_ref/libfragment_ref.so).

Run here (needs /root/reference for oracle/_ref and the staged models):  python tools/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import build_ref   # noqa: E402
import oracle      # noqa: E402
from opengl_raytracer_b200 import configs   # noqa: E402
from opengl_raytracer_b200.scene import Scene   # noqa: E402

PARAM_FIELDS = [f for f in configs.RenderParams.__dataclass_fields__]

CASES = {
    # name: (scene, width, height, spp, frames, overrides)
    "rects_box_2f": ("rects_box", 96, 54, 2, 2, {}),
    "rects_box_nolambert": ("rects_box", 64, 36, 1, 1, {"lambertian": 0}),
    "rects_box_tiles": ("rects_box", 64, 36, 1, 2, {"tile_size": 2}),
    "cfg1_open": ("cfg1_open", 128, 72, 1, 1, {}),
    "knight_box": (None, 96, 54, 2, 1, {}),
}


def knight_box_config():
    objs = [configs.mesh([10, -24.9, -15], [270, 0, 0], "knight", [0.8, 0.6, 0.2], roughness=0.5, scale=4)] + configs.box_room()
    return configs.SceneConfig("knight_box", objs, 96, 54, 4, 2, **configs.BOX_CAMERA)


def main():
    assert build_ref.build(quiet=True), "oracle/_ref needs /root/reference"
    cfgs = configs.scene_configs()
    out_dir = os.path.join(ROOT, "tests", "golden")
    for name, (scene, w, h, spp, frames, over) in CASES.items():
        cfg = (cfgs[scene] if scene else knight_box_config()).scaled(w, h, spp)
        tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
        over = dict(over)
        lamb = over.pop("lambertian", 1)
        params = cfg.params(lambertian=bool(lamb), **over)
        acc = oracle.render_frames(tris, boxes, indices, params, 0, frames, impl="ref")
        blob = {k: np.asarray(getattr(params, k)) for k in PARAM_FIELDS}
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), tris=tris.view(np.uint8), boxes=boxes.view(np.uint8),
                            indices=indices, frames=np.int32(frames), expected=acc, **{"p_" + k: v for k, v in blob.items()})
        print(name, acc.shape, float(acc[..., :3].mean()))


if __name__ == "__main__":
    main()
