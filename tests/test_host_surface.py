"""CPU tests of the kept host surface: OBJ loader, Mesh/Rect, Scene packing, BVH builder, camera/uniforms.
Where /root/reference is mounted the packed buffers are compared byte for byte with what the reference's
own object.py / scene.py / *.pyx upload (tools/ref_host_dump.py runs them under a stub OpenGL.GL)."""
import json
import math
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import REFERENCE_ROOT, ROOT, have_reference
from opengl_raytracer_b200 import configs
from opengl_raytracer_b200.loadObject import loadObj
from opengl_raytracer_b200.object import Rect
from opengl_raytracer_b200.scene import Scene, boundingBoxStruct, triangleStruct


def _write(tmp_path, text):
    p = tmp_path / "m.obj"
    p.write_text(text)
    return str(p)


def test_loader_face_forms_and_fan(tmp_path):
    # reference loadObject.pyx:53-111: p/t/n, p//n, p/t/, p/t and p corners; quads become fans
    obj = ("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvt 0.25 0.75\nvn 0 1 0\n"
           "f 1/1/1 2//1 3/1/ 4/1\nf 1 2 3\n")
    v = np.array(loadObj(_write(tmp_path, obj))).reshape(-1, 8)
    assert v.shape == (9, 8)                                   # quad -> 2 triangles, + 1 triangle
    assert v[:, :3].tolist()[:6] == [[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 0, 0], [1, 1, 0], [0, 1, 0]]
    assert v[0, 3:6].tolist() == [0, 1, 0] and v[0, 6:].tolist() == [0.25, 0.25]     # v flipped (:109)
    assert v[1, 3:6].tolist() == [0, 1, 0] and v[1, 6:].tolist() == [0, 1]           # p//n: uv (0,0) -> (0,1)
    assert v[2, 3:6].tolist() == [0, 0, 1]                                            # p/t/ : default normal
    assert v[5, 3:6].tolist() == [0, 0, 1] and v[5, 6:].tolist() == [0.25, 0.25]     # p/t
    assert v[6, 3:6].tolist() == [0, 0, 1] and v[6, 6:].tolist() == [0, 1]           # p


def test_loader_quirks(tmp_path):
    # v keeps the LAST three tokens (:113-118); tabs are not separators (:23); blank lines and comments skipped
    obj = "# c\n\nv 9 1 2 3\nv 0 0 1\nv\t5 5 5\nv 0 1 0\nf 1 2 3\n"
    v = np.array(loadObj(_write(tmp_path, obj))).reshape(-1, 8)
    assert v[:, :3].tolist() == [[1, 2, 3], [0, 0, 1], [0, 1, 0]]
    # relative index -1 reaches the second-to-last vertex, exactly like the reference's int(tok)-1 lookup
    obj = "v 1 0 0\nv 2 0 0\nv 3 0 0\nf -1 1 2\n"
    v = np.array(loadObj(_write(tmp_path, obj))).reshape(-1, 8)
    assert v[0, :3].tolist() == [2, 0, 0]


def test_rect_box_geometry():
    r = Rect([2, 4, 6], [0, 0, 0], [0, 0, 0])
    assert r.pos.shape == (36, 3) and r.pos.dtype == np.float32
    assert np.abs(r.pos).max(axis=0).tolist() == [1, 2, 3]
    assert np.allclose(np.linalg.norm(r.normals, axis=1), 1)
    # first face is +Z, triangles (0,1,2),(0,2,3) (object.py:266-303)
    assert r.pos[:6].tolist() == [[-1, -2, 3], [1, -2, 3], [1, 2, 3], [-1, -2, 3], [1, 2, 3], [-1, 2, 3]]
    # rotation order: R = Rx @ Ry @ Rz (Z applied first) — 90 deg about Z then 90 about X
    r2 = Rect([2, 2, 2], [10, 0, 0], [90, 0, 90], scale=2)
    rz = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1]], dtype=np.float64)
    rx = np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0]], dtype=np.float64)
    want = (rx @ rz @ (2 * Rect([2, 2, 2], [0, 0, 0], [0, 0, 0]).pos.T.astype(np.float64))).T + [10, 0, 0]
    assert np.allclose(r2.pos, want, atol=1e-5)


def test_scene_layout_and_tree_invariants():
    cfg = configs.scene_configs()["rects_box"]
    sc = Scene(configs.build_objects(cfg.objects))
    tris, boxes, indices = sc.pack()
    assert tris.dtype == triangleStruct and tris.dtype.itemsize == 144
    assert boxes.dtype == boundingBoxStruct and boxes.dtype.itemsize == 48
    assert indices.dtype == np.uint32
    n = len(tris)
    assert n == 12 * 9 and sc.total_triangles == n
    depth = math.ceil(math.log(n, 2)) - 1
    # material of the emitter (object 8 of 9) landed on its 12 triangles (scene.py:116-133)
    assert (tris["surface"][:, 0] > 0).sum() == 12
    # walk the tree as the shader does: every triangle is reached exactly once through reachable leaves
    seen, stack, max_depth = [], [(0, 0)], 0
    while stack:
        i, d = stack.pop()
        max_depth = max(max_depth, d)
        b = boxes[i]
        if b["childA"] == -1:
            seen.extend(indices[b["triangleOffset"]:b["triangleOffset"] + b["numTriangles"]].tolist())
        else:
            assert 0 < b["childA"] < len(boxes) and 0 < b["childB"] < len(boxes)
            for ch in (b["childA"], b["childB"]):
                assert (boxes[ch]["posMin"] >= b["posMin"]).all() and (boxes[ch]["posMax"] <= b["posMax"]).all()
                stack.append((int(ch), d + 1))
    assert sorted(seen) == list(range(n))
    assert max_depth <= depth
    # bounds of the root cover every vertex
    allv = np.concatenate([tris[v]["pos"] for v in ("v0", "v1", "v2")])
    assert (boxes[0]["posMin"] == allv.min(axis=0)).all() and (boxes[0]["posMax"] == allv.max(axis=0)).all()


def test_camera_and_uniforms():
    right, fwd, up = configs.camera_basis((0, 0))
    assert right.tolist() == [1, 0, 0] and fwd.tolist() == [0, 0, 1] and up.tolist() == [0, 1, 0]
    p = configs.make_params(640, 360, 2, 1, (0, 1.5, -4), (0, -10))
    assert p.nBounces == 3                                   # main.py:186: bounces + 1
    assert p.numTilesX == 1 and p.tileSizeX == 640 and p.tileSizeY == 360
    assert math.isclose(p.xStep, math.pi / 2 * 640 / 360) and math.isclose(p.dirStartY, -math.pi / 4)
    p4 = configs.make_params(640, 360, 2, 1, (0, 0, 0), (0, 0), tile_size=4)
    assert (p4.tileSizeX, p4.tileSizeY, p4.numTilesX, p4.numTilesY) == (160, 90, 4, 4)
    r, f, u = configs.camera_basis((65, -25.4))
    for a, b in ((r, f), (r, u), (f, u)):
        assert abs(float(np.dot(a, b))) < 1e-6


@pytest.mark.parametrize("scene", ["rects_box", "cfg1_open", "cfg3_box"])
def test_buffers_byte_identical_to_reference_host_code(scene, tmp_path, scene_buffers):
    if not have_reference():
        pytest.skip("/root/reference not mounted")
    tris, boxes, indices, cfg = scene_buffers(scene)
    spec, out = tmp_path / "scene.json", tmp_path / "ref.npz"
    spec.write_text(json.dumps(cfg.objects))
    home = os.path.join(ROOT, "oracle", "_ref", "home")
    os.makedirs(home, exist_ok=True)
    proc = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ref_host_dump.py"), str(spec), str(out),
                           REFERENCE_ROOT], env=dict(os.environ, HOME=home), capture_output=True, text=True)
    assert proc.returncode == 0, proc.stderr[-2000:]
    ref = np.load(out)
    assert tris.tobytes() == ref["tris"].tobytes()
    assert boxes.tobytes() == ref["boxes"].tobytes()
    assert indices.tobytes() == ref["indices"].tobytes()
    assert int(ref["total_triangles"]) == len(tris) and int(ref["total_boxes"]) == len(boxes)
