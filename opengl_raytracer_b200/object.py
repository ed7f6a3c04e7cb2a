"""Scene objects — host surface with the reference's constructor signatures.

``Mesh(pos, eulers, dirPath, color, emission_color, emission, roughness, scale)`` mirrors
``/root/reference/object.py:8`` and ``Rect(size, pos, eulers, color, emission_color, emission,
roughness, scale)`` mirrors ``/root/reference/object.py:241``.  Both expose ``.pos`` and ``.normals``
as world-space ``(n, 3)`` float32 arrays, ``.uvs``, and the per-object material
(``.color``, ``.emission_color``, ``.emission``, ``.roughness``) that ``Scene`` packs per triangle.

World transform (follow the reference's code, not its docstring — SURVEY §8 f3):

* degrees -> radians in float32 (``object.py:89``), rotation ``R = Rx @ Ry @ Rz`` — the loop at
  ``object.py:74-79`` left-multiplies in reversed "XYZ" order, so Z is applied to a vertex first;
* ``M = R @ diag(scale)``; ``pos' = (M @ pos.T).T + t`` in float32 (``object.py:92-98,45-48``);
* normals by ``inv(M).T`` then renormalised, zero-length rows left as they are (``object.py:49-52,102-106``).

The float32 matrix products go through the same numpy calls as the reference so the bytes agree on the
same machine (the parity tests check that against the reference's own ``object.py`` when it is mounted).
This is one-off host preprocessing, not the hot path.
"""
from __future__ import annotations

import os

import numpy as np

from . import loadObject

#: directory that holds ``<dirPath>/*.obj``; the reference hard-codes "models/" relative to the cwd
#: (``object.py:9``).  Resolution order: $RT_MODELS_DIR, ./models, <repo>/models.
_REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def models_root():
    env = os.environ.get("RT_MODELS_DIR")
    if env:
        return env
    if os.path.isdir("models"):
        return "models"
    return os.path.join(_REPO_ROOT, "models")


class _Placed:
    """Shared placement code: Euler XYZ rotation, uniform scale, translation."""

    def _place(self, pos, eulers, scale):
        self.position = np.array(pos, dtype=np.float32)
        self.eulers = np.array(eulers, dtype=np.float32)
        self.scale = np.array([scale, scale, scale], dtype=np.float32)

    def _take_vertices(self, table):
        table = np.asarray(table).reshape(-1, 8).astype(np.float32)
        self.total_vertices = table
        self.pos = table[:, 0:3].astype(np.float32)
        self.normals = table[:, 3:6].astype(np.float32)
        self.uvs = table[:, 6:8].astype(np.float32)
        self.pos, self.normals = self.getWorld()

    def rotation_matrix_from_euler(self, rx, ry, rz, order="XYZ"):
        # scalar calls on float32 angles, one per angle, like the reference (object.py:58-60)
        c = (np.cos(rx), np.cos(ry), np.cos(rz))
        s = (np.sin(rx), np.sin(ry), np.sin(rz))
        per_axis = {
            "X": np.array([[1, 0, 0], [0, c[0], -s[0]], [0, s[0], c[0]]], dtype=np.float32),
            "Y": np.array([[c[1], 0, s[1]], [0, 1, 0], [-s[1], 0, c[1]]], dtype=np.float32),
            "Z": np.array([[c[2], -s[2], 0], [s[2], c[2], 0], [0, 0, 1]], dtype=np.float32),
        }
        rot = np.eye(3, dtype=np.float32)
        for axis in reversed(order):
            rot = per_axis[axis] @ rot
        return rot

    def make_model_and_normal_matrices(self, position, euler_deg, scale=(1.0, 1.0, 1.0), order="XYZ"):
        rx, ry, rz = np.deg2rad(euler_deg)
        rot = self.rotation_matrix_from_euler(rx, ry, rz, order)
        linear = rot @ np.diag([scale[0], scale[1], scale[2]]).astype(np.float32)
        model = np.eye(4, dtype=np.float32)
        model[:3, :3] = linear
        model[:3, 3] = np.array([position[0], position[1], position[2]], dtype=np.float32)
        try:
            normal_mat = np.linalg.inv(model[:3, :3]).T.astype(np.float32)
        except np.linalg.LinAlgError:
            normal_mat = rot.astype(np.float32)
        return model, normal_mat

    def getWorld(self):
        model, normal_mat = self.make_model_and_normal_matrices(self.position, self.eulers, self.scale, "XYZ")
        world_pos = (model[:3, :3] @ self.pos.T).T + model[:3, 3]
        world_normals = (normal_mat @ self.normals.T).T
        lengths = np.linalg.norm(world_normals, axis=1, keepdims=True)
        lengths[lengths == 0] = 1.0
        return world_pos, world_normals / lengths


class Mesh(_Placed):
    """Triangle mesh read from ``models/<dirPath>/*.obj`` (last ``.obj`` in listdir order wins,
    reference ``object.py:17-23``)."""

    def __init__(self, pos, eulers, dirPath, color=[0, 0, 0], emission_color=[0, 0, 0], emission=0,
                 roughness=0, scale=1):
        directory = os.path.join(models_root(), dirPath)
        self._place(pos, eulers, scale)
        obj_path = None
        for name in os.listdir(directory):
            if name.endswith(".obj"):
                obj_path = os.path.join(directory, name)
        if obj_path is None:
            raise FileNotFoundError(f"no .obj file in {directory}")
        self._take_vertices(np.array(self.loadObjCompiled(obj_path)))
        self.color = color
        self.emission_color = emission_color
        self.emission = emission
        self.roughness = roughness

    def loadObjCompiled(self, filePath):
        return loadObject.loadObj(filePath)


# Box faces as (corner sign pattern x4, normal); corners are multiplied by the half extents.
# Order and winding follow the reference's box (+Z, -Z, +Y, -Y, +X, -X; two triangles (0,1,2),(0,2,3)
# per quad — object.py:266-303) because triangle order feeds the BVH and therefore tie-breaks.
_BOX_FACES = (
    (((-1, -1, 1), (1, -1, 1), (1, 1, 1), (-1, 1, 1)), (0.0, 0.0, 1.0)),
    (((1, -1, -1), (-1, -1, -1), (-1, 1, -1), (1, 1, -1)), (0.0, 0.0, -1.0)),
    (((-1, 1, 1), (1, 1, 1), (1, 1, -1), (-1, 1, -1)), (0.0, 1.0, 0.0)),
    (((-1, -1, -1), (1, -1, -1), (1, -1, 1), (-1, -1, 1)), (0.0, -1.0, 0.0)),
    (((1, -1, 1), (1, -1, -1), (1, 1, -1), (1, 1, 1)), (1.0, 0.0, 0.0)),
    (((-1, -1, -1), (-1, -1, 1), (-1, 1, 1), (-1, 1, -1)), (-1.0, 0.0, 0.0)),
)
_QUAD_UV = ((0.0, 0.0), (1.0, 0.0), (1.0, 1.0), (0.0, 1.0))


class Rect(_Placed):
    """Axis-aligned box primitive of extents ``size`` (12 triangles), then placed like a mesh."""

    def __init__(self, size, pos, eulers, color=[0, 0, 0], emission_color=[0, 0, 0], emission=0,
                 roughness=0, scale=1):
        self._place(pos, eulers, scale)
        self._take_vertices(self.make_cube_vertices(size))
        self.color = color
        self.emission_color = np.array(emission_color)
        self.emission = emission
        self.roughness = roughness

    def make_cube_vertices(self, size):
        half = (float(size[0]) / 2.0, float(size[1]) / 2.0, float(size[2]) / 2.0)
        rows = []
        for corners, normal in _BOX_FACES:
            quad = [tuple(sign * h for sign, h in zip(corner, half)) for corner in corners]
            for k in (0, 1, 2, 0, 2, 3):
                rows.append(quad[k] + normal + _QUAD_UV[k])
        return np.array(rows, dtype=np.float32)
