"""Render parameters, camera and the named BASELINE.json scenes.

``camera_basis`` restates ``/root/reference/main.py:211-237`` (float64 trig -> float32 vectors,
``right = cross(world_up, forward)``, ``up = cross(forward, right)``) and ``make_params`` restates the
uniform derivation of ``main.py:166-202``: the "fov" of pi/2 is used as a *linear* extent (no tan),
``dirStart = -fov/2 (* aspect)``, ``step = fov (* aspect)``, and the shader's ``nBounces`` uniform is
``bounces + 1`` (``main.py:186``).  The scene table follows SURVEY.md §8(d); a scene is a list of plain
dict object specs so the same description can be instantiated with this package's ``Mesh``/``Rect`` and
(in the parity tests) with the reference's own classes.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np


def camera_basis(cam_dir):
    """(yaw_deg, pitch_deg) -> (camRight, camForward, camUp) float32, reference ``main.py:211-237``."""
    yaw_deg, pitch_deg = cam_dir
    yaw, pitch = math.radians(yaw_deg), math.radians(pitch_deg)
    forward = np.array([math.sin(yaw) * math.cos(pitch), math.sin(pitch), math.cos(yaw) * math.cos(pitch)],
                       dtype=np.float32)
    forward /= np.linalg.norm(forward)
    right = np.cross(np.array([0.0, 1.0, 0.0], dtype=np.float32), forward)
    right /= np.linalg.norm(right)
    up = np.cross(forward, right)
    up /= np.linalg.norm(up)
    return right, forward, up


@dataclass
class RenderParams:
    """Mirror of the shader's uniform block (``fragment.glsl:7-38``), same names."""
    width: int
    height: int
    tileX: int = 0
    tileY: int = 0
    tileSizeX: int = 1
    tileSizeY: int = 1
    numTilesX: int = 1
    numTilesY: int = 1
    total_frames: int = 0
    lambertian: int = 1
    skyBrightness: float = 1.0
    fov: float = math.pi / 2
    xStep: float = 0.0
    yStep: float = 0.0
    dirStartX: float = 0.0
    dirStartY: float = 0.0
    camPos: tuple = (0.0, 0.0, 0.0)
    camRight: tuple = (1.0, 0.0, 0.0)
    camUp: tuple = (0.0, 1.0, 0.0)
    camForward: tuple = (0.0, 0.0, 1.0)
    nBounces: int = 1
    rays_per_pixel: int = 1
    jitterAmount: float = 0.0
    frameNumber: int = 0
    trisCount: int = 0
    boundingBoxCount: int = 0

    def replace(self, **kw):
        out = RenderParams(**self.__dict__)
        for k, v in kw.items():
            setattr(out, k, v)
        return out


def make_params(width, height, bounces, rays_per_pixel, cam_pos, cam_dir, jitter_amount=0.001,
                lambertian=True, sky_brightness=1.0, aspect=None, tile_size=1):
    """Derive the uniform values the way ``App.__init__`` does (``main.py:125-137,151-202``).

    ``tile_size`` is the reference's ``tileSize`` constructor argument: the frame is cut into
    ``tile_size x tile_size`` tiles drawn one per pass.  ``aspect`` defaults to width/height (the
    reference uses the display's aspect, ``main.py:137``).
    """
    aspect = (width / height) if aspect is None else aspect
    fov = np.radians(90)
    right, forward, up = camera_basis(cam_dir)
    tsx, tsy = width // tile_size, height // tile_size
    return RenderParams(
        width=width, height=height,
        tileSizeX=tsx, tileSizeY=tsy,
        numTilesX=int((width + tsx - 1) / tsx), numTilesY=int((height + tsy - 1) / tsy),
        lambertian=int(bool(lambertian)), skyBrightness=float(sky_brightness), fov=float(fov),
        xStep=float(fov * aspect), yStep=float(fov),
        dirStartX=float(-fov / 2 * aspect), dirStartY=float(-fov / 2),
        camPos=tuple(float(x) for x in np.array(cam_pos, dtype=np.float32)),
        camRight=tuple(float(x) for x in right), camUp=tuple(float(x) for x in up),
        camForward=tuple(float(x) for x in forward),
        nBounces=bounces + 1, rays_per_pixel=rays_per_pixel, jitterAmount=float(jitter_amount),
    )


# ----------------------------------------------------------------------------- scene descriptions
def mesh(pos, eulers, dirPath, color=(0, 0, 0), emission_color=(0, 0, 0), emission=0, roughness=0, scale=1):
    return dict(kind="Mesh", pos=list(pos), eulers=list(eulers), dirPath=dirPath, color=list(color),
                emission_color=list(emission_color), emission=emission, roughness=roughness, scale=scale)


def rect(size, pos, eulers, color=(0, 0, 0), emission_color=(0, 0, 0), emission=0, roughness=0, scale=1):
    return dict(kind="Rect", size=list(size), pos=list(pos), eulers=list(eulers), color=list(color),
                emission_color=list(emission_color), emission=emission, roughness=roughness, scale=scale)


def box_room():
    """The seven ``Rect``s of ``main.py:37-99`` in the order ``main.py:101-111`` lists them."""
    red = rect([8, 5, 0.1], [0, 0, 30], [0, 0, 0], [1, 0.25, 0.3], roughness=1, scale=10)
    blue = rect([8, 5, 0.1], [0, 0, -30], [0, 0, 0], [0.3, 0.25, 1], roughness=1, scale=10)
    green = rect([8, 6, 0.1], [0, -25, 0], [90, 0, 0], [0.25, 1, 0.3], roughness=1, scale=10)
    back = rect([6, 8, 0.1], [-35, 0, 0], [0, 90, 0], [0.9, 0.9, 0.9], roughness=1, scale=10)
    front = rect([6, 8, 0.1], [25, 0, 0], [0, 90, 0], [0.9, 0.9, 0.9], roughness=0, scale=10)
    floor = rect([8, 6, 0.1], [0, 25, 0], [90, 0, 0], [1, 1, 1], roughness=1, scale=10)
    light = rect([5, 5, 0.25], [0, 23.9, 0], [-90, 0, 0], [0, 0, 0], [1, 1, 1], 1.5, scale=5)
    return [red, blue, green, front, floor, light, back]


BOX_CAMERA = dict(cam_pos=(-33.7, 14.8, -21.1), cam_dir=(65, -25.4))      # main.py:151-152
_MINI = dict(color=[0.96, 0.96, 0.86], roughness=1)


@dataclass
class SceneConfig:
    name: str
    objects: list
    width: int
    height: int
    bounces: int
    rays_per_pixel: int
    cam_pos: tuple
    cam_dir: tuple
    frames: int = 1
    note: str = ""
    extra: dict = field(default_factory=dict)

    def params(self, **overrides):
        kw = dict(width=self.width, height=self.height, bounces=self.bounces,
                  rays_per_pixel=self.rays_per_pixel, cam_pos=self.cam_pos, cam_dir=self.cam_dir)
        kw.update(overrides)
        return make_params(**kw)

    def scaled(self, width, height, rays_per_pixel=None, frames=None):
        return SceneConfig(self.name, self.objects, width, height, self.bounces,
                           self.rays_per_pixel if rays_per_pixel is None else rays_per_pixel,
                           self.cam_pos, self.cam_dir, self.frames if frames is None else frames, self.note)


def _main_py_pair():
    """Dragon + mirror sphere of ``main.py:19-35`` with the missing medium dragon replaced."""
    dragon = mesh([-5, -10, 0], [270, 0, -90], "stanford_minidragon", [0.96, 0.96, 0.86], roughness=1, scale=0.25)
    sphere = mesh([-25, -20, 20], [0, 0, 0], "sphere", color=[1, 1, 1], roughness=0, scale=7)
    return [dragon, sphere]


def scene_configs():
    """name -> SceneConfig for every row of SURVEY.md §8(d) (open and box variants)."""
    cfg = {}
    cfg["cfg1_open"] = SceneConfig(
        "cfg1_open",
        [mesh([0, 1, 0], [0, 0, 0], "sphere", [0.9, 0.3, 0.3], roughness=1, scale=1),
         mesh([0, 0, 0], [0, 0, 0], "ground", [0.8, 0.8, 0.8], roughness=1)],
        640, 360, 2, 1, (0, 1.5, -4), (0, -10), note="models/sphere + models/ground")
    cfg["cfg2_open"] = SceneConfig(
        "cfg2_open", [mesh([0, 0, 0], [0, 0, 0], "stanford_minidragon", **_MINI)],
        1280, 720, 4, 8, (1, -4, -160), (0, 0), note="models/stanford_minidragon")
    cfg["cfg2_box"] = SceneConfig(
        "cfg2_box", _main_py_pair() + box_room(), 1280, 720, 4, 8, **BOX_CAMERA,
        note="main.py scene with stanford_minidragon substituted for the missing medium dragon")
    knight_open = mesh([30, 0, -60], [0, 0, 0], "knight", [0.8, 0.6, 0.2], roughness=0.5, scale=4)
    cfg["cfg3_open"] = SceneConfig(
        "cfg3_open",
        [knight_open,
         mesh([-40, 10, -60], [0, 0, 0], "sphere", [1, 1, 1], roughness=0, scale=7),
         mesh([0, 0, 0], [0, 0, 0], "stanford_minidragon", **_MINI),
         mesh([0, -45, 0], [0, 0, 0], "ground", [0.8, 0.8, 0.8], roughness=1, scale=20)],
        1920, 1080, 4, 16, (1, -4, -160), (0, 0),
        note="car/airplane OBJ missing: knight + sphere + minidragon + ground substituted")
    cfg["cfg3_box"] = SceneConfig(
        "cfg3_box",
        _main_py_pair() + [mesh([10, -24.9, -15], [270, 0, 0], "knight", [0.8, 0.6, 0.2], roughness=0.5, scale=4)]
        + box_room(), 1920, 1080, 4, 16, **BOX_CAMERA, note="boxed multi-object scene")
    cfg["cfg4_open"] = SceneConfig(
        "cfg4_open", [mesh([0, 0, 0], [0, 0, 0], "dragon", **_MINI)],
        3840, 2160, 4, 64, (0, 30, -90), (0, 0),
        note="models/dragon (stands in for the missing stanford_mediumdragon)")
    cfg["cfg4_box"] = SceneConfig(
        "cfg4_box", [mesh([0, -19.1, 0], [0, 0, 0], "dragon", scale=0.3, **_MINI)] + box_room(),
        3840, 2160, 4, 64, **BOX_CAMERA, note="dragon inside the main.py box")
    cfg["cfg5_open"] = SceneConfig(
        "cfg5_open", cfg["cfg4_open"].objects, 3840, 2160, 4, 1, (0, 30, -90), (0, 0), frames=1024,
        note="models/dragon, progressive accumulation")
    cfg["cfg5_box"] = SceneConfig(
        "cfg5_box", cfg["cfg4_box"].objects, 3840, 2160, 4, 1, frames=1024, **BOX_CAMERA,
        note="boxed dragon, progressive accumulation")
    # asset-free scene (Rect only): used by smoke() and tests that must not depend on staged models
    cfg["rects_box"] = SceneConfig(
        "rects_box",
        [rect([3, 3, 3], [-5, -18, 5], [0, 30, 0], [0.9, 0.8, 0.3], roughness=1, scale=3),
         rect([2, 6, 2], [8, -16, -8], [0, -20, 0], [0.8, 0.8, 0.9], roughness=0.2, scale=3)] + box_room(),
        320, 180, 4, 2, **BOX_CAMERA, note="two boxes inside the main.py room; no OBJ assets needed")
    return cfg


def build_objects(specs, mesh_cls=None, rect_cls=None):
    """Instantiate object specs with this package's classes (or the ones given)."""
    if mesh_cls is None or rect_cls is None:
        from .object import Mesh, Rect
        mesh_cls, rect_cls = mesh_cls or Mesh, rect_cls or Rect
    out = []
    for spec in specs:
        spec = dict(spec)
        kind = spec.pop("kind")
        if kind == "Mesh":
            out.append(mesh_cls(spec.pop("pos"), spec.pop("eulers"), spec.pop("dirPath"), **spec))
        else:
            out.append(rect_cls(spec.pop("size"), spec.pop("pos"), spec.pop("eulers"), **spec))
    return out
