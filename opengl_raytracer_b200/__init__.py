"""B200-native path tracer behind the host surface of pirosow/openGL_raytracer.

Host surface (same names as the reference): ``Mesh``, ``Rect`` (object.py), ``Scene`` (scene.py),
``BoundingBoxes`` (boundingBoxes.py), ``loadObj`` (loadObject.py), ``Screen`` (screen.py).
The hot path — the per-pixel loop of the reference's ``shaders/fragment.glsl`` — runs as hand-written
sm_100a CUDA kernels inside ``librt_b200.so`` (C ABI in ``include/rt_b200.h``), reached through
``rtlib`` with ctypes.  There is no CPU fallback: without the CUDA library every render call raises.
"""
from .object import Mesh, Rect          # noqa: F401
from .scene import Scene                # noqa: F401
from .boundingBoxes import BoundingBoxes  # noqa: F401
from .loadObject import loadObj         # noqa: F401

__all__ = ["Mesh", "Rect", "Scene", "BoundingBoxes", "loadObj"]
