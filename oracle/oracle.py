"""ctypes bindings of the two CPU checkers.  TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py's
cpu_baseline / --impl reference legs); the product package never imports this module.

* ``port``  — oracle/librt_oracle.so, the hand-written restatement (rt_oracle.cpp) with work counters and
  the primary-hit id image;
* ``ref``   — oracle/_ref/libfragment_ref.so, the reference's own fragment.glsl compiled as C++
  (build_ref.py).  Image only (the shader has no other outputs).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from opengl_raytracer_b200.configs import RenderParams           # noqa: E402
from opengl_raytracer_b200.rtlib import COUNTER_NAMES, RtParams   # noqa: E402

PORT_LIB = os.path.join(HERE, "librt_oracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libfragment_ref.so")
_cache = {}


def build_port():
    subprocess.run(["make", "-C", HERE, "-s"], check=True)
    return PORT_LIB


def _load(path):
    if path not in _cache:
        _cache[path] = ctypes.CDLL(path)
    return _cache[path]


def have_ref():
    return os.path.isfile(REF_LIB)


def _as_struct(params):
    return RtParams.from_render_params(params) if isinstance(params, RenderParams) else params


def render_pass(tris, boxes, indices, params, prev=None, impl="port", want_ids=False, want_counters=False):
    """One draw call.  Returns (next_rgba32f[H,W,4], ids[H,W] or None, counters dict or None)."""
    p = _as_struct(params)
    w, h = params.width, params.height
    tris = np.ascontiguousarray(tris)
    boxes = np.ascontiguousarray(boxes)
    indices = np.ascontiguousarray(indices, dtype=np.uint32)
    if prev is None:
        prev = np.zeros((h, w, 4), dtype=np.float32)
    prev = np.ascontiguousarray(prev, dtype=np.float32)
    nxt = np.zeros((h, w, 4), dtype=np.float32)
    n_tris, n_boxes = tris.nbytes // 144, boxes.nbytes // 48
    vp = ctypes.c_void_p
    if impl == "ref":
        lib = _load(REF_LIB)
        lib.ref_render_pass.argtypes = [vp, ctypes.c_uint32, vp, ctypes.c_uint32, vp, ctypes.c_uint32,
                                        ctypes.POINTER(RtParams), ctypes.c_int, ctypes.c_int, vp, vp]
        rc = lib.ref_render_pass(tris.ctypes.data, n_tris, boxes.ctypes.data, n_boxes, indices.ctypes.data,
                                 indices.size, ctypes.byref(p), w, h, prev.ctypes.data, nxt.ctypes.data)
        assert rc == 0
        return nxt, None, None
    lib = _load(PORT_LIB)
    lib.oracle_render_pass.argtypes = [vp, ctypes.c_uint32, vp, ctypes.c_uint32, vp, ctypes.c_uint32,
                                       ctypes.POINTER(RtParams), ctypes.c_int, ctypes.c_int, vp, vp, vp, vp]
    ids = np.full((h, w), -2, dtype=np.int32) if want_ids else None
    cnt = np.zeros(len(COUNTER_NAMES), dtype=np.uint64) if want_counters else None
    rc = lib.oracle_render_pass(tris.ctypes.data, n_tris, boxes.ctypes.data, n_boxes, indices.ctypes.data,
                                indices.size, ctypes.byref(p), w, h, prev.ctypes.data, nxt.ctypes.data,
                                ids.ctypes.data if want_ids else None, cnt.ctypes.data if want_counters else None)
    assert rc == 0
    counters = {k: int(v) for k, v in zip(COUNTER_NAMES, cnt)} if want_counters else None
    return nxt, ids, counters


def render_frames(tris, boxes, indices, params, first_frame, n_frames, impl="port", accum=None):
    """The reference's frame loop (main.py:376-430) with every tile drawn per frame; returns the accumulation."""
    cur = accum
    for frame in range(first_frame, first_frame + n_frames):
        for ty in range(params.numTilesY):
            for tx in range(params.numTilesX):
                p = params.replace(frameNumber=frame, tileX=tx, tileY=ty)
                cur, _, _ = render_pass(tris, boxes, indices, p, prev=cur, impl=impl)
    return cur


def num_threads(impl="port"):
    lib = _load(REF_LIB if impl == "ref" else PORT_LIB)
    fn = lib.ref_num_threads if impl == "ref" else lib.oracle_num_threads
    fn.restype = ctypes.c_int
    return fn()


def random_values(seed, n):
    lib = _load(PORT_LIB)
    out = np.zeros(n, dtype=np.float32)
    states = np.zeros(n, dtype=np.uint32)
    lib.oracle_random_values.restype = ctypes.c_uint32
    lib.oracle_random_values.argtypes = [ctypes.c_uint32, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.oracle_random_values(seed, n, out.ctypes.data, states.ctypes.data)
    return out, states


def to_rgba8(rgba32f):
    """round(clamp(x,0,1)*255) — the 8-bit read-back of main.py:434 (GL float->unorm conversion)."""
    x = np.clip(np.nan_to_num(rgba32f, nan=0.0), 0.0, 1.0).astype(np.float32)
    return np.floor(x * np.float32(255.0) + np.float32(0.5)).astype(np.uint8)
