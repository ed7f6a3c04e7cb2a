"""ctypes binding of ``librt_b200.so`` (C ABI: ``include/rt_b200.h``).

This is the drop-in boundary: where the reference's host code calls PyOpenGL (``glBufferData``,
``glUniform*``, ``glDrawArrays``, ``glReadPixels`` — /root/reference/scene.py:223-236,
main.py:175-202,376-399,434) this module calls the device library.  Errors surface as ``RuntimeError``
carrying ``rt_last_error`` — the counterpart of PyOpenGL raising ``GLError``.  There is no CPU path:
if the CUDA library is missing or no GPU is present the calls raise.

Like a GL context, a :class:`Context` can be made *current* so that ``Scene.send()`` and ``Screen``
find it without being handed one (the reference relies on the thread's current GL context).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

from .configs import RenderParams

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "csrc", "librt_b200.so")

c_f3 = ctypes.c_float * 3


class RtParams(ctypes.Structure):
    """``struct rt_params`` — the shader's uniform block (fragment.glsl:7-38)."""
    _fields_ = [
        ("tileX", ctypes.c_int32), ("tileY", ctypes.c_int32),
        ("tileSizeX", ctypes.c_int32), ("tileSizeY", ctypes.c_int32),
        ("numTilesX", ctypes.c_int32), ("numTilesY", ctypes.c_int32),
        ("total_frames", ctypes.c_int32), ("lambertian", ctypes.c_int32),
        ("skyBrightness", ctypes.c_float), ("fov", ctypes.c_float),
        ("xStep", ctypes.c_float), ("yStep", ctypes.c_float),
        ("dirStartX", ctypes.c_float), ("dirStartY", ctypes.c_float),
        ("camPos", c_f3), ("camRight", c_f3), ("camUp", c_f3), ("camForward", c_f3),
        ("nBounces", ctypes.c_int32), ("rays_per_pixel", ctypes.c_int32),
        ("jitterAmount", ctypes.c_float), ("frameNumber", ctypes.c_int32),
        ("trisCount", ctypes.c_int32), ("boundingBoxCount", ctypes.c_int32),
    ]

    @classmethod
    def from_render_params(cls, p: RenderParams) -> "RtParams":
        out = cls()
        for name, _ in cls._fields_:
            value = getattr(p, name)
            if name in ("camPos", "camRight", "camUp", "camForward"):
                setattr(out, name, c_f3(*[float(x) for x in value]))
            else:
                setattr(out, name, value)
        return out


COUNTER_NAMES = ("rays", "rays_traced", "internal", "leaves", "tri_tests", "hits", "pops", "paths")
VARIANTS = {"auto": 0, "pixel": 1, "persistent": 2, "wavefront": 3, "pool": 4, "samples": 5, "chain": 6}
PASS_COUNTERS, PASS_PRIMARY_IDS, PASS_SKIP_MASKED, PASS_TO_PEER = 1, 2, 4, 8
IPC_HANDLE_BYTES = 64
STATIC_MAX_PIXELS = 1 << 20
SUM_CONTINUE = 1

_lib = None


def load_library(path: str | None = None):
    """dlopen ``librt_b200.so`` and declare every prototype of ``include/rt_b200.h``."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.isfile(path):
        raise RuntimeError(f"{path} not built — run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback for the render path)")
    lib = ctypes.CDLL(path)
    vp, i, u32, P = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint32, ctypes.POINTER
    protos = {
        "rt_create": (i, [P(vp), i]),
        "rt_destroy": (None, [vp]),
        "rt_last_error": (ctypes.c_char_p, [vp]),
        "rt_version": (ctypes.c_char_p, []),
        "rt_set_stream": (i, [vp, vp]),
        "rt_get_stream": (vp, [vp]),
        "rt_sync": (i, [vp]),
        "rt_scene_upload": (i, [vp, vp, u32, vp, u32, vp, u32]),
        "rt_scene_free": (i, [vp]),
        "rt_set_params": (i, [vp, P(RtParams)]),
        "rt_set_variant": (i, [vp, i]),
        "rt_last_pass_launches": (i, [vp]),
        "rt_last_pass_variant": (i, [vp]),
        "rt_measure_peaks": (i, [vp, P(ctypes.c_double)]),
        "rt_target_create": (i, [vp, i, i]),
        "rt_target_clear": (i, [vp]),
        "rt_target_free": (i, [vp]),
        "rt_render_pass": (i, [vp, i, i, i]),
        "rt_render_pass_ex": (i, [vp, i, i, i, i]),
        "rt_render_frames": (i, [vp, i, i]),
        "rt_render_frames_sum": (i, [vp, i, i, i]),
        "rt_render_frames_sum_ex": (i, [vp, i, i, i, i]),
        "rt_sum_resume": (i, [vp, i]),
        "rt_read_accum_f32": (i, [vp, vp]),
        "rt_write_accum_f32": (i, [vp, vp]),
        "rt_read_accum_f32_async": (i, [vp, vp]),
        "rt_read_wait": (i, [vp]),
        "rt_read_rgba8": (i, [vp, vp]),
        "rt_read_primary_ids": (i, [vp, vp]),
        "rt_read_counters": (i, [vp, vp, i]),
        "rt_reset_counters": (i, [vp]),
        "rt_accum_device_ptr": (vp, [vp, i]),
        "rt_accum_index": (i, [vp]),
        "rt_set_accum_index": (i, [vp, i]),
        "rt_band_rows": (i, [vp, i]),
        "rt_band_gather": (i, [vp, i, vp]),
        "rt_band_scatter": (i, [vp, i, vp]),
        "rt_render_pass_static": (i, [vp, i, i, i, i, i, vp, vp]),
        "rt_peer_export": (i, [vp, vp]),
        "rt_peer_open": (i, [vp, vp]),
        "rt_peer_close": (i, [vp]),
        "rt_peer_is_open": (i, [vp]),
        "rt_target_wait_reads": (i, [vp]),
        "rt_finish_sum": (i, [vp, i]),
        "rt_timer_start": (i, [vp]),
        "rt_timer_stop": (i, [vp, P(ctypes.c_float)]),
    }
    for name, (res, args) in protos.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    if path == LIB_PATH:
        _lib = lib
    return lib


_current = None


def current_context() -> "Context":
    if _current is None:
        raise RuntimeError("no current rt context — create Context(...) first (the reference needs a GL context here)")
    return _current


class Context:
    """One device context = the reference's GL context + compiled shader program (main.py:115-147)."""

    def __init__(self, device: int = 0, make_current: bool = True, lib_path: str | None = None):
        self.lib = load_library(lib_path)
        handle = ctypes.c_void_p()
        rc = self.lib.rt_create(ctypes.byref(handle), int(device))
        if rc != 0 or not handle:
            raise RuntimeError(f"rt_create(device={device}) failed with {rc}: is a CUDA GPU present?")
        self.handle = handle
        self.device = device
        self.width = self.height = 0
        if make_current:
            self.make_current()

    # -- plumbing
    def make_current(self):
        global _current
        _current = self

    def _check(self, rc):
        if rc != 0:
            msg = self.lib.rt_last_error(self.handle)
            raise RuntimeError(f"librt_b200 error {rc}: {msg.decode() if msg else '?'}")

    def close(self):
        global _current
        if getattr(self, "handle", None):
            self.lib.rt_destroy(self.handle)
            self.handle = None
            if _current is self:
                _current = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- scene (scene.py:223-236, 423-424)
    def scene_upload(self, tris: np.ndarray, boxes: np.ndarray, indices: np.ndarray):
        tris = np.ascontiguousarray(tris)
        boxes = np.ascontiguousarray(boxes)
        indices = np.ascontiguousarray(indices, dtype=np.uint32)
        if tris.dtype.itemsize != 144 and tris.dtype != np.uint8:
            raise ValueError("tris records must be 144 bytes")
        n_tris = tris.nbytes // 144
        n_boxes = boxes.nbytes // 48
        self._check(self.lib.rt_scene_upload(self.handle, tris.ctypes.data, n_tris, boxes.ctypes.data, n_boxes,
                                             indices.ctypes.data, indices.size))
        self.n_tris, self.n_boxes = n_tris, n_boxes

    def scene_free(self):
        self._check(self.lib.rt_scene_free(self.handle))

    # -- uniforms (main.py:175-202)
    def set_params(self, params):
        if isinstance(params, RenderParams):
            params = RtParams.from_render_params(params)
        self._params = params
        self._check(self.lib.rt_set_params(self.handle, ctypes.byref(params)))

    def set_variant(self, variant):
        self._check(self.lib.rt_set_variant(self.handle, VARIANTS.get(variant, variant)))

    def last_pass_launches(self) -> int:
        return int(self.lib.rt_last_pass_launches(self.handle))

    def last_pass_variant(self) -> str:
        code = int(self.lib.rt_last_pass_variant(self.handle))
        if code < 0:
            self._check(code)
        return {v: k for k, v in VARIANTS.items()}.get(code, str(code))

    def measure_peaks(self) -> dict:
        """Micro-benchmarked ceilings of this GPU: warp issue, L1 data-pipe wavefronts, L2 -> SM bandwidth."""
        out = (ctypes.c_double * 3)()
        self._check(self.lib.rt_measure_peaks(self.handle, out))
        return {"issue_gwarpinst_s": out[0], "l1_gwavefronts_s": out[1], "l2_gbs": out[2]}

    def set_stream(self, cuda_stream: int | None):
        """Launch on ``cuda_stream`` (a cudaStream_t as int; None = the context's own stream).  The new stream is
        ordered behind the work already queued on the old one, so switching needs no host synchronisation."""
        self._check(self.lib.rt_set_stream(self.handle, ctypes.c_void_p(cuda_stream or 0)))

    def get_stream(self) -> int:
        return int(self.lib.rt_get_stream(self.handle) or 0)

    def sync(self):
        self._check(self.lib.rt_sync(self.handle))

    # -- targets (screen.py:37-59)
    def target_create(self, width: int, height: int):
        self._check(self.lib.rt_target_create(self.handle, width, height))
        self.width, self.height = width, height

    def target_clear(self):
        self._check(self.lib.rt_target_clear(self.handle))

    def target_free(self):
        self._check(self.lib.rt_target_free(self.handle))
        self.width = self.height = 0

    # -- draw (main.py:376-394)
    def render_pass(self, frameNumber: int, tileX: int = 0, tileY: int = 0, flags: int = 0):
        self._check(self.lib.rt_render_pass_ex(self.handle, frameNumber, tileX, tileY, flags))

    def render_pass_static(self, frameNumber: int, tileX: int, tileY: int, width: int, height: int, prev_dev_ptr, next_dev_ptr: int):
        """One draw into caller-provided RGBA32F device buffers without any allocation (``rt_render_pass_static``: the
        wavefront queue lives in a static device array of ``STATIC_MAX_PIXELS`` entries shared by all calls on the GPU —
        calls that overlap on the GPU race).  ``prev_dev_ptr`` may be None for frame 0 of a tile that covers the frame."""
        self._check(self.lib.rt_render_pass_static(self.handle, frameNumber, tileX, tileY, width, height,
                                                   ctypes.c_void_p(prev_dev_ptr or None), ctypes.c_void_p(next_dev_ptr)))

    def render_frames(self, first_frame: int, n_frames: int):
        self._check(self.lib.rt_render_frames(self.handle, first_frame, n_frames))

    def render_frames_sum(self, first_frame: int, n_frames: int, frame_stride: int = 1, resume: bool = False):
        """Frames first, first + stride, ... kept as a SUM in the targets.  ``resume``: add to what the current target
        holds (see :meth:`sum_resume`) instead of starting a new sum."""
        self._check(self.lib.rt_render_frames_sum_ex(self.handle, first_frame, n_frames, frame_stride, SUM_CONTINUE if resume else 0))

    def sum_resume(self, prior_frames: int):
        """Current target: running mean of ``prior_frames`` frames -> their sum (0 clears it)."""
        self._check(self.lib.rt_sum_resume(self.handle, prior_frames))

    def finish_sum(self, total_frames: int):
        self._check(self.lib.rt_finish_sum(self.handle, total_frames))

    # -- read-backs (main.py:397-399, 434)
    def read_accum_f32(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._check(self.lib.rt_read_accum_f32(self.handle, out.ctypes.data))
        return out

    def read_accum_f32_async(self, host_ptr: int):
        """Start the read-back of the current target into pinned host memory at ``host_ptr`` (width*height*16 bytes);
        passes issued afterwards overlap it.  ``read_wait()`` returns when every started read-back has landed."""
        self._check(self.lib.rt_read_accum_f32_async(self.handle, ctypes.c_void_p(host_ptr)))

    def read_wait(self):
        self._check(self.lib.rt_read_wait(self.handle))

    def write_accum_f32(self, image: np.ndarray):
        """Upload an (H, W, 4) float32 accumulation into the current target (checkpoint restore)."""
        image = np.ascontiguousarray(image, dtype=np.float32)
        if image.shape != (self.height, self.width, 4):
            raise ValueError(f"accumulation must be {(self.height, self.width, 4)}, got {image.shape}")
        self._check(self.lib.rt_write_accum_f32(self.handle, image.ctypes.data))

    def read_rgba8(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.uint8)
        self._check(self.lib.rt_read_rgba8(self.handle, out.ctypes.data))
        return out

    def read_primary_ids(self) -> np.ndarray:
        out = np.empty((self.height, self.width), dtype=np.int32)
        self._check(self.lib.rt_read_primary_ids(self.handle, out.ctypes.data))
        return out

    def read_counters(self) -> dict:
        buf = np.zeros(len(COUNTER_NAMES), dtype=np.uint64)
        self._check(self.lib.rt_read_counters(self.handle, buf.ctypes.data, buf.size))
        return {k: int(v) for k, v in zip(COUNTER_NAMES, buf)}

    def reset_counters(self):
        self._check(self.lib.rt_reset_counters(self.handle))

    # -- zero-copy hand-off / sharding helpers
    def accum_device_ptr(self, which: int = -1) -> int:
        return int(self.lib.rt_accum_device_ptr(self.handle, which) or 0)

    @property
    def accum_index(self) -> int:
        return self.lib.rt_accum_index(self.handle)

    @accum_index.setter
    def accum_index(self, index: int):
        self._check(self.lib.rt_set_accum_index(self.handle, index))

    def band_rows(self, tileY: int) -> int:
        rows = self.lib.rt_band_rows(self.handle, tileY)
        if rows < 0:
            self._check(rows)
        return rows

    def band_gather(self, tileY: int, dense_dev_ptr: int):
        self._check(self.lib.rt_band_gather(self.handle, tileY, ctypes.c_void_p(dense_dev_ptr)))

    def band_scatter(self, tileY: int, dense_dev_ptr: int):
        self._check(self.lib.rt_band_scatter(self.handle, tileY, ctypes.c_void_p(dense_dev_ptr)))

    # -- peer tile mode (include/rt_b200.h: the exchange step folded into the render kernels) --
    def peer_export(self) -> bytes:
        """CUDA IPC handles of both targets (assembling rank); hand them to the other ranks' ``peer_open``."""
        buf = ctypes.create_string_buffer(2 * IPC_HANDLE_BYTES)
        self._check(self.lib.rt_peer_export(self.handle, buf))
        return buf.raw

    def peer_open(self, handles: bytes):
        if len(handles) != 2 * IPC_HANDLE_BYTES:
            raise ValueError(f"expected {2 * IPC_HANDLE_BYTES} bytes of IPC handles, got {len(handles)}")
        self._check(self.lib.rt_peer_open(self.handle, ctypes.c_char_p(handles)))

    def peer_close(self):
        self._check(self.lib.rt_peer_close(self.handle))

    @property
    def peer_is_open(self) -> bool:
        return bool(self.lib.rt_peer_is_open(self.handle))

    def target_wait_reads(self):
        self._check(self.lib.rt_target_wait_reads(self.handle))

    def timer_start(self):
        self._check(self.lib.rt_timer_start(self.handle))

    def timer_stop(self) -> float:
        ms = ctypes.c_float()
        self._check(self.lib.rt_timer_stop(self.handle, ctypes.byref(ms)))
        return ms.value

    @property
    def torch_device(self) -> str:
        return f"cuda:{self.device}"

    def use_torch_stream(self) -> int:
        """Run on torch's current CUDA stream so torch events and NCCL collectives order with the kernels.  Returns the
        stream the context was on (hand it to :meth:`set_stream` to go back; the switch keeps the order both ways)."""
        import torch
        previous = self.get_stream()
        handle = torch.cuda.current_stream(self.device).cuda_stream
        # torch reports the legacy default stream as 0, which rt_set_stream reads as "back to the context's own
        # stream"; cudaStreamLegacy (0x1) is the explicit handle of that stream
        self.set_stream(handle if handle else 1)
        return previous

    def accum_tensor(self, which: int = -1):
        """Zero-copy torch view (H, W, 4) float32 of an accumulation target (optional tensor hand-off)."""
        import torch

        class _Dev:
            pass

        holder = _Dev()
        holder.__cuda_array_interface__ = {
            "shape": (self.height, self.width, 4), "typestr": "<f4", "version": 3,
            "data": (self.accum_device_ptr(which), False), "strides": None,
        }
        return torch.as_tensor(holder, device=f"cuda:{self.device}")
