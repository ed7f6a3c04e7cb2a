"""GPU test (-m gpu) of rt_render_pass_static, the entry point that allocates nothing (include/rt_b200.h: the wavefront
queue in a static device array; no counterpart in the reference).  Named to be collected last.  Through the C ABI, two
calls in a row, against the existing entry point (rt_render_pass_ex) on the same inputs: bit-exact, the tolerance of
every parity test here.  Self-contained: its own context, the asset-free scene."""
import numpy as np
import pytest

from opengl_raytracer_b200 import configs, rtlib
from opengl_raytracer_b200.scene import Scene

pytestmark = pytest.mark.gpu

W, H, SPP = 200, 120, 3


def _setup(tile_size=1):
    cfg = configs.scene_configs()["rects_box"].scaled(W, H, SPP)
    tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
    params = cfg.params(tile_size=tile_size)
    ctx = rtlib.Context(0)
    ctx.scene_upload(tris, boxes, indices)
    ctx.set_params(params)
    return ctx, params


def _existing(ctx, params, frames, variant):
    """The frames as the existing entry point accumulates them, read back after each."""
    ctx.set_variant(variant)
    ctx.target_create(W, H)
    out = []
    for frame in range(frames):
        for ty in range(params.numTilesY):
            for tx in range(params.numTilesX):
                ctx.render_pass(frame, tx, ty)
        out.append(ctx.read_accum_f32())
    return out


def _static(ctx, params, frames, bufs, cur=None):
    """`cur`: the buffer that stands for the previous target at the start (None: frame 0 of an untiled frame reads none)."""
    import torch
    out = []
    for frame in range(frames):
        for ty in range(params.numTilesY):
            for tx in range(params.numTilesX):
                nxt = bufs[1] if cur is bufs[0] else bufs[0]
                ctx.render_pass_static(frame, tx, ty, W, H, cur.data_ptr() if cur is not None else None, nxt.data_ptr())
                cur = nxt
        ctx.sync()
        out.append(cur.cpu().numpy().copy())
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("variant", ["wavefront", "auto"])
def test_static_pass_equals_existing_entry_point_twice_in_a_row(variant):
    import torch
    ctx, params = _setup()
    try:
        want = _existing(ctx, params, 2, variant)
        # garbage in the caller's buffers: every pixel of the frame must be written
        bufs = [torch.full((H, W, 4), float("nan"), device="cuda:0") for _ in range(2)]
        torch.cuda.synchronize()
        for attempt in range(2):          # the static array keeps the last call's contents: run the pair twice
            got = _static(ctx, params, 2, bufs)
            for frame in range(2):
                assert got[frame].tobytes() == want[frame].tobytes(), \
                    f"attempt {attempt}, frame {frame}: {(got[frame] != want[frame]).any(-1).sum()} pixels differ"
    finally:
        ctx.close()


def test_static_pass_with_the_reference_tiling():
    """3 x 3 tiles drawn one per call (main.py:409-418): partial tiles copy the previous buffer outside the tile."""
    import torch
    ctx, params = _setup(tile_size=3)
    try:
        assert params.numTilesX > 1 and params.numTilesY > 1
        want = _existing(ctx, params, 2, "wavefront")
        bufs = [torch.zeros((H, W, 4), device="cuda:0") for _ in range(2)]     # the cleared targets of rt_target_create
        torch.cuda.synchronize()
        got = _static(ctx, params, 2, bufs, cur=bufs[1])
        for frame in range(2):
            assert got[frame].tobytes() == want[frame].tobytes()
    finally:
        ctx.close()


def test_static_pass_allocates_nothing_and_refuses_what_does_not_fit():
    import torch
    ctx, params = _setup()
    try:
        bufs = [torch.zeros((H, W, 4), device="cuda:0") for _ in range(2)]
        torch.cuda.synchronize()
        _static(ctx, params, 1, bufs)                      # first use loads the kernels
        free_before, _ = torch.cuda.mem_get_info(0)
        _static(ctx, params, 2, bufs)
        free_after, _ = torch.cuda.mem_get_info(0)
        assert free_after == free_before, (free_before, free_after)
        # a tile of 2048 x 1024 = 2 Mi pixels: over RT_STATIC_MAX_PIXELS, refused before any buffer is touched
        big = configs.scene_configs()["rects_box"].scaled(2048, 1024, 1).params()
        ctx.set_params(big)
        with pytest.raises(RuntimeError, match="RT_STATIC_MAX_PIXELS"):
            ctx.render_pass_static(0, 0, 0, 2048, 1024, None, bufs[0].data_ptr())
        ctx.set_params(params)
        with pytest.raises(RuntimeError, match="prev_dev"):
            ctx.render_pass_static(1, 0, 0, W, H, None, bufs[0].data_ptr())
        with pytest.raises(RuntimeError, match="65535"):
            ctx.render_pass_static(0, 0, 0, 70000, 4, None, bufs[0].data_ptr())
        assert rtlib.STATIC_MAX_PIXELS == 1 << 20
    finally:
        ctx.close()
