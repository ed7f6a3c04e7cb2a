"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI (ctypes), against the CPU oracle on the
same inputs.  Bar: RGBA32F accumulation bit-exact, primary-hit triangle ids bit-exact, work counters equal,
8-bit read-back equal.  Nothing here reads /root/reference."""
import numpy as np
import pytest

from conftest import GOLDEN_CASES, load_golden
from opengl_raytracer_b200 import rtlib

pytestmark = pytest.mark.gpu

VARIANTS = ["pixel", "persistent", "pool", "wavefront", "samples", "chain"]
EXACT_COUNTERS = ("rays", "internal", "leaves", "tri_tests", "hits", "pops", "paths")


@pytest.fixture(scope="module")
def ctx():
    c = rtlib.Context(0)
    yield c
    c.close()


def _render_gpu(ctx, tris, boxes, indices, params, frames, variant, flags=0):
    ctx.set_variant(variant)
    ctx.scene_upload(tris, boxes, indices)
    ctx.set_params(params)
    ctx.target_create(params.width, params.height)
    ctx.reset_counters()
    for frame in range(frames):
        for ty in range(params.numTilesY):
            for tx in range(params.numTilesX):
                ctx.render_pass(frame, tx, ty, flags=flags)
    return ctx.read_accum_f32()


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_golden_frames(ctx, case, variant):
    """Frames rendered by the reference's own shader (compiled as C++) — committed fixtures."""
    g = load_golden(case)
    got = _render_gpu(ctx, g["tris"], g["boxes"], g["indices"], g["params"], g["frames"], variant)
    assert got.tobytes() == g["expected"].tobytes(), f"{(got != g['expected']).any(-1).sum()} pixels differ"


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("scene,w,h,spp,frames,over", [
    ("rects_box", 160, 90, 3, 2, {}),
    ("rects_box", 97, 61, 1, 1, {"jitter_amount": 0.0}),               # ragged size: partial warps / blocks
    ("rects_box", 64, 36, 2, 1, {"lambertian": False, "sky_brightness": 0.5}),
    ("rects_box", 64, 36, 1, 1, {"bounces": 0}),                        # nBounces = 1: primary only
    ("cfg1_open", 320, 180, 1, 2, {}),
    ("cfg2_open", 320, 180, 4, 1, {}),
    ("cfg2_box", 192, 108, 2, 2, {}),
    ("cfg3_box", 192, 108, 2, 1, {}),
    ("cfg4_open", 256, 144, 2, 1, {}),
    ("cfg4_box", 192, 108, 2, 1, {}),
])
def test_image_ids_counters_match_oracle(ctx, oracle_mod, scene_buffers, scene, w, h, spp, frames, over, variant):
    tris, boxes, indices, cfg = scene_buffers(scene)
    params = cfg.scaled(w, h, spp).params(**over)
    got = _render_gpu(ctx, tris, boxes, indices, params, frames, variant,
                      flags=rtlib.PASS_COUNTERS | rtlib.PASS_PRIMARY_IDS)
    ids = ctx.read_primary_ids()
    cnt = ctx.read_counters()
    acc, want = None, {k: 0 for k in EXACT_COUNTERS}
    for frame in range(frames):
        acc, ref_ids, c = oracle_mod.render_pass(tris, boxes, indices, params.replace(frameNumber=frame), prev=acc,
                                                 want_ids=True, want_counters=True)
        for k in EXACT_COUNTERS:
            want[k] += c[k]
    assert np.array_equal(ids, ref_ids), f"{(ids != ref_ids).sum()} primary ids differ"
    assert got.tobytes() == acc.tobytes(), f"{(got != acc).any(-1).sum()} pixels differ"
    assert {k: cnt[k] for k in EXACT_COUNTERS} == want
    assert cnt["rays_traced"] == cnt["rays"] - (spp - 1) * w * h * frames * (1 if params.nBounces > 0 else 0)
    assert np.array_equal(ctx.read_rgba8(), oracle_mod.to_rgba8(acc))


@pytest.mark.parametrize("mode", ["0,0,0", "0,0,1", "0,0,2"], ids=["auto", "speculative", "sequential"])
@pytest.mark.parametrize("scene,w,h,spp,frames,over", [
    ("rects_box", 96, 54, 5, 2, {}),                      # closed room: forced speculation leaves its table -> on-demand tail
    ("rects_box", 64, 36, 3, 1, {"lambertian": False, "sky_brightness": 0.5}),
    ("rects_box", 64, 36, 4, 1, {"bounces": 0}),          # every path ends at bounce 0
    ("cfg1_open", 160, 90, 4, 2, {}),
    ("cfg4_open", 192, 108, 24, 1, {}),                   # open sky: the case the variant is for
    ("cfg2_box", 96, 54, 3, 1, {}),                       # emitter + mirror sphere
    ("cfg4_open", 96, 54, 96, 1, {}),                     # many samples: chains that run through several size classes
    ("rects_box", 40, 24, 40, 1, {"bounces": 7}),         # 8 hits per sample: the last round's bound, a full table buffer
    ("cfg2_box", 64, 36, 48, 2, {}),
])
@pytest.mark.parametrize("variant", ["samples", "chain"])
def test_samples_variant_matches_oracle(ctx, oracle_mod, scene_buffers, monkeypatch, scene, w, h, spp, frames, over, mode, variant):
    """The speculative variants (parallel evaluation of F(k) + per-pixel chain; SAMPLES through a global table in rounds,
    CHAIN inside one warp) without counting flags, in their device-chosen mode and with either mode forced: RGBA32F
    bit-exact against the oracle, primary ids exact."""
    monkeypatch.setenv("RT_VOTE_TUNE", mode)
    tris, boxes, indices, cfg = scene_buffers(scene)
    params = cfg.scaled(w, h, spp).params(**over)
    got = _render_gpu(ctx, tris, boxes, indices, params, frames, variant, flags=rtlib.PASS_PRIMARY_IDS)
    ids = ctx.read_primary_ids()
    acc = None
    for frame in range(frames):
        acc, ref_ids, _ = oracle_mod.render_pass(tris, boxes, indices, params.replace(frameNumber=frame), prev=acc,
                                                 want_ids=True)
    assert np.array_equal(ids, ref_ids)
    assert got.tobytes() == acc.tobytes(), f"{(got != acc).any(-1).sum()} pixels differ"


@pytest.mark.parametrize("variant", VARIANTS)
def test_tile_passes_match_oracle_and_full_frame(ctx, oracle_mod, scene_buffers, variant):
    """The reference's time-sliced tiling (main.py:409-418): 3x3 tiles drawn one per pass; pixels outside the
    tile copy the previous target (fragment.glsl:382-386).  Same result as the oracle's tile loop."""
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    params = cfg.scaled(100, 58, 1).params(tile_size=3)
    assert params.numTilesX == 4 and params.numTilesY == 4        # 100//3 = 33 -> ceil(100/33) = 4 tiles
    got = _render_gpu(ctx, tris, boxes, indices, params, 2, variant)
    want = oracle_mod.render_frames(tris, boxes, indices, params, 0, 2)
    assert got.tobytes() == want.tobytes()
    full = _render_gpu(ctx, tris, boxes, indices, cfg.scaled(100, 58, 1).params(), 2, variant)
    assert got.tobytes() == full.tobytes()


def test_render_frames_equals_pass_loop(ctx, scene_buffers):
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    params = cfg.scaled(80, 45, 1).params(tile_size=2)
    want = _render_gpu(ctx, tris, boxes, indices, params, 3, "pixel")
    ctx.target_clear()
    ctx.render_frames(0, 3)
    assert ctx.read_accum_f32().tobytes() == want.tobytes()


def test_frame_sharded_sum_mode(ctx, scene_buffers):
    """spp-batch sharding: two 'ranks' render frames {0,2} and {1,3} as sums; sum / 4 differs
    from the sequential running mean only by float re-association."""
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    params = cfg.scaled(80, 45, 1).params()
    mean = _render_gpu(ctx, tris, boxes, indices, params, 4, "pixel")
    shares = []
    for rank in range(2):
        ctx.target_clear()
        ctx.render_frames_sum(rank, 2, 2)
        shares.append(ctx.read_accum_f32())
    total = (shares[0] + shares[1])[..., :3] / np.float32(4)
    assert np.abs(total - mean[..., :3]).max() < 1e-5
    # finish_sum: IEEE division by the frame count, alpha forced to 1
    ctx.target_clear()
    ctx.render_frames_sum(0, 4, 1)
    summed = ctx.read_accum_f32()
    ctx.finish_sum(4)
    fin = ctx.read_accum_f32()
    assert np.array_equal(fin[..., :3], summed[..., :3] / np.float32(4)) and (fin[..., 3] == 1).all()
    assert np.abs(fin[..., :3] - mean[..., :3]).max() < 1e-5


def test_band_sharding_reassembles_bit_exact(ctx, scene_buffers):
    """Multi-GPU tile mode on one device: G 'ranks' each render tileY = g of G interleaved bands with
    SKIP_MASKED, their rows are gathered densely and scattered into one image == the single-pass image."""
    import torch
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    w, h, bands, rows = 96, 70, 3, 8
    full = _render_gpu(ctx, tris, boxes, indices, cfg.scaled(w, h, 2).params(), 1, "pixel")
    params = cfg.scaled(w, h, 2).params().replace(tileSizeX=w, numTilesX=1, tileSizeY=rows, numTilesY=bands)
    ctx.set_params(params)
    dense = []
    for g in range(bands):
        ctx.target_clear()
        ctx.render_pass(0, 0, g, flags=rtlib.PASS_SKIP_MASKED)
        n = ctx.band_rows(g)
        buf = torch.empty((n, w, 4), dtype=torch.float32, device="cuda")
        ctx.band_gather(g, buf.data_ptr())
        ctx.sync()
        dense.append(buf)
    assert sum(d.shape[0] for d in dense) == h
    ctx.target_clear()
    for g in range(bands):
        ctx.band_scatter(g, dense[g].data_ptr())
    assert ctx.read_accum_f32().tobytes() == full.tobytes()


def test_invalid_scenes_are_rejected(ctx, scene_buffers):
    tris, boxes, indices, _ = scene_buffers("rects_box")
    bad = boxes.copy()
    bad["childA"][0] = len(boxes) + 5
    with pytest.raises(RuntimeError, match="child index"):
        ctx.scene_upload(tris, bad, indices)
    bad = boxes.copy()
    leaf = int(np.nonzero(boxes["childA"] == -1)[0][0])
    bad["triangleOffset"][leaf] = len(indices)
    with pytest.raises(RuntimeError, match="exceeds"):
        ctx.scene_upload(tris, bad, indices)
    bad_idx = indices.copy()
    bad_idx[:] = len(tris)
    with pytest.raises(RuntimeError, match="triangle index"):
        ctx.scene_upload(tris, boxes, bad_idx)
    bad = boxes.copy()
    inner = int(np.nonzero(boxes["childA"] != -1)[0][-1])
    bad["childA"][inner] = 0
    with pytest.raises(RuntimeError, match="cycle"):
        ctx.scene_upload(tris, bad, indices)


def test_call_order_errors(scene_buffers):
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    with rtlib.Context(0, make_current=False) as c:
        with pytest.raises(RuntimeError, match="rt_scene_upload"):
            c.render_pass(0)
        c.scene_upload(tris, boxes, indices)
        with pytest.raises(RuntimeError, match="rt_set_params"):
            c.render_pass(0)
        c.set_params(cfg.params())
        with pytest.raises(RuntimeError, match="rt_target_create"):
            c.render_pass(0)


def test_empty_scene_renders_sky(ctx):
    from opengl_raytracer_b200 import configs
    from opengl_raytracer_b200.scene import boundingBoxStruct, triangleStruct
    params = configs.make_params(32, 16, 2, 1, (0, 0, 0), (0, 0))
    ctx.scene_upload(np.zeros(0, dtype=triangleStruct), np.zeros(0, dtype=boundingBoxStruct), np.zeros(0, dtype=np.uint32))
    ctx.set_params(params)
    ctx.target_create(32, 16)
    ctx.render_pass(0)
    img = ctx.read_accum_f32()
    assert np.array_equal(img[0, 0], np.array([0.1, 0.6, 0.92, 1.0], dtype=np.float32)) and (img == img[0, 0]).all()


@pytest.mark.parametrize("scene", ["cfg4_open", "cfg4_box"])
def test_full_size_properties(ctx, scene_buffers, scene):
    """BASELINE size (3840x2160): size-independent properties instead of the (slow) oracle — the two kernel
    variants agree bit for bit, a 2x2 tile sequence equals the single pass, and an oracle spot-check of a
    64-row band of the same frame is bit-exact."""
    tris, boxes, indices, cfg = scene_buffers(scene)
    params = cfg.scaled(3840, 2160, 1).params()
    a = _render_gpu(ctx, tris, boxes, indices, params, 1, "pixel")
    b = _render_gpu(ctx, tris, boxes, indices, params, 1, "persistent")
    assert a.tobytes() == b.tobytes()
    t = _render_gpu(ctx, tris, boxes, indices, cfg.scaled(3840, 2160, 1).params(tile_size=2), 1, "pixel")
    assert a.tobytes() == t.tobytes()
    assert np.isfinite(a).all() and (a[..., 3] == 1).all()


def test_full_size_band_against_oracle(ctx, oracle_mod, scene_buffers):
    tris, boxes, indices, cfg = scene_buffers("cfg4_box")
    params = cfg.scaled(3840, 2160, 1).params()
    got = _render_gpu(ctx, tris, boxes, indices, params, 1, "pixel")
    # the oracle renders only tile row 17 of 135 (16 rows) of the same 4K frame via the shader's own tile mask
    band = params.replace(tileSizeX=3840, numTilesX=1, tileSizeY=16, numTilesY=135, tileY=17)
    want, _, _ = oracle_mod.render_pass(tris, boxes, indices, band)
    rows = slice(17 * 16, 18 * 16)
    assert got[rows].tobytes() == want[rows].tobytes()


def test_full_size_auto_runs_the_chain_variant(ctx, oracle_mod, scene_buffers):
    """The bench workload's shape (dragon under open sky, 3840x2160, here 16 spp): AUTO carries the chain variant and the
    persistent kernel, stage 1's count picks the chain variant on the device; the frame equals the persistent kernel's
    bit for bit, and an oracle spot-check of one 8-row band of it is bit-exact.  In the closed room the same launch
    sequence leaves the pass to the persistent kernel."""
    tris, boxes, indices, cfg = scene_buffers("cfg4_open")
    params = cfg.scaled(3840, 2160, 16).params()
    a = _render_gpu(ctx, tris, boxes, indices, params, 1, "auto")
    assert ctx.last_pass_variant() == "chain" and ctx.last_pass_launches() == 3
    b = _render_gpu(ctx, tris, boxes, indices, params, 1, "persistent")
    assert a.tobytes() == b.tobytes()
    band = params.replace(tileSizeX=3840, numTilesX=1, tileSizeY=8, numTilesY=270, tileY=140)
    want, _, _ = oracle_mod.render_pass(tris, boxes, indices, band)
    rows = slice(140 * 8, 141 * 8)
    assert a[rows].tobytes() == want[rows].tobytes()
    tris, boxes, indices, cfg = scene_buffers("cfg4_box")
    _render_gpu(ctx, tris, boxes, indices, cfg.scaled(3840, 2160, 16).params(), 1, "auto")
    assert ctx.last_pass_variant() == "persistent"


def test_headless_app_loop_and_checkpoint(tmp_path, oracle_mod):
    """The headless App (main.py's loop without a window): tile/frame sequencing equals the oracle's frame loop, and a
    run interrupted by save_checkpoint / load_checkpoint continues bit-identically."""
    from opengl_raytracer_b200 import configs
    from opengl_raytracer_b200.app import App
    from opengl_raytracer_b200.scene import Scene
    cfg = configs.scene_configs()["rects_box"]
    objs = lambda: configs.build_objects(cfg.objects)
    tris, boxes, indices = Scene(objs()).pack()
    app = App((96, 54), (96, 54), bounces=3, rays_per_pixel=2, tileSize=2, objects=objs(), cam_pos=cfg.cam_pos,
              cam_dir=cfg.cam_dir)
    assert app.params.nBounces == 4 and app.params.numTilesX == 2 and app.params.numTilesY == 2
    app.main(frames=3)
    assert app.screen.frame_count == 3 and app.total_frames == 12
    whole = app.screen.read_accum()
    want = oracle_mod.render_frames(tris, boxes, indices, app.params, 0, 3)
    assert whole.tobytes() == want.tobytes()
    # interrupted run: 1 frame + 1 extra pass, checkpoint, fresh app, resume
    app.resetFrames()
    assert app.screen.frame_count == 0 and not app.screen.read_accum().any()
    app.main(frames=1)
    app.step()
    ck = str(tmp_path / "ck.npz")
    app.save_checkpoint(ck)
    app.close()
    app2 = App((96, 54), (96, 54), bounces=3, rays_per_pixel=2, tileSize=2, objects=objs(), cam_pos=cfg.cam_pos,
               cam_dir=cfg.cam_dir)
    app2.load_checkpoint(ck)
    assert (app2.screen.frame_count, app2.tileX, app2.tileY) == (1, 1, 0)
    app2.main(frames=2)                     # finishes frame 1, then frame 2
    assert app2.screen.frame_count == 3
    assert app2.screen.read_accum().tobytes() == whole.tobytes()
    app2.save(str(tmp_path / "out.png"))
    app2.close()


def test_scene_reupload_reuses_buffers_and_rebuilds_records(ctx, scene_buffers):
    """rt_scene_upload keeps its device buffers and pinned staging between uploads and builds the triangle records on
    the device: a large scene, then a small one, then the large one again must render exactly as the first time."""
    big = scene_buffers("cfg2_box")
    small = scene_buffers("rects_box")
    frames = []
    for tris, boxes, indices, cfg in (big, small, big, small):
        params = cfg.scaled(80, 45, 2).params()
        frames.append(_render_gpu(ctx, tris, boxes, indices, params, 1, "pixel"))
    assert frames[0].tobytes() == frames[2].tobytes()
    assert frames[1].tobytes() == frames[3].tobytes()
    assert frames[0].tobytes() != frames[1].tobytes()


def test_launch_count_and_measured_peaks(ctx, scene_buffers):
    """The two measurement aids of the C ABI: kernels launched by the last pass, micro-benchmarked ceilings."""
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    params = cfg.scaled(64, 36, 2).params()
    _render_gpu(ctx, tris, boxes, indices, params, 1, "pixel")
    assert ctx.last_pass_launches() == 1
    _render_gpu(ctx, tris, boxes, indices, params, 1, "wavefront")
    assert ctx.last_pass_launches() == 2
    _render_gpu(ctx, tris, boxes, indices, params, 1, "samples")
    assert ctx.last_pass_launches() == 12                       # stage 1 + round 0 (2) + 3 later rounds x 3
    peaks = ctx.measure_peaks()
    assert 100.0 < peaks["issue_gwarpinst_s"] < 3000.0           # B200 nominal: 148 SMs x 4 x 1.965 GHz = 1163
    assert 20.0 < peaks["l1_gwavefronts_s"] < 1000.0             # nominal: 1 per cycle per SM = 291
    assert 1000.0 < peaks["l2_gbs"] < 100000.0


def test_async_readback_and_upload_overlap_keep_results(ctx, scene_buffers):
    """rt_read_accum_f32_async copies on its own stream while later passes run, and rt_scene_upload builds the next
    scene in the second buffer set while a pass still reads the first: both must leave every result untouched."""
    import torch
    big = scene_buffers("cfg2_box")
    small = scene_buffers("rects_box")
    w, h = 320, 180
    p_big, p_small = big[3].scaled(w, h, 6).params(), small[3].scaled(w, h, 6).params()
    want_big = [None, None]
    ctx.set_variant("auto")
    ctx.scene_upload(*big[:3]); ctx.set_params(p_big); ctx.target_create(w, h)
    ctx.render_pass(0, 0, 0); want_big[0] = ctx.read_accum_f32()
    ctx.render_pass(1, 0, 0); want_big[1] = ctx.read_accum_f32()
    ctx.scene_upload(*small[:3]); ctx.set_params(p_small); ctx.target_clear()
    ctx.render_pass(0, 0, 0); want_small = ctx.read_accum_f32()

    bufs = [torch.empty((h, w, 4), dtype=torch.float32).pin_memory() for _ in range(3)]
    ctx.scene_upload(*big[:3]); ctx.set_params(p_big); ctx.target_clear()
    ctx.render_pass(0, 0, 0)
    ctx.read_accum_f32_async(bufs[0].data_ptr())           # copy of frame 0 ...
    ctx.render_pass(1, 0, 0)                               # ... while frame 1 renders into the other target
    ctx.read_accum_f32_async(bufs[1].data_ptr())
    ctx.scene_upload(*small[:3])                           # next scene uploaded while frame 1 may still be in flight
    ctx.set_params(p_small)
    ctx.target_clear()                                     # must wait (on the device) for both copies
    ctx.render_pass(0, 0, 0)
    ctx.read_accum_f32_async(bufs[2].data_ptr())
    ctx.read_wait()
    assert bufs[0].numpy().tobytes() == want_big[0].tobytes()
    assert bufs[1].numpy().tobytes() == want_big[1].tobytes()
    assert bufs[2].numpy().tobytes() == want_small.tobytes()


@pytest.mark.parametrize("scene,w,h,spp,expect", [
    ("cfg4_open", 192, 108, 24, "samples"),        # sky behind the mesh: stage 1 queues few pixels -> the rounds render
    ("rects_box", 96, 54, 16, "persistent"),       # closed room: every primary ray hits -> the persistent kernel renders
    ("cfg2_box", 96, 54, 16, "persistent"),
])
def test_auto_decides_on_the_device(ctx, oracle_mod, scene_buffers, scene, w, h, spp, expect):
    """AUTO at >= 16 samples per pixel launches the samples rounds AND the persistent kernel; stage 1's count picks the
    one that runs.  Either way the frame is the oracle's, bit for bit, over two accumulated frames."""
    tris, boxes, indices, cfg = scene_buffers(scene)
    params = cfg.scaled(w, h, spp).params()
    got = _render_gpu(ctx, tris, boxes, indices, params, 2, "auto", flags=rtlib.PASS_PRIMARY_IDS)
    assert ctx.last_pass_variant() == expect
    assert ctx.last_pass_launches() == 13
    ids = ctx.read_primary_ids()
    acc = None
    for frame in range(2):
        acc, ref_ids, _ = oracle_mod.render_pass(tris, boxes, indices, params.replace(frameNumber=frame), prev=acc, want_ids=True)
    assert np.array_equal(ids, ref_ids)
    assert got.tobytes() == acc.tobytes(), f"{(got != acc).any(-1).sum()} pixels differ"


class _OneDeviceDist:
    """torch.distributed stand-in that plays a W-rank job on ONE device, rank after rank: `gather` / `reduce` are
    torch ops on torch's CURRENT stream (as NCCL's are ordered with it), so the renderer's kernels and the 'collective'
    only see each other's results if distributed.py orders the two streams — there is no host synchronisation here."""

    def __init__(self, world):
        self.world, self.dense, self.sums = world, {}, []
        self.rank = 0

    def gather(self, tensor, gather_list=None, dst=0):
        if self.rank != dst:
            self.dense[self.rank] = tensor.clone()
            return
        for g in range(self.world):
            gather_list[g].copy_(tensor if g == dst else self.dense[g])

    def reduce(self, tensor, dst=0):
        if self.rank != dst:
            self.sums.append(tensor.clone())
            return
        for other in self.sums:
            tensor.add_(other)


def test_distributed_functions_order_their_own_streams(ctx, scene_buffers):
    """render_tile_mode / render_frame_mode on a context that keeps ITS OWN stream while torch's current
    stream is a side stream: the functions themselves order renderer kernels and collectives (rt_set_stream keeps the
    order across the switch), no use_torch_stream() / sync() by the caller.  Tile mode reassembles bit-exactly, frame
    mode (run in two instalments: first_frame > 0 continues the mean) matches the sequential running mean."""
    import torch
    from opengl_raytracer_b200 import distributed
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    w, h, world = 640, 360, 3
    params = cfg.scaled(w, h, 2).params()
    full = _render_gpu(ctx, tris, boxes, indices, params, 1, "auto")
    mean7 = _render_gpu(ctx, tris, boxes, indices, cfg.scaled(w, h, 1).params(), 7, "auto")
    ctx.set_stream(None)
    own = ctx.get_stream()
    side = torch.cuda.Stream()
    fake = _OneDeviceDist(world)
    with torch.cuda.stream(side):
        for rank in list(range(1, world)) + [0]:            # rank 0 last: it needs the others' bands
            fake.rank = rank
            ctx.target_clear()
            distributed.render_tile_mode(ctx, fake, params, rank, world, frame=0, band_rows=8)
            assert ctx.get_stream() == own                   # put back where it was
        assert ctx.read_accum_f32().tobytes() == full.tobytes()
        # frame mode, frames 0..2 then 3..6 continued; every 'rank' uses the one context, so rank 0's mean is kept aside
        p1 = cfg.scaled(w, h, 1).params()
        kept = None
        for first, n in ((0, 3), (3, 4)):
            fake.sums = []
            for rank in list(range(1, world)) + [0]:
                fake.rank = rank
                ctx.target_clear()
                if rank == 0 and kept is not None:
                    previous = ctx.use_torch_stream()        # (scaffolding: the copy is a torch op on the side stream)
                    ctx.accum_tensor().copy_(kept)
                    ctx.set_stream(previous)
                distributed.render_frame_mode(ctx, fake, p1, rank, world, first, n, lambda c: c.accum_tensor())
            kept = ctx.accum_tensor().clone()
    got = ctx.read_accum_f32()
    assert np.abs(got[..., :3] - mean7[..., :3]).max() < 2e-5 and (got[..., 3] == 1).all()


def test_sum_mode_continuation_and_partial_tiles(ctx, scene_buffers):
    """rt_render_frames_sum_ex(RT_SUM_CONTINUE) adds to the sum the target holds; rt_sum_resume turns a mean back
    into a sum; a tile that does not cover the target is refused (masked pixels would carry stale values)."""
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    params = cfg.scaled(80, 45, 1).params()
    _render_gpu(ctx, tris, boxes, indices, params, 0, "pixel")
    ctx.target_clear()
    ctx.render_frames_sum(0, 5, 1)
    whole = ctx.read_accum_f32()
    ctx.target_clear()
    ctx.render_frames_sum(0, 2, 1)
    ctx.render_frames_sum(2, 3, 1, resume=True)
    assert ctx.read_accum_f32().tobytes() == whole.tobytes()
    # mean of 2 frames -> sum -> 3 more frames -> mean of 5
    ctx.target_clear()
    ctx.render_frames(0, 2)
    ctx.sum_resume(2)
    ctx.render_frames_sum(2, 3, 1, resume=True)
    ctx.finish_sum(5)
    a = ctx.read_accum_f32()
    ctx.target_clear()
    ctx.render_frames(0, 5)
    assert np.abs(a[..., :3] - ctx.read_accum_f32()[..., :3]).max() < 1e-5
    ctx.set_params(params.replace(tileSizeY=8, numTilesY=2))
    with pytest.raises(RuntimeError, match="covers the target"):
        ctx.render_frames_sum(0, 1, 1)
    ctx.set_params(params)


@pytest.mark.parametrize("variant", ["samples", "chain"])
@pytest.mark.parametrize("mode", ["0,0,1"], ids=["speculative"])
def test_samples_offsets_beyond_the_jump_tables(ctx, oracle_mod, scene_buffers, monkeypatch, mode, variant):
    """rays_per_pixel x hits beyond kMaxK = 16384 (closed room, 11 bounces, 2100 samples): the chain's RNG
    offsets leave the jump tables and are composed instead; still bit-exact."""
    monkeypatch.setenv("RT_VOTE_TUNE", mode)
    tris, boxes, indices, cfg = scene_buffers("rects_box")
    params = cfg.scaled(6, 4, 2100).params(bounces=10)
    got = _render_gpu(ctx, tris, boxes, indices, params, 1, variant)
    acc, _, cnt = oracle_mod.render_pass(tris, boxes, indices, params, want_counters=True)
    assert cnt["hits"] // (6 * 4) > 16384                     # the offsets really do leave the tables
    assert got.tobytes() == acc.tobytes(), f"{(got != acc).any(-1).sum()} pixels differ"


@pytest.mark.parametrize("seed", range(24))
@pytest.mark.parametrize("variant", ["auto"] + VARIANTS)
def test_random_scenes_match_oracle(ctx, oracle_mod, seed, variant):
    """Seeded random scenes (conftest.random_scene: boxes, slabs, mirrors, emitters; random bounces, samples, lambertian,
    sky, jitter, ragged sizes, tiling), two progressive frames in the reference's tile order: every variant bit-exact."""
    from conftest import random_scene
    tris, boxes, indices, params = random_scene(seed)
    got = _render_gpu(ctx, tris, boxes, indices, params, 2, variant)
    want = oracle_mod.render_frames(tris, boxes, indices, params, 0, 2)
    assert got.tobytes() == want.tobytes(), f"{(got != want).any(-1).sum()} pixels differ"


@pytest.mark.parametrize("seed", range(24))
@pytest.mark.parametrize("variant", ["samples", "chain"])
def test_random_scenes_with_speculation_forced(ctx, oracle_mod, monkeypatch, seed, variant):
    """The same scenes with the device-side gate overridden: the speculating variants evaluate samples ahead of the chain
    even where most pixels hit (emitters end paths early, mirrors and rooms make them long) and stay bit-exact."""
    from conftest import random_scene
    monkeypatch.setenv("RT_VOTE_TUNE", "0,0,1")
    tris, boxes, indices, params = random_scene(seed)
    got = _render_gpu(ctx, tris, boxes, indices, params, 2, variant)
    want = oracle_mod.render_frames(tris, boxes, indices, params, 0, 2)
    assert got.tobytes() == want.tobytes(), f"{(got != want).any(-1).sum()} pixels differ"


@pytest.mark.parametrize("scene,w,h,spp,over", [("cfg4_open", 192, 108, 24, {}), ("cfg1_open", 160, 90, 16, {}),
                                                ("rects_box", 64, 36, 12, {"bounces": 5})])
def test_chain_rings_are_never_read_before_they_are_cleared(ctx, oracle_mod, scene_buffers, monkeypatch, scene, w, h, spp, over):
    """The chain variant's rings live in uninitialised global memory; a slot clears its own ring at its first help.
    RT_CHAIN_POISON fills them beforehand with entries that look like valid results of small offsets (radiance 1e9): any
    read before the clear, or of an entry left by an earlier pixel, would show in the image.  Speculation forced on."""
    monkeypatch.setenv("RT_CHAIN_POISON", "1")
    monkeypatch.setenv("RT_VOTE_TUNE", "0,0,1")
    tris, boxes, indices, cfg = scene_buffers(scene)
    params = cfg.scaled(w, h, spp).params(**over)
    got = _render_gpu(ctx, tris, boxes, indices, params, 2, "chain")
    want = oracle_mod.render_frames(tris, boxes, indices, params, 0, 2)
    assert got.tobytes() == want.tobytes(), f"{(got != want).any(-1).sum()} pixels differ"


@pytest.mark.parametrize("variant", ["auto", "pool", "chain"])
def test_node_numbering_does_not_change_results(ctx, oracle_mod, scene_buffers, monkeypatch, variant):
    """rt_scene_upload numbers the internal nodes depth-first (subtrees contiguous); RT_NODE_ORDER=bfs keeps round 1's
    breadth-first numbering.  Only addresses change: image, primary ids and the work counters are the oracle's either way
    (POOL stages the first nodes of the array in shared memory, whichever they are)."""
    tris, boxes, indices, cfg = scene_buffers("cfg4_open")
    params = cfg.scaled(160, 90, 6).params()
    want, want_ids, want_cnt = oracle_mod.render_pass(tris, boxes, indices, params, want_ids=True, want_counters=True)
    flags = rtlib.PASS_PRIMARY_IDS | (0 if variant == "chain" else rtlib.PASS_COUNTERS)
    for order in ("dfs", "bfs"):
        monkeypatch.setenv("RT_NODE_ORDER", order)
        got = _render_gpu(ctx, tris, boxes, indices, params, 1, variant, flags=flags)
        assert got.tobytes() == want.tobytes(), order
        assert np.array_equal(ctx.read_primary_ids(), want_ids), order
        if variant != "chain":
            cnt = ctx.read_counters()
            assert {k: cnt[k] for k in EXACT_COUNTERS} == {k: want_cnt[k] for k in EXACT_COUNTERS}, order
