"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in opengl_raytracer_b200/distributed.py, driven
with a stand-in renderer that calls the CPU oracle (test infrastructure) instead of the CUDA library."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT
from opengl_raytracer_b200 import configs, distributed


def test_band_partition_covers_every_row_once():
    for height, rows, world in [(2160, 8, 8), (2160, 8, 4), (90, 8, 3), (7, 8, 2), (61, 4, 5)]:
        owned = [distributed.band_row_indices(height, rows, world, r) for r in range(world)]
        assert sorted(np.concatenate(owned).tolist()) == list(range(height))
        assert distributed.max_band_rows(height, rows, world) == max(len(o) for o in owned)


def test_frame_share_partitions_frames():
    for n, world in [(1024, 8), (10, 4), (3, 8), (1, 2)]:
        frames = []
        for r in range(world):
            first, count, stride = distributed.frame_share(5, n, r, world)
            frames += [first + k * stride for k in range(count)]
        assert sorted(frames) == list(range(5, 5 + n))


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from opengl_raytracer_b200.scene import Scene
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = configs.scene_configs()["rects_box"].scaled(48, 30, 1)
    tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
    params = cfg.params()
    # tile mode: every rank renders its bands with the oracle, dense rows are gathered to rank 0
    p = distributed.band_params(params, rank, world, band_rows=4)
    img, _, _ = oracle.render_pass(tris, boxes, indices, p)
    rows = distributed.band_row_indices(cfg.height, 4, world, rank)
    dense = torch.zeros((distributed.max_band_rows(cfg.height, 4, world), cfg.width, 4))
    dense[:len(rows)] = torch.from_numpy(img[rows])
    bufs = distributed.gather_bands(dist, dense, rank, world)
    # frame mode: sums of this rank's frames, reduced to rank 0
    first, count, stride = distributed.frame_share(0, 5, rank, world)
    total = torch.zeros((cfg.height, cfg.width, 4))
    for k in range(count):
        n = first + k * stride
        f, _, _ = oracle.render_pass(tris, boxes, indices, params.replace(frameNumber=n))
        # with an all-zero previous target the shader's running mean leaves curr / (n + 1): undo it
        total += torch.from_numpy(f) * float(n + 1)
    dist.reduce(total, dst=0)
    if rank == 0:
        frame = np.zeros((cfg.height, cfg.width, 4), dtype=np.float32)
        for g in range(world):
            r = distributed.band_row_indices(cfg.height, 4, world, g)
            frame[r] = bufs[g][:len(r)].numpy()
        np.save(os.path.join(out_dir, "tile.npy"), frame)
        np.save(os.path.join(out_dir, "sum.npy"), total.numpy())
    dist.destroy_process_group()


def test_two_rank_tile_and_frame_modes(tmp_path, oracle_mod):
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    from opengl_raytracer_b200.scene import Scene
    cfg = configs.scene_configs()["rects_box"].scaled(48, 30, 1)
    tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
    params = cfg.params()
    single, _, _ = oracle_mod.render_pass(tris, boxes, indices, params)
    assert np.load(tmp_path / "tile.npy").tobytes() == single.tobytes()          # tile mode: bit-identical
    mean = oracle_mod.render_frames(tris, boxes, indices, params, 0, 5)
    got = np.load(tmp_path / "sum.npy")[..., :3] / np.float32(5)
    assert np.abs(got - mean[..., :3]).max() < 2e-5                               # frame mode: re-association only


class _OracleRenderer:
    """Stand-in for rtlib.Context on the CPU: same method names, the oracle (test infrastructure) does the rendering,
    'device pointers' are host addresses.  Records the order of the calls so the sequencing can be checked."""

    torch_device = "cpu"

    def __init__(self, oracle, tris, boxes, indices, width, height):
        self.oracle, self.scene = oracle, (tris, boxes, indices)
        self.width, self.height = width, height
        self.image = np.zeros((height, width, 4), dtype=np.float32)
        self.params = None
        self.calls = []

    def set_params(self, p):
        self.params = p

    def _rows(self, tile):
        return distributed.band_row_indices(self.height, self.params.tileSizeY, self.params.numTilesY, tile)

    def render_pass(self, frame, tx, ty, flags=0):
        self.calls.append("render_pass")
        p = self.params.replace(frameNumber=frame, tileX=tx, tileY=ty)
        img, _, _ = self.oracle.render_pass(*self.scene, p, prev=self.image)
        rows = self._rows(ty)
        self.image[rows] = img[rows]                           # RT_PASS_SKIP_MASKED: only the owned rows are written
        if flags & 8:                                          # RT_PASS_TO_PEER
            self.calls[-1] = "render_pass(to_peer)"
            self.peer_image[rows] = img[rows]
            self.peer_image.flush()

    # peer tile mode: the 'IPC handle' is the name of a file both processes map
    accum_index = 0
    peer_image = None

    def peer_export(self):
        import tempfile
        fd, path = tempfile.mkstemp(suffix=".peer")
        os.close(fd)
        shared = np.memmap(path, dtype=np.float32, mode="w+", shape=self.image.shape)
        shared[:] = self.image
        self.image = shared
        return path.encode().ljust(128, b"\0")

    def peer_open(self, handles):
        assert len(handles) == 128
        self.peer_image = np.memmap(handles.rstrip(b"\0").decode(), dtype=np.float32, mode="r+", shape=self.image.shape)

    def peer_close(self):
        self.peer_image = None

    def target_wait_reads(self):
        self.calls.append("target_wait_reads")

    def _view(self, ptr, rows):
        import ctypes
        n = rows * self.width * 4
        return np.ctypeslib.as_array((ctypes.c_float * n).from_address(ptr)).reshape(rows, self.width, 4)

    def band_gather(self, tile, ptr):
        self.calls.append("band_gather")
        rows = self._rows(tile)
        self._view(ptr, len(rows))[:] = self.image[rows]

    def band_scatter(self, tile, ptr):
        self.calls.append("band_scatter")
        rows = self._rows(tile)
        self.image[rows] = self._view(ptr, len(rows))

    def sum_resume(self, prior):
        self.calls.append(f"sum_resume({prior})")
        self.image[..., :3] *= np.float32(prior)
        if prior == 0:
            self.image[:] = 0

    def render_frames_sum(self, first, count, stride, resume=False):
        self.calls.append(f"render_frames_sum(resume={resume})")
        if not resume:
            self.image[:] = 0
        for k in range(count):
            n = first + k * stride
            f, _, _ = self.oracle.render_pass(*self.scene, self.params.replace(frameNumber=n))
            # with an all-zero previous target the shader's running mean leaves curr / (n + 1): undo it
            self.image[..., :3] += (f * np.float32(n + 1))[..., :3]

    def finish_sum(self, total):
        self.calls.append(f"finish_sum({total})")
        self.image[..., :3] /= np.float32(total)
        self.image[..., 3] = 1


def _mode_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from opengl_raytracer_b200.scene import Scene
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = configs.scene_configs()["rects_box"].scaled(40, 27, 1)
    tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
    params = cfg.params()
    ctx = _OracleRenderer(oracle, tris, boxes, indices, cfg.width, cfg.height)
    distributed.render_tile_mode(ctx, dist, params, rank, world, frame=0, band_rows=4)
    if rank == 0:
        np.save(os.path.join(out_dir, "tile_mode.npy"), ctx.image)
        assert ctx.calls == ["render_pass", "band_gather", "band_scatter"], ctx.calls
    # tile mode with the peer exchange: two progressive frames, no gather; the fences are the group's all-reduces
    ctx = _OracleRenderer(oracle, tris, boxes, indices, cfg.width, cfg.height)
    peer = distributed.open_peer_targets(ctx, dist, rank, world)
    assert peer is not None
    for frame in (0, 1):
        distributed.render_tile_mode(ctx, dist, params, rank, world, frame=frame, band_rows=4, peer=peer)
    if rank == 0:
        np.save(os.path.join(out_dir, "tile_mode_peer.npy"), np.asarray(ctx.image))
        assert ctx.calls == ["target_wait_reads", "render_pass"] * 2, ctx.calls
    else:
        assert ctx.calls == ["render_pass(to_peer)"] * 2, ctx.calls
    peer.close()
    # a mapping refused on ONE rank (ranks on different nodes, no peer access): every rank learns it, nobody keeps a
    # half-open exchange, and the same frame then goes through the gather exchange
    ctx = _OracleRenderer(oracle, tris, boxes, indices, cfg.width, cfg.height)
    if rank == 1:
        def refuse(handles):
            raise RuntimeError("cudaIpcOpenMemHandle: invalid device context")
        ctx.peer_open = refuse
    assert distributed.open_peer_targets(ctx, dist, rank, world) is None
    assert ctx.peer_image is None
    distributed.render_tile_mode(ctx, dist, params, rank, world, frame=0, band_rows=4, peer=None)
    if rank == 0:
        np.save(os.path.join(out_dir, "tile_mode_fallback.npy"), np.asarray(ctx.image))
    # unequal targets are a caller's error, reported on every rank before anything is mapped
    ctx = _OracleRenderer(oracle, tris, boxes, indices, cfg.width, cfg.height + rank)
    try:
        distributed.open_peer_targets(ctx, dist, rank, world)
        raise AssertionError("unequal targets accepted")
    except RuntimeError as err:
        assert "lockstep" in str(err)
    # frame mode in two instalments: frames 0..2, then 3..6 continued from rank 0's running mean
    ctx = _OracleRenderer(oracle, tris, boxes, indices, cfg.width, cfg.height)
    view = lambda c: torch.from_numpy(c.image)
    distributed.render_frame_mode(ctx, dist, params, rank, world, 0, 3, view)
    distributed.render_frame_mode(ctx, dist, params, rank, world, 3, 4, view)
    if rank == 0:
        np.save(os.path.join(out_dir, "frame_mode.npy"), ctx.image)
        assert "sum_resume(3)" in ctx.calls and ctx.calls[-1] == "finish_sum(7)", ctx.calls
    else:
        assert not any(c.startswith("sum_resume") for c in ctx.calls), ctx.calls
    dist.destroy_process_group()


def test_render_modes_through_the_public_functions(tmp_path, oracle_mod):
    """render_tile_mode / render_frame_mode themselves (world 2, gloo), incl. a continued progressive render
    (first_frame > 0): the earlier mean is turned back into a sum on rank 0 and the division is by all frames."""
    port = 31500 + os.getpid() % 2000
    mp.spawn(_mode_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    from opengl_raytracer_b200.scene import Scene
    cfg = configs.scene_configs()["rects_box"].scaled(40, 27, 1)
    tris, boxes, indices = Scene(configs.build_objects(cfg.objects)).pack()
    params = cfg.params()
    single, _, _ = oracle_mod.render_pass(tris, boxes, indices, params)
    assert np.load(tmp_path / "tile_mode.npy").tobytes() == single.tobytes()
    assert np.load(tmp_path / "tile_mode_fallback.npy").tobytes() == single.tobytes()
    two = oracle_mod.render_frames(tris, boxes, indices, params, 0, 2)
    assert np.load(tmp_path / "tile_mode_peer.npy").tobytes() == two.tobytes()
    mean = oracle_mod.render_frames(tris, boxes, indices, params, 0, 7)
    got = np.load(tmp_path / "frame_mode.npy")
    assert np.abs(got[..., :3] - mean[..., :3]).max() < 2e-5 and (got[..., 3] == 1).all()
