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
