"""GPU tests (-m gpu) of peer tile mode: the bands of a frame written straight into the assembling rank's target through
CUDA IPC mappings by the render kernels themselves (include/rt_b200.h: rt_peer_export / rt_peer_open / RT_PASS_TO_PEER;
it stands for the one framebuffer every tile of a frame is drawn into, main.py:376-401 + fragment.glsl:382).

Two processes, as the library is deployed (one per rank).  With one GPU both use cuda:0 and the process group is gloo
(NCCL refuses two ranks on one device): the fences then synchronise the device before a host all-reduce.  With two or
more GPUs the same test also runs one rank per GPU over NCCL — the production path.  The assembled frames must be
bit-identical to the same frames rendered by one context alone."""
import os

import numpy as np
import pytest
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

W, H, SPP, FRAMES, BANDS = 320, 200, 4, 3, 8
SCENE = "rects_box"


class _HostFencedGroup:
    """torch.distributed over gloo for ranks that share a device: object collectives pass through, the tensor all-reduce
    used as a fence becomes 'finish the device work queued so far, then meet on the host'."""

    def __init__(self, dist):
        import torch
        self.dist, self.torch = dist, torch
        self.cpu_token = torch.zeros(1, dtype=torch.int32)

    def all_gather_object(self, out, obj):
        self.dist.all_gather_object(out, obj)

    def broadcast_object_list(self, box, src=0):
        self.dist.broadcast_object_list(box, src=src)

    def all_reduce(self, token):
        self.torch.cuda.synchronize()
        self.dist.all_reduce(self.cpu_token)


def _free_port():
    import socket
    with socket.socket() as sock:
        sock.bind(("127.0.0.1", 0))
        return sock.getsockname()[1]


def _scene():
    from opengl_raytracer_b200 import configs
    from opengl_raytracer_b200.scene import Scene
    cfg = configs.scene_configs()[SCENE].scaled(W, H, SPP)
    return Scene(configs.build_objects(cfg.objects)).pack() + (cfg.params(),)


def _worker(rank, world, port, out_dir, backend, variant):
    import torch
    import torch.distributed as dist
    from opengl_raytracer_b200 import distributed, rtlib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    device = rank if backend == "nccl" else 0
    torch.cuda.set_device(device)
    dist.init_process_group(backend, rank=rank, world_size=world)
    group = dist if backend == "nccl" else _HostFencedGroup(dist)
    tris, boxes, indices, params = _scene()
    ctx = rtlib.Context(device)
    ctx.set_variant(variant)
    ctx.scene_upload(tris, boxes, indices)
    ctx.target_create(W, H)
    peer = distributed.open_peer_targets(ctx, group, rank, world)
    assert peer is not None, "cudaIpcOpenMemHandle refused between two processes of one box"
    assert ctx.peer_is_open == (rank != 0)
    # rank 0 reads every frame back asynchronously while the next ones render: frame f + 2 is written into the target
    # frame f is being copied out of, by the peers too, so the fence before the pass must wait for that copy
    stage = [torch.empty((H, W, 4), dtype=torch.float32).pin_memory() for _ in range(FRAMES)] if rank == 0 else []
    for frame in range(FRAMES):
        distributed.render_tile_mode(ctx, group, params, rank, world, frame=frame, band_rows=BANDS, peer=peer)
        if rank == 0:
            ctx.read_accum_f32_async(stage[frame].data_ptr())
    if rank == 0:
        ctx.read_wait()
        np.save(os.path.join(out_dir, "frames.npy"), np.stack([t.numpy() for t in stage]))
        np.save(os.path.join(out_dir, "peer.npy"), ctx.read_accum_f32())
    else:
        # a pass that asks for the peer copy after the mapping is gone is refused, not silently local
        peer.close()
        with pytest.raises(RuntimeError, match="rt_peer_open"):
            ctx.render_pass(0, 0, rank, flags=rtlib.PASS_SKIP_MASKED | rtlib.PASS_TO_PEER)
    group.all_reduce(peer.token)          # rank 0's mapping source outlives the peers' use of it
    ctx.close()
    dist.destroy_process_group()


def _single(variant):
    from opengl_raytracer_b200 import rtlib
    tris, boxes, indices, params = _scene()
    ctx = rtlib.Context(0)
    ctx.set_variant(variant)
    ctx.scene_upload(tris, boxes, indices)
    ctx.set_params(params)
    ctx.target_create(W, H)
    frames = []
    for frame in range(FRAMES):
        ctx.render_pass(frame, 0, 0)
        frames.append(ctx.read_accum_f32())
    ctx.close()
    return np.stack(frames)


@pytest.mark.parametrize("variant", ["auto", "pixel", "persistent", "wavefront", "pool", "samples", "chain"])
def test_peer_tile_mode_two_processes_one_device(tmp_path, variant):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path), "gloo", variant), nprocs=2, join=True)
    single = _single(variant)
    assert np.load(tmp_path / "peer.npy").tobytes() == single[-1].tobytes()
    assert np.load(tmp_path / "frames.npy").tobytes() == single.tobytes()      # each frame as read back on the way


def test_peer_tile_mode_over_nccl(tmp_path):
    import torch
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("one GPU: the NCCL leg needs a rank per device")
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path), "nccl", "auto"), nprocs=world, join=True)
    single = _single("auto")
    assert np.load(tmp_path / "peer.npy").tobytes() == single[-1].tobytes()
    assert np.load(tmp_path / "frames.npy").tobytes() == single.tobytes()
