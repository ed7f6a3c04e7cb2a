"""Multi-GPU host logic: one process per GPU, torch.distributed for the plumbing.

The path shards two ways, both taken from the reference's own mechanisms:

* **tile mode** (one frame, many GPUs) — the shader's tile mask (fragment.glsl:382) with ``numTilesX = 1``,
  ``numTilesY = world`` and ``tileSizeY = band_rows`` makes rank g own the interleaved row bands
  ``(y // band_rows) % world == g``: load-balanced, and each pixel is computed exactly as on one GPU, so the
  assembled frame is bit-identical.  Exchange step, two ways: **peer** (``open_peer_targets``; same node) — rank 0's
  targets are mapped into every rank through CUDA IPC and the render kernels store each finished pixel into them over
  NVLink as well as into their own, so the exchange rides on the pass and all that is left is a 4-byte all-reduce on
  either side of it as a fence; or **gather** — rank 0 gathers every rank's dense band buffer (``rt_band_gather`` ->
  NCCL gather -> ``rt_band_scatter``), which works across nodes too.
* **frame mode** (progressive accumulation) — frames are independent (the seed depends on px, py and
  frameNumber only, fragment.glsl:390): rank g renders frames g, g+world, ... as a running SUM
  (``rt_render_frames_sum``), the sums are added with an NCCL reduce and rank 0 divides by the frame count
  (``rt_finish_sum``).  Differs from the sequential running mean only by float re-association.

Samples inside one pixel-frame are never split: the RNG state is carried across them (fragment.glsl:358-363).
The functions take the renderer as an object with the rtlib.Context methods they call, so the CPU tests can
drive the same logic with gloo and a stand-in renderer.

Stream order: the renderer's kernels (pass, band pack / unpack, sum, finish) and torch's collectives must run in
program order.  ``render_tile_mode`` and ``render_frame_mode`` therefore move the context onto torch's current stream for
their duration (``rt_set_stream`` orders the new stream behind the work already queued on the old one, both ways) and
put it back afterwards — the caller needs no ``use_torch_stream()`` or ``sync()`` of its own.
"""
from __future__ import annotations

import contextlib

import numpy as np


@contextlib.contextmanager
def on_torch_stream(ctx):
    """Run ``ctx`` on torch's current CUDA stream inside the block (no-op for a renderer without streams, e.g. the
    CPU tests' stand-in), so collectives issued through torch.distributed order with the renderer's kernels."""
    switch = getattr(ctx, "use_torch_stream", None)
    if switch is None:
        yield
        return
    previous = switch()
    try:
        yield
    finally:
        ctx.set_stream(previous)


def band_params(params, rank: int, world: int, band_rows: int = 8):
    """Uniform block for rank ``rank`` of ``world`` in tile mode."""
    return params.replace(tileSizeX=params.width, numTilesX=1, tileX=0,
                          tileSizeY=band_rows, numTilesY=world, tileY=rank)


def band_row_indices(height: int, band_rows: int, world: int, rank: int) -> np.ndarray:
    """Rows y of the frame owned by ``rank``: (y // band_rows) % world == rank, ascending."""
    y = np.arange(height)
    return y[(y // band_rows) % world == rank]


def max_band_rows(height: int, band_rows: int, world: int) -> int:
    return max(len(band_row_indices(height, band_rows, world, r)) for r in range(world))


def frame_share(first_frame: int, n_frames: int, rank: int, world: int):
    """(first, count, stride) of the frames rank renders in frame mode."""
    count = (n_frames - rank + world - 1) // world if n_frames > rank else 0
    return first_frame + rank, count, world


def gather_bands(dist, local_dense, rank: int, world: int, dst: int = 0):
    """Gather equally sized (padded) dense band buffers to ``dst``; returns the list on dst, else None."""
    import torch
    if world == 1:
        return [local_dense]
    bufs = [torch.empty_like(local_dense) for _ in range(world)] if rank == dst else None
    dist.gather(local_dense, gather_list=bufs, dst=dst)
    return bufs


class PeerTargets:
    """Peer tile mode for one renderer and process group: rank ``dst``'s targets mapped into every other rank."""

    def __init__(self, ctx, dst: int, token):
        self.ctx, self.dst, self.token = ctx, dst, token

    def close(self):
        self.ctx.peer_close()


def open_peer_targets(ctx, dist, rank: int, world: int, dst: int = 0):
    """Collective: rank ``dst`` exports its targets (``rt_peer_export``), the others map them (``rt_peer_open``).
    All ranks must hold targets of one size and have flipped them equally often (checked here; keep issuing the same
    passes / clears on every rank afterwards).  Returns the ``PeerTargets`` to hand to ``render_tile_mode``, or None
    where the mapping is refused (different nodes, no peer access) on ANY rank — then use the gather exchange."""
    import torch
    state = [None] * world
    dist.all_gather_object(state, (ctx.width, ctx.height, ctx.accum_index))
    if len(set(state)) != 1:
        raise RuntimeError(f"peer tile mode needs equal targets in lockstep on every rank, got {state}")
    box = [ctx.peer_export() if rank == dst else None]
    dist.broadcast_object_list(box, src=dst)
    ok = True
    if rank != dst:
        try:
            ctx.peer_open(box[0])
        except Exception:                       # RtError: cudaIpcOpenMemHandle refused
            ok = False
    flags = [None] * world
    dist.all_gather_object(flags, ok)
    if not all(flags):
        ctx.peer_close()
        return None
    token = torch.zeros(1, dtype=torch.int32, device=getattr(ctx, "torch_device", "cpu"))
    return PeerTargets(ctx, dst, token)


def render_tile_mode(ctx, dist, params, rank: int, world: int, frame: int = 0, band_rows: int = 8, dense=None,
                     peer: PeerTargets | None = None):
    """One frame in tile mode.  ``ctx`` already holds the scene and a width x height target.  Returns the
    dense buffer used (reusable); on rank 0 the current target then holds the whole frame.  With ``peer`` (from
    ``open_peer_targets``) no dense buffer is used: the bands land in rank 0's target during the pass."""
    import torch
    p = band_params(params, rank, world, band_rows)
    ctx.set_params(p)
    if world == 1:
        ctx.render_pass(frame, 0, rank, flags=4)                # RT_PASS_SKIP_MASKED
        return dense
    if peer is not None:
        with on_torch_stream(ctx):
            if rank == peer.dst:
                ctx.target_wait_reads()                         # a read-back of the target about to be written
            dist.all_reduce(peer.token)                         # fence: rank dst is done clearing / reading it
            ctx.render_pass(frame, 0, rank, flags=4 | (8 if rank != peer.dst else 0))   # SKIP_MASKED | TO_PEER
            dist.all_reduce(peer.token)                         # fence: every rank's pass (and its stores) completed
        return dense
    with on_torch_stream(ctx):
        ctx.render_pass(frame, 0, rank, flags=4)
        rows = max_band_rows(params.height, band_rows, world)
        if dense is None:
            dense = torch.empty((rows, params.width, 4), dtype=torch.float32, device=getattr(ctx, "torch_device", "cuda"))
        ctx.band_gather(rank, dense.data_ptr())
        bufs = gather_bands(dist, dense, rank, world)
        if rank == 0:
            for g in range(1, world):
                ctx.band_scatter(g, bufs[g].data_ptr())
    return dense


def render_frame_mode(ctx, dist, params, rank: int, world: int, first_frame: int, n_frames: int, accum_tensor):
    """Progressive accumulation sharded by frames.  ``accum_tensor(ctx)`` returns a torch view of the current
    target; after the reduce rank 0's target holds the mean of frames 0 .. first_frame + n_frames - 1.

    ``first_frame > 0`` continues a render: rank 0's current target must hold the running mean of frames
    0 .. first_frame - 1 (left by an earlier call, by ``render_frames``, or restored from a checkpoint); it is turned
    back into a sum (``rt_sum_resume``) that rank 0's new frames are added to, the other ranks start from zero, and the
    final division is by ``first_frame + n_frames``.  ``params`` must describe the whole frame (numTilesX = numTilesY =
    1): ``rt_render_frames_sum`` refuses partial tiles."""
    first, count, stride = frame_share(first_frame, n_frames, rank, world)
    ctx.set_params(params)
    with on_torch_stream(ctx):
        resume = rank == 0 and first_frame > 0
        if resume:
            ctx.sum_resume(first_frame)
        if count:
            ctx.render_frames_sum(first, count, stride, resume=resume)
        view = accum_tensor(ctx)
        if count == 0 and not resume:
            view.zero_()
        if world > 1:
            dist.reduce(view, dst=0)
        if rank == 0:
            ctx.finish_sum(first_frame + n_frames)
