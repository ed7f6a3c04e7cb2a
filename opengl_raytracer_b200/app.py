"""Headless ``App`` — the reference's ``main.py`` driver without a window.

``App(window_size, screen_size, bounces, rays_per_pixel, jitter_amount, lambertian, skyIllumination, tileSize)``
keeps the constructor of ``/root/reference/main.py:16`` and the sequencing of its loop:

* uniforms derived as ``main.py:125-137,166-202`` (``configs.make_params``; note ``nBounces = bounces + 1``);
* one *pass* per loop iteration = one tile (``main.py:376-394``), tiles advance x-fastest and
  ``frame_count`` increments after the last tile (``main.py:409-430``);
* ``resetFrames()`` = ``main.py:252-271`` (both targets cleared, ``frame_count = accum_index = 0``) and is
  called, like there, whenever the camera or ``lambertian`` changes;
* ``save()`` = the RGBA8 read-back + PNG of ``main.py:432-439`` (without the 10-minute condition).

What a window provided — events, the blit, the caption — has no counterpart here; the scene is passed in (or one
of the named BASELINE.json scenes) instead of being hard-coded as in ``main.py:19-111``.  New: the accumulation
can be checkpointed and resumed (``save_checkpoint`` / ``load_checkpoint``), which the reference cannot do
(its accumulation lives in GL textures only).

CLI:  python -m opengl_raytracer_b200.app --scene cfg2_box --width 1280 --height 720 --frames 16 --out render.png
"""
from __future__ import annotations

import argparse
import time

import numpy as np

from . import configs, rtlib
from .scene import Scene
from .screen import Screen


class App:
    def __init__(self, window_size, screen_size=None, bounces=7, rays_per_pixel=1, jitter_amount=0.001,
                 lambertian=True, skyIllumination=1, tileSize=1, objects=None, cam_pos=(-33.7, 14.8, -21.1),
                 cam_dir=(65, -25.4), device=0, variant="auto", verbose=False):
        self.window_size = (int(window_size[0]), int(window_size[1]))
        self.screen_size = tuple(int(x) for x in (screen_size if screen_size is not None else window_size))
        self.bounces, self.rays_per_pixel = int(bounces), int(rays_per_pixel)
        self.jitter_amount, self.lambertian = float(jitter_amount), bool(lambertian)
        self.skyIllumination, self.tileSize = float(skyIllumination), int(tileSize)
        self.camPos, self.camDir = tuple(cam_pos), tuple(cam_dir)
        self.context = rtlib.Context(device)
        self.context.set_variant(variant)
        if objects is None:                                   # main.py:19-111 with the missing dragon replaced
            objects = configs.build_objects(configs.scene_configs()["cfg2_box"].objects)
        self.scene = Scene(objects, verbose=verbose) if not isinstance(objects, Scene) else objects
        self.scene.send(self.context)                         # main.py:149
        w, h = self.window_size
        self.screen = Screen(w, h, *self.screen_size, context=self.context)   # main.py:161
        self.tileX = self.tileY = 0
        self.total_frames = 0                                 # counts draws, like main.py:290,430
        self.render_seconds = 0.0
        self._upload_uniforms()

    # -- uniforms (main.py:166-202) ---------------------------------------------------------------------------
    def _upload_uniforms(self):
        w, h = self.window_size
        sw, sh = self.screen_size
        self.params = configs.make_params(w, h, self.bounces, self.rays_per_pixel, self.camPos, self.camDir,
                                          jitter_amount=self.jitter_amount, lambertian=self.lambertian,
                                          sky_brightness=self.skyIllumination, aspect=sw / sh, tile_size=self.tileSize)
        self.context.set_params(self.params)

    def set_camera(self, cam_pos=None, cam_dir=None):
        """Camera change: new uniforms and ``resetFrames()`` (main.py:326-337)."""
        if cam_pos is not None:
            self.camPos = tuple(cam_pos)
        if cam_dir is not None:
            self.camDir = tuple(cam_dir)
        self._upload_uniforms()
        self.resetFrames()

    def set_lambertian(self, value):
        """Key L of the reference (main.py:352-360)."""
        self.lambertian = bool(value)
        self._upload_uniforms()
        self.resetFrames()

    def resetFrames(self):
        self.screen.clear()
        self.tileX = self.tileY = 0

    # -- the loop (main.py:289-430) ------------------------------------------------------------------------------
    def step(self):
        """One loop iteration: draw the current tile, advance tileX/tileY, count the frame after the last tile."""
        self.context.render_pass(self.screen.frame_count, self.tileX, self.tileY)
        self.total_frames += 1
        self.tileX += 1
        if self.tileX >= self.params.numTilesX:
            self.tileX = 0
            self.tileY += 1
        if self.tileY >= self.params.numTilesY:
            self.tileY = 0
            self.screen.frame_count += 1

    def main(self, frames=1):
        """Render ``frames`` whole frames (every tile of each); returns the seconds spent on the device."""
        self.context.timer_start()
        passes = frames * self.params.numTilesX * self.params.numTilesY - (self.tileY * self.params.numTilesX + self.tileX)
        for _ in range(max(passes, 0)):
            self.step()
        seconds = self.context.timer_stop() * 1e-3
        self.render_seconds += seconds
        return seconds

    # -- outputs -------------------------------------------------------------------------------------------------
    def save(self, path, reference_orientation=False):
        self.screen.save_png(path, reference_orientation=reference_orientation)

    def save_checkpoint(self, path):
        """Accumulation + loop position: enough to continue the progressive render bit-identically."""
        np.savez_compressed(path, accum=self.screen.read_accum(), frame_count=self.screen.frame_count,
                            tile=np.array([self.tileX, self.tileY]), total_frames=self.total_frames,
                            window_size=np.array(self.window_size))

    def load_checkpoint(self, path):
        with np.load(path) as z:
            if tuple(z["window_size"]) != self.window_size:
                raise ValueError(f"checkpoint is {tuple(z['window_size'])}, this app renders {self.window_size}")
            self.context.write_accum_f32(np.ascontiguousarray(z["accum"], dtype=np.float32))
            self.screen.frame_count = int(z["frame_count"])
            self.tileX, self.tileY = (int(v) for v in z["tile"])
            self.total_frames = int(z["total_frames"])

    def close(self):
        """main.py:440-445."""
        self.screen.delete()
        self.scene.clearMemory()
        self.context.close()


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--scene", default="cfg2_box", choices=sorted(configs.scene_configs()))
    ap.add_argument("--width", type=int)
    ap.add_argument("--height", type=int)
    ap.add_argument("--spp", type=int, help="rays_per_pixel")
    ap.add_argument("--bounces", type=int)
    ap.add_argument("--frames", type=int, default=1)
    ap.add_argument("--tile-size", type=int, default=1)
    ap.add_argument("--variant", default="auto", choices=sorted(rtlib.VARIANTS))
    ap.add_argument("--out", default="render.png")
    ap.add_argument("--resume", help="checkpoint (.npz) to continue from")
    ap.add_argument("--checkpoint", help="write a checkpoint (.npz) at the end")
    args = ap.parse_args(argv)
    cfg = configs.scene_configs()[args.scene]
    size = (args.width or cfg.width, args.height or cfg.height)
    started = time.time()
    app = App(size, size, args.bounces if args.bounces is not None else cfg.bounces,
              args.spp if args.spp is not None else cfg.rays_per_pixel, objects=configs.build_objects(cfg.objects),
              cam_pos=cfg.cam_pos, cam_dir=cfg.cam_dir, tileSize=args.tile_size, variant=args.variant)
    if args.resume:
        app.load_checkpoint(args.resume)
    seconds = app.main(args.frames)
    app.save(args.out)
    if args.checkpoint:
        app.save_checkpoint(args.checkpoint)
    print(f"{args.scene} {size[0]}x{size[1]}: {args.frames} frame(s) in {seconds * 1e3:.1f} ms on the device "
          f"({time.time() - started:.1f} s with scene build), frame_count = {app.screen.frame_count} -> {args.out}")
    app.close()


if __name__ == "__main__":
    main()
