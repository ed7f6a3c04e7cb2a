"""Headless ``Screen`` — replaces the reference's ``screen.py``.

The reference's Screen owns a full-screen quad and two RGBA32F textures + FBOs used as a ping-pong
accumulation pair (``/root/reference/screen.py:37-66``).  There is no graphics stack here, so the pair
lives in device memory inside the rt context (``rt_target_create``); the attribute names the app loop uses
are kept: ``frame_count``, ``accum_index``, ``width``, ``height``, ``delete()``.  ``accum_tex`` /
``accum_fbo`` hold the two device pointers instead of GL object names.  Presentation (the blit to the
window, ``main.py:397-399``, and the RGBA8 ``glReadPixels`` of ``main.py:434``) becomes ``read_rgba8()``
and ``save_png()``; the saved image is rotated by 180 degrees like ``main.py:436`` when
``reference_orientation=True``.
"""
from __future__ import annotations

import numpy as np

from . import rtlib


class Screen:
    def __init__(self, width, height, sw=None, sh=None, context=None):
        self.width, self.height = int(width), int(height)
        self.sw, self.sh = sw or self.width, sh or self.height
        self.context = context if context is not None else rtlib.current_context()
        self.context.target_create(self.width, self.height)
        self.accum_tex = [self.context.accum_device_ptr(0), self.context.accum_device_ptr(1)]
        self.accum_fbo = list(self.accum_tex)
        self.frame_count = 0

    @property
    def accum_index(self):
        return self.context.accum_index

    @accum_index.setter
    def accum_index(self, value):
        self.context.accum_index = int(value)

    def clear(self):
        """``App.resetFrames`` (main.py:252-271): zero both targets, frame_count = accum_index = 0."""
        self.context.target_clear()
        self.frame_count = 0

    def read_accum(self) -> np.ndarray:
        """RGBA32F accumulation, shape (H, W, 4), row 0 = bottom row (GL orientation)."""
        return self.context.read_accum_f32()

    def read_rgba8(self) -> np.ndarray:
        return self.context.read_rgba8()

    def save_png(self, path, reference_orientation=False):
        """Write the current accumulation as an 8-bit PNG, top row first.  With
        ``reference_orientation`` the image is rotated by 180 degrees exactly like the reference's save
        path (``pg.image.fromstring`` of bottom-up rows, then ``rotate(180)``, main.py:434-436), which
        nets a horizontal mirror."""
        from PIL import Image

        rgba = self.read_rgba8()
        img = rgba[::-1, ::-1] if reference_orientation else rgba[::-1]
        Image.fromarray(np.ascontiguousarray(img), "RGBA").save(path)

    def delete(self):
        if self.context is not None and self.context.handle:
            self.context.target_free()
        self.context = None
