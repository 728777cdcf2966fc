"""Canvas compositing and post-FX on the device: what ``Window.draw_tile`` / ``Window.draw`` do with pygame surfaces
after the trace loop (reference init.py:185-190, 207-254).  Thin ctypes layer over ``vt_fx_compose`` / ``vt_fx_draw``.
The text overlay and ``display.flip`` need a window and are not part of it.  pygame's rounding is unpinned
(DESIGN.md §2); the formulas are stated in csrc/vt_postfx.cuh.
"""
from __future__ import annotations

import ctypes as C

import torch

from .abi import VtFx


class Canvas:
    """The window's SRCALPHA canvas (init.py:159) as a CUDA tensor [H, W, 4] u8, plus the iris state (init.py:167)."""

    def __init__(self, cam, scale=1):
        self.cam = cam
        s = cam.settings
        self.w, self.h = int(s.width), int(s.height)
        self.out_w, self.out_h = self.w * int(scale), self.h * int(scale)          # data.py:65
        self.canvas = torch.zeros((self.h, self.w, 4), dtype=torch.uint8, device=cam.device)
        self.out = torch.empty((self.out_h, self.out_w, 4), dtype=torch.uint8, device=cam.device)
        self.iris = 0.0
        self.iris_target = 0.0

    def draw_tile(self, tile_rgba8):
        """init.py:185-187: alpha-blit a tile (CUDA tensor [H, W, 4] u8; alpha 0 = pixel not rendered)."""
        cam = self.cam
        cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
        t = tile_rgba8.contiguous()
        cam._check(cam._lib.vt_fx_compose(cam._ctx, self.canvas.data_ptr(), t.data_ptr(), self.w * self.h))

    def draw(self, spill=0.0, iris=0.0, iris_time=0.0, bloom=0.0, bloom_blur=0.0, subsamples=0.0, smooth=0.0):
        """init.py:207-254 up to (not including) the text overlay; returns the window image [out_h, out_w, 4] and
        updates ``iris_target`` (init.py:229).  The caller advances ``iris`` like init.py:462."""
        cam = self.cam
        fx = VtFx()
        fx.width, fx.height, fx.out_width, fx.out_height = self.w, self.h, self.out_w, self.out_h
        fx.iris_on = 1 if (iris and iris_time) else 0
        fx.spill, fx.iris, fx.bloom, fx.bloom_blur = float(spill), float(self.iris), float(bloom), float(bloom_blur)
        fx.subsamples, fx.smooth = float(subsamples), float(smooth)
        avg = (C.c_int32 * 4)()
        cam._check(cam._lib.vt_set_stream(cam._ctx, torch.cuda.current_stream(cam.device).cuda_stream))
        cam._check(cam._lib.vt_fx_draw(cam._ctx, C.byref(fx), self.canvas.data_ptr(), self.out.data_ptr(), avg))
        self.average = [int(v) for v in avg]
        if fx.iris_on:
            self.iris_target = 1 - (max(self.average[:3]) / 255) * 2
        return self.out
