"""Minimal stand-in for pygame, used ONLY by the oracle harness (test infrastructure).

The reference hot path (init.py:126-150) uses pygame purely as an RGBA pixel container:
``pg.Surface(size, pg.SRCALPHA)``, ``Surface.set_at``, ``pg.image.tobytes`` — plus
``pg.time.get_ticks`` in sprite animation (data.py:306).  pygame is not installed in this image
and there is no display, so this stub provides just those names.

``set_at`` receives *float* channel means (lib.py:419-428 -> init.py:145-146).  Real pygame
quantises them to u8 in C code that is not under /root/reference ("parity unpinned" at that one
conversion); the stub keeps the floats (``Surface.f64``) and truncates in
``tobytes`` (documented choice, same as the CUDA epilogue).
"""
import numpy as np

SRCALPHA = 0x00010000
LAST_SURFACE = None  # the harness reads the float buffer of the most recent surface


class Surface:
    def __init__(self, size, flags=0):
        global LAST_SURFACE
        self.size = (int(size[0]), int(size[1]))
        self.px = {}                # (x, y) -> the float RGBA handed to set_at; untouched pixels are (0, 0, 0, 0)
        LAST_SURFACE = self

    def set_at(self, pos, color):
        self.px[(int(pos[0]), int(pos[1]))] = tuple(color)

    @property
    def f64(self):
        """[H, W, 4] float64 view of the surface (built on demand: a worker that renders a few thousand pixels of a
        1080p window must not pay for a 66 MB array per tile, which real pygame does not either)."""
        a = np.zeros((self.size[1], self.size[0], 4), dtype=np.float64)
        for (x, y), c in self.px.items():
            a[y, x, :] = c
        return a


class image:
    @staticmethod
    def tobytes(surface, fmt):
        assert fmt == "RGBA"
        a = np.zeros((surface.size[1], surface.size[0], 4), dtype=np.uint8)      # fresh SRCALPHA surface: all zero
        for (x, y), c in surface.px.items():
            a[y, x, :] = np.clip(np.trunc(np.asarray(c, dtype=np.float64)), 0, 255).astype(np.uint8)
        return a.tobytes()


class time:
    @staticmethod
    def get_ticks():
        return 0
