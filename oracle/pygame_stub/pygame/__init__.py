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
        # [H, W, 4] float64; untouched pixels stay (0,0,0,0) like a fresh SRCALPHA surface
        self.f64 = np.zeros((self.size[1], self.size[0], 4), dtype=np.float64)
        LAST_SURFACE = self

    def set_at(self, pos, color):
        self.f64[pos[1], pos[0], :] = color


class image:
    @staticmethod
    def tobytes(surface, fmt):
        assert fmt == "RGBA"
        return np.clip(np.trunc(surface.f64), 0, 255).astype(np.uint8).tobytes()


class time:
    @staticmethod
    def get_ticks():
        return 0
