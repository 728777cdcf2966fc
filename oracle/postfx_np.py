"""numpy restatement of the image-space steps after the trace loop.  TEST INFRASTRUCTURE ONLY.

Follows Window.draw_tile (reference init.py:185-190) and Window.draw (init.py:207-254) statement by statement.  The
reference performs every step with a pygame call; pygame is not part of /root/reference and cannot be installed
here, so **parity of this file is UNPINNED**: the per-pixel formulas are pygame 2.x's documented ones (surface.h
ALPHA_BLEND and BLEND_RGBA_* macros, transform.average_color(consider_alpha=True), transform.grayscale's luminosity
weights); transform.smoothscale is replaced by an area average (shrinking) / 8.8 fixed-point bilinear (enlarging).
The CUDA kernels (csrc/vt_postfx.cuh) are checked against this file bit for bit.
"""
from __future__ import annotations

import numpy as np


def _mult(d, s):
    return np.where((d != 0) & (s != 0), (d * s + 255) >> 8, 0)


def _add(d, s):
    return np.minimum(255, d + s)


def _sub(d, s):
    return np.maximum(0, d - s)


def compose(canvas, tile):
    """canvas.blit(tile) on SRCALPHA surfaces (init.py:187): pygame ALPHA_BLEND."""
    d = canvas.astype(np.int64)
    s = tile.astype(np.int64)
    sa, da = s[..., 3:4], d[..., 3:4]
    rgb = ((d[..., :3] << 8) + (s[..., :3] - d[..., :3]) * sa + s[..., :3]) >> 8
    a = sa + da - (sa * da) // 255
    blended = np.concatenate([rgb, a], axis=-1)
    out = np.where(da != 0, blended, s)
    out = np.where(sa != 0, out, d)
    return out.astype(np.uint8)


def average_color(canvas):
    """pg.transform.average_color(canvas, consider_alpha=True) (init.py:211)."""
    c = canvas.reshape(-1, 4).astype(np.int64)
    at = int(c[:, 3].sum())
    rgb = [int((c[:, k] * c[:, 3]).sum()) // at if at else 0 for k in range(3)]
    return rgb + [at // len(c)]


def shrink(src, dw, dh):
    sh, sw = src.shape[:2]
    out = np.zeros((dh, dw, 4), np.uint8)
    s = src.astype(np.int64)
    for y in range(dh):
        y0, y1 = y * sh // dh, (y + 1) * sh // dh
        y1 = max(y1, y0 + 1)
        for x in range(dw):
            x0, x1 = x * sw // dw, (x + 1) * sw // dw
            x1 = max(x1, x0 + 1)
            blk = s[y0:y1, x0:x1].reshape(-1, 4)
            out[y, x] = blk.sum(axis=0) // len(blk)
    return out


def resample(src, dw, dh, smooth):
    sh, sw = src.shape[:2]
    s = src.astype(np.int64)
    ys, xs = np.arange(dh, dtype=np.int64), np.arange(dw, dtype=np.int64)
    if not smooth:
        return src[(ys * sh // dh)[:, None], (xs * sw // dw)[None, :]].copy()
    sx = ((sw - 1) << 16) // (dw - 1) if dw > 1 else 0
    sy = ((sh - 1) << 16) // (dh - 1) if dh > 1 else 0
    px, py = xs * sx, ys * sy
    x0, y0 = px >> 16, py >> 16
    fx, fy = ((px & 0xffff) >> 8)[None, :, None], ((py & 0xffff) >> 8)[:, None, None]
    x1, y1 = np.minimum(x0 + 1, sw - 1), np.minimum(y0 + 1, sh - 1)
    p00, p01 = s[y0[:, None], x0[None, :]], s[y0[:, None], x1[None, :]]
    p10, p11 = s[y1[:, None], x0[None, :]], s[y1[:, None], x1[None, :]]
    top = p00 * (256 - fx) + p01 * fx
    bot = p10 * (256 - fx) + p11 * fx
    return ((top * (256 - fy) + bot * fy) >> 16).astype(np.uint8)


def scale(src, dw, dh, smooth):
    sh, sw = src.shape[:2]
    if smooth and dw < sw and dh < sh:
        return shrink(src, dw, dh)
    return resample(src, dw, dh, smooth)


def draw(canvas, width, height, out_width, out_height, spill=0.0, iris_on=False, iris=0.0, bloom=0.0, bloom_blur=0.0,
         subsamples=0.0, smooth=0.0):
    """Window.draw's canvas processing (init.py:209-254).  Returns (image [out_h, out_w, 4] u8, average colour)."""
    color = average_color(canvas)
    c = canvas.astype(np.int64)
    if spill:                                                   # init.py:213-216
        fac = 255 - round(spill * 255)
        tint = np.array([min(255, v + fac) for v in color], np.int64)
        c = _mult(c, tint)
    if iris_on:                                                 # init.py:219-228
        add = iris > 0
        col = 0 if add else 255
        fac = round(abs(iris * 255))
        gray = (0.299 * c[..., 0] + 0.587 * c[..., 1] + 0.114 * c[..., 2]).astype(np.int64)
        g4 = np.stack([gray, gray, gray, c[..., 3]], axis=-1)
        mask = _add(col, g4) if add else _sub(col, g4)
        mask = _mult(mask, fac)
        c = _add(c, mask) if add else _sub(c, mask)
    if bloom and bloom_blur:                                    # init.py:232-239
        bw = max(1, round(width / max(1, bloom_blur)))
        bh = max(1, round(height / max(1, bloom_blur)))
        fac = round((1 - bloom) * 255)
        dark = _sub(c, np.array([fac, fac, fac, 255], np.int64)).astype(np.uint8)
        blur = scale(dark, bw, bh, True)
        blur = resample(blur, width, height, True).astype(np.int64)
        c = _add(c, blur)
    img = c.astype(np.uint8)
    if subsamples:                                              # init.py:242-244
        f = 1 + subsamples
        img = scale(img, int(width * f), int(height * f), True)
    if smooth == 0:                                             # init.py:247-254
        img = scale(img, out_width, out_height, False)
    elif smooth == 1:
        img = scale(img, out_width, out_height, True)
    else:
        f = max(1, int(1 / smooth))
        img = scale(img, img.shape[1] * f, img.shape[0] * f, False)
        img = scale(img, out_width, out_height, True)
    return img, color
