// vt_postfx.cuh — the image-space steps right after the trace loop, included at the end of
// voxtrace.cu: tile compositing (Window.draw_tile, reference init.py:185-190) and the canvas effects of Window.draw
// (init.py:207-253: colour spill, iris adaptation, bloom, subsampling / window scaling).
//
// The reference does these with pygame surface operations; pygame's C code is not part of /root/reference and is
// not installable here, so its per-pixel rounding is UNPINNED.  The integer formulas below follow pygame 2.x's
// documented blend macros (surface.h: ALPHA_BLEND, BLEND_RGBA_ADD / SUB / MULT), transform.average_color with
// consider_alpha, transform.grayscale's luminosity weights, and replace transform.smoothscale by an area average
// (shrinking) and a fixed-point bilinear filter (enlarging).  oracle/postfx_np.py restates exactly these formulas
// in numpy; the GPU tests compare against it bit for bit.
#pragma once

namespace vtfx {

__device__ __forceinline__ int blend_mult(int d, int s) { return (d && s) ? ((d * s) + 255) >> 8 : 0; }   // BLEND_RGBA_MULT
__device__ __forceinline__ int blend_add(int d, int s) { int v = d + s; return v > 255 ? 255 : v; }          // BLEND_RGBA_ADD
__device__ __forceinline__ int blend_sub(int d, int s) { int v = d - s; return v < 0 ? 0 : v; }              // BLEND_RGBA_SUB

// Window.draw_tile: canvas.blit(tile) with per-pixel alpha on a SRCALPHA canvas (ALPHA_BLEND).  The tile's alpha is
// round(min(1, energy + shutter) * 255) (init.py:141): the motion-blur weight.
__global__ void compose_kernel(uchar4 *__restrict__ canvas, const uchar4 *__restrict__ tile, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uchar4 s = tile[i];
        uchar4 d = canvas[i];
        if (s.w == 0) continue;                      // pixels the tile did not render stay untouched (alpha 0)
        if (d.w) {
            d.x = (unsigned char)((((int)d.x << 8) + ((int)s.x - (int)d.x) * (int)s.w + (int)s.x) >> 8);
            d.y = (unsigned char)((((int)d.y << 8) + ((int)s.y - (int)d.y) * (int)s.w + (int)s.y) >> 8);
            d.z = (unsigned char)((((int)d.z << 8) + ((int)s.z - (int)d.z) * (int)s.w + (int)s.z) >> 8);
            d.w = (unsigned char)((int)s.w + (int)d.w - (((int)s.w * (int)d.w) / 255));
        } else d = s;
        canvas[i] = d;
    }
}

// transform.average_color(canvas, consider_alpha=True): sums of r*a, g*a, b*a, a  -> sums[4]
__global__ void average_kernel(const uchar4 *__restrict__ canvas, long long n, unsigned long long *sums) {
    unsigned long long r = 0, g = 0, b = 0, a = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uchar4 p = canvas[i];
        r += (unsigned)p.x * p.w; g += (unsigned)p.y * p.w; b += (unsigned)p.z * p.w; a += p.w;
    }
    for (int o = 16; o; o >>= 1) {
        r += __shfl_xor_sync(0xffffffffu, r, o); g += __shfl_xor_sync(0xffffffffu, g, o);
        b += __shfl_xor_sync(0xffffffffu, b, o); a += __shfl_xor_sync(0xffffffffu, a, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(sums + 0, r); atomicAdd(sums + 1, g); atomicAdd(sums + 2, b); atomicAdd(sums + 3, a); }
}

struct FxParams {
    int spill_fac;        // 255 - round(spill * 255), or -1 when spill is off (init.py:213-216)
    int iris_on;          // settings.iris and settings.iris_time (init.py:219)
    int iris_add;         // 1: self.iris > 0 -> BLEND_RGBA_ADD with col 0, 0: BLEND_RGBA_SUB with col 255 (init.py:220-221)
    int iris_fac;         // round(abs(self.iris * 255)) (init.py:222)
    int bloom_fac;        // round((1 - bloom) * 255), or -1 when bloom is off (init.py:232-236)
};

// average colour (r, g, b, a) as pygame returns it from the sums: r = sum(r*a) / sum(a), a = sum(a) / pixels
__device__ __forceinline__ void average_from_sums(const unsigned long long *sums, long long n, int avg[4]) {
    const unsigned long long at = sums[3];
    avg[0] = at ? (int)(sums[0] / at) : 0; avg[1] = at ? (int)(sums[1] / at) : 0; avg[2] = at ? (int)(sums[2] / at) : 0;
    avg[3] = (int)(at / (unsigned long long)n);
}

// colour spill + iris on one pixel (init.py:211-229); writes the canvas after both and, for bloom, the darkened copy
__global__ void spill_iris_kernel(const uchar4 *__restrict__ canvas, long long n, const unsigned long long *sums, FxParams fx,
                                  uchar4 *__restrict__ out, uchar4 *__restrict__ dark, int *avg_out) {
    int avg[4];
    average_from_sums(sums, n, avg);
    if (blockIdx.x == 0 && threadIdx.x == 0 && avg_out) { avg_out[0] = avg[0]; avg_out[1] = avg[1]; avg_out[2] = avg[2]; avg_out[3] = avg[3]; }
    int tint[4];
    for (int k = 0; k < 4; k++) { int t = avg[k] + fx.spill_fac; tint[k] = t > 255 ? 255 : t; }
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uchar4 p = canvas[i];
        int c[4] = {p.x, p.y, p.z, p.w};
        if (fx.spill_fac >= 0)
            for (int k = 0; k < 4; k++) c[k] = blend_mult(c[k], tint[k]);
        if (fx.iris_on) {
            // transform.grayscale keeps alpha; mask = fill(col) (+|-) gray, then * fac, then canvas (+|-) mask
            const int gray = (int)(0.299 * c[0] + 0.587 * c[1] + 0.114 * c[2]);
            const int g4[4] = {gray, gray, gray, c[3]};
            for (int k = 0; k < 4; k++) {
                int m = fx.iris_add ? blend_add(0, g4[k]) : blend_sub(255, g4[k]);
                m = blend_mult(m, fx.iris_fac);
                c[k] = fx.iris_add ? blend_add(c[k], m) : blend_sub(c[k], m);
            }
        }
        out[i] = make_uchar4((unsigned char)c[0], (unsigned char)c[1], (unsigned char)c[2], (unsigned char)c[3]);
        if (dark) {
            // canvas_blur.fill((fac, fac, fac), BLEND_RGBA_SUB): a 3-tuple colour has alpha 255 (init.py:235)
            dark[i] = make_uchar4((unsigned char)blend_sub(c[0], fx.bloom_fac), (unsigned char)blend_sub(c[1], fx.bloom_fac),
                                  (unsigned char)blend_sub(c[2], fx.bloom_fac), (unsigned char)blend_sub(c[3], 255));
        }
    }
}

// area average: dst pixel (x, y) = floor mean of the source pixels [x*sw/dw, (x+1)*sw/dw) x [y*sh/dh, (y+1)*sh/dh)
__global__ void shrink_kernel(const uchar4 *__restrict__ src, int sw, int sh, uchar4 *__restrict__ dst, int dw, int dh) {
    const long long n = (long long)dw * dh;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % dw), y = (int)(i / dw);
        int x0 = (int)((long long)x * sw / dw), x1 = (int)((long long)(x + 1) * sw / dw);
        int y0 = (int)((long long)y * sh / dh), y1 = (int)((long long)(y + 1) * sh / dh);
        if (x1 <= x0) x1 = x0 + 1;
        if (y1 <= y0) y1 = y0 + 1;
        unsigned r = 0, g = 0, b = 0, a = 0;
        for (int yy = y0; yy < y1; yy++)
            for (int xx = x0; xx < x1; xx++) {
                const uchar4 p = src[(size_t)yy * sw + xx];
                r += p.x; g += p.y; b += p.z; a += p.w;
            }
        const unsigned cnt = (unsigned)((x1 - x0) * (y1 - y0));
        dst[i] = make_uchar4((unsigned char)(r / cnt), (unsigned char)(g / cnt), (unsigned char)(b / cnt), (unsigned char)(a / cnt));
    }
}

// enlarging / resampling: nearest (transform.scale) or 8.8 fixed-point bilinear; `add` != NULL blends the result onto
// `add` with BLEND_RGBA_ADD (the bloom blit, init.py:239) instead of storing it
__global__ void resample_kernel(const uchar4 *__restrict__ src, int sw, int sh, uchar4 *__restrict__ dst, int dw, int dh, int smooth,
                                const uchar4 *__restrict__ add) {
    const long long n = (long long)dw * dh;
    const long long sx = dw > 1 ? (((long long)(sw - 1)) << 16) / (dw - 1) : 0, sy = dh > 1 ? (((long long)(sh - 1)) << 16) / (dh - 1) : 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % dw), y = (int)(i / dw);
        int c[4];
        if (!smooth) {
            const uchar4 p = src[(size_t)((long long)y * sh / dh) * sw + (size_t)((long long)x * sw / dw)];
            c[0] = p.x; c[1] = p.y; c[2] = p.z; c[3] = p.w;
        } else {
            const long long px = x * sx, py = y * sy;
            const int x0 = (int)(px >> 16), y0 = (int)(py >> 16), fx = (int)((px & 0xffff) >> 8), fy = (int)((py & 0xffff) >> 8);
            const int x1 = x0 + 1 < sw ? x0 + 1 : x0, y1 = y0 + 1 < sh ? y0 + 1 : y0;
            const uchar4 p00 = src[(size_t)y0 * sw + x0], p01 = src[(size_t)y0 * sw + x1], p10 = src[(size_t)y1 * sw + x0], p11 = src[(size_t)y1 * sw + x1];
            const int a00[4] = {p00.x, p00.y, p00.z, p00.w}, a01[4] = {p01.x, p01.y, p01.z, p01.w};
            const int a10[4] = {p10.x, p10.y, p10.z, p10.w}, a11[4] = {p11.x, p11.y, p11.z, p11.w};
            for (int k = 0; k < 4; k++) {
                const int top = a00[k] * (256 - fx) + a01[k] * fx, bot = a10[k] * (256 - fx) + a11[k] * fx;
                c[k] = (top * (256 - fy) + bot * fy) >> 16;
            }
        }
        if (add) {
            const uchar4 d = add[i];
            c[0] = blend_add(d.x, c[0]); c[1] = blend_add(d.y, c[1]); c[2] = blend_add(d.z, c[2]); c[3] = blend_add(d.w, c[3]);
        }
        dst[i] = make_uchar4((unsigned char)c[0], (unsigned char)c[1], (unsigned char)c[2], (unsigned char)c[3]);
    }
}

}  // namespace vtfx
