// vt_device.cuh — device-side ray march + shading for sm_100a.
//
// Restates, as compiled device functions, what the reference does per ray in Python:
//   Camera.trace (reference init.py:37-121), Frame.get_voxel (data.py:136-145), lib.material
//   (lib.py:448-460), lib.material_background (lib.py:463-476) and the vec3/quaternion arithmetic they
//   use (lib.py:166-376).  All ray state is fp64 like Python floats and the file is compiled with
//   -fmad=false so that every a*b+c rounds twice, as the interpreter does.
//
// Sampling rule: positions are p_{k+1} = p_k + vel*step accumulated in fp64 and only
// floor(p_k) is tested -- a fixed-increment DDA, not a cell-exact Amanatides-Woo walk.  Brick-level
// reasoning is used only to avoid memory traffic, never to choose sample positions.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/voxtrace.h"

#define VT_LOD_TAB 16
#define VT_POW_BITS 16
#define VT_POW_SLOTS (1 << VT_POW_BITS)               // Frame.resolution values served by the multiply-high snap table

namespace vt {

struct DevWorld {
    int ox, oy, oz;                 // world cell of the minimum corner (multiple of chunk_size)
    int nx, ny, nz;                 // cells per axis (multiple of chunk_size)
    int nbx, nby, nbz;              // 8^3 bricks per axis (ceil)
    int ncx, ncy, ncz;              // chunk slots per axis
    const uint32_t *brick_tab;      // [nbx][nby][nbz]  0 = empty brick, else pool slot + 1
    const uint8_t *pool;            // [slots][512]     material ids, (lx*8 + ly)*8 + lz
    const uint8_t *chunk_res;       // [ncx][ncy][ncz]  0 = no frame, else Frame.resolution
    const double *mats;             // [256][8]         vt_material rows
    unsigned int pool_slots;        // bricks held by the pool (bound for the checked build)
};

// Checked build (-DVT_CHECKED=1, tests/check_bounds.sh): every table index of the render kernel is tested against the
// table's size; a violation records its site code and the access is skipped.  vt_render then fails with that code.
#ifndef VT_CHECKED
#define VT_CHECKED 0
#endif
#if VT_CHECKED
__device__ unsigned int vt_oob_code = 0;
#define VT_IN(cond, code) ((cond) ? true : (atomicMax(&vt_oob_code, (unsigned)(code)), false))
#else
#define VT_IN(cond, code) true
#endif

struct DevCamera {
    double pos[3];
    double rot[4];
    double lens;
    const double *col_sc;           // [width][2]  sin, cos of the column's half yaw  (dof == 0 only)
    const double *row_sc;           // [height][2] sin, cos of the row's half pitch   (dof == 0 only)
    const double *dir_x;            // [width]  -1 + (x / width) * 2   (init.py:131), host division
    const double *dir_y;            // [height] -1 + (y / height) * 2  (init.py:132)
};

struct DevParams {
    vt_settings s;
    DevWorld w;
    DevCamera cam;
    int cs_shift;                   // log2(chunk_size) if it is a power of two, else -1
    uint32_t seed_lo, seed_hi;
    // region
    int x0, y0, x1, y1, stripe_rows, stripe_count, stripe_index, band_rows;
    const int32_t *pixels;
    long long n_pixels;
    // warp tiling: `group` lanes per pixel, 32/group pixels per warp arranged tile_w x tile_h
    int group, group_shift, tile_w, tile_w_shift, tile_h, tiles_x;
    long long n_warps;
    // persistent kernel (vt_march.cuh): ray queue, per-pixel accumulators, LOD snap tables
    unsigned long long total_rays;  // n_tiles * rounds * 32 ray slots (invalid slots included)
    int rounds;                     // ceil(max samples per pixel / group)
    int acc_wide;                   // 0: one u64 per pixel (<= 64 samples), 1: two u64 per pixel
    unsigned long long *acc;        // zero between launches (the last sample of a pixel clears it)
    unsigned long long *queue;      // [0] next ray slot, [1] warps finished; both zero between launches
    unsigned int total_warps;       // warps launched
    int refill_min;                 // refill once this many lanes are idle
    unsigned long long *pow_tab;    // [VT_POW_SLOTS][2] {bits of x, bits of pow(x, 1 + falloff)} filled as values occur; 0 = empty
    int brick_runs;                 // 1: chunk_size % 8 == 0, empty-brick runs allowed (VT_RUNS=0 disables, A/B)
    uint32_t lod_magic[VT_LOD_TAB + 1];   // floor(n / r) == umulhi(n, magic[r]) for n < 2^24 / r ... (r >= 2)
    int32_t lod_bias[VT_LOD_TAB + 1][3];  // world origin mod r per axis
    // multi-GPU frame target: full [height][width] RGBA8 frame (possibly in a peer GPU's memory) + arrival flag
    uint8_t *frame_target;
    unsigned int *arrive_flag;
    unsigned int arrive_value;
    // outputs
    uint8_t *rgba8;
    float *mean;
    double *ray0;
    uint32_t *visited;
    unsigned long long *counters;
};

// ------------------------------------------------------------------------------------------------------
// Philox4x32-10, keyed by the global pixel index, counter = (block, sample, frame seed).  One block gives two
// 53-bit uniforms built like CPython's random.random(): (a >> 5, b >> 6) -> (a * 2^26 + b) / 2^53.
struct Rng {
    uint32_t key0, key1, sample, seed_lo, seed_hi, draw;
    uint32_t w2, w3;
};

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// ------------------------------------------------------------------------------------------------------
// Conversions on the fp64 pipe.  F2I/I2F/FRND with a 64-bit operand issue on the XU pipe (4 lanes/clk/SMSP on
// sm_100); the v1 kernel saturated it (profiles/r1_v2a: xu "realtime" 136 %).  Adding 2^52 + 2^51 puts the
// integer part of |p| < 2^31 into the low mantissa word with the rounding of the add itself.
#define VT_MAGIC 6755399441055744.0            /* 2^52 + 2^51 */
__device__ __forceinline__ int floor_i(double p) { return __double2loint(__dadd_rd(p, VT_MAGIC)); }
__device__ __forceinline__ int rint_i(double p) { return __double2loint(__dadd_rn(p, VT_MAGIC)); }      // ties to even
__device__ __forceinline__ double rint_d(double p) { return __dadd_rn(__dadd_rn(p, VT_MAGIC), -VT_MAGIC); }
__device__ __forceinline__ double floor_d(double p) { return __dadd_rn(__dadd_rd(p, VT_MAGIC), -VT_MAGIC); }
__device__ __forceinline__ double i2d(int x) {                                                          // exact
    return __dadd_rn(__hiloint2double(0x43300000, x ^ (int)0x80000000), -4503601774854144.0);          // 2^52 + 2^31
}
__device__ __forceinline__ double u2d(uint32_t x) { return __dadd_rn(__hiloint2double(0x43300000, (int)x), -4503599627370496.0); }

__device__ __forceinline__ double u53m(uint32_t hi, uint32_t lo) {    // u53 without XU conversions, same value
    return (u2d(hi >> 5) * 67108864.0 + u2d(lo >> 6)) * (1.0 / 9007199254740992.0);
}

__device__ __forceinline__ double u53(uint32_t hi, uint32_t lo) {
    uint32_t a = hi >> 5, b = lo >> 6;
    return ((double)a * 67108864.0 + (double)b) * (1.0 / 9007199254740992.0);
}

__device__ __forceinline__ double rng_next(Rng &g) {
    uint32_t d = g.draw++;
    if ((d & 1u) == 0u) {
        uint32_t w[4];
        philox4x32_10(d >> 1, g.sample, g.seed_lo, g.seed_hi, g.key0, g.key1, w);
        g.w2 = w[2]; g.w3 = w[3];
        return u53(w[0], w[1]);
    }
    return u53(g.w2, g.w3);
}

// lib.py:431-434: rand(amp) returns 0 without consuming a draw when amp is falsy
__device__ __forceinline__ double rng_rand(Rng &g, double amp) {
    if (amp == 0.0) return 0.0;
    return (-1.0 + rng_next(g) * 2.0) * amp;
}

// ------------------------------------------------------------------------------------------------------
// integer helpers (Python // on ints, data.py:137; vec3.snapped on floats reduces to it because the unit is a
// positive integer: floor(p / u) == floor(floor(p) / u))
__device__ __forceinline__ int ifloordiv(int a, int b) {
    int q = a / b, r = a - q * b;
    return (r < 0) ? q - 1 : q;
}
__device__ __forceinline__ int snap_chunk(int ip, int cs, int cs_shift) {
    return cs_shift >= 0 ? ((ip >> cs_shift) << cs_shift) : ifloordiv(ip, cs) * cs;
}
// Python float % with a positive integer modulus (init.py:114): fmod, then wrap negatives up
__device__ __forceinline__ double py_mod_pos(double a, double m) {
    double r = fmod(a, m);
    if (r < 0.0) r += m;
    return r;
}

// chunk slot of the chunk whose origin is (cx, cy, cz), or -1 when outside the table (`post in self.chunks`)
__device__ __forceinline__ int chunk_slot(const DevWorld &w, int cs, int cs_shift, int cx, int cy, int cz) {
    int rx = cx - w.ox, ry = cy - w.oy, rz = cz - w.oz;
    if (rx < 0 || ry < 0 || rz < 0 || rx >= w.nx || ry >= w.ny || rz >= w.nz) return -1;
    int sx, sy, sz;
    if (cs_shift >= 0) { sx = rx >> cs_shift; sy = ry >> cs_shift; sz = rz >> cs_shift; }
    else { sx = rx / cs; sy = ry / cs; sz = rz / cs; }
    return (sx * w.ncy + sy) * w.ncz + sz;
}

// Frame.get_voxel (data.py:136-145) on the LOD-`res` frame of the chunk at (cx, cy, cz): key floor(p) // res
// exists iff the full-resolution cell res*key lies inside this chunk (init.py:417-424, data.py:165-169).
__device__ __forceinline__ int get_voxel(const DevWorld &w, int cs, int cx, int cy, int cz, int res, int px, int py,
                                         int pz) {
    int qx = px, qy = py, qz = pz;
    if (res > 1) { qx = ifloordiv(px, res) * res; qy = ifloordiv(py, res) * res; qz = ifloordiv(pz, res) * res; }
    if (qx < cx || qy < cy || qz < cz || qx >= cx + cs || qy >= cy + cs || qz >= cz + cs) return 0;
    int rx = qx - w.ox, ry = qy - w.oy, rz = qz - w.oz;
    uint32_t b = __ldg(&w.brick_tab[((rx >> 3) * w.nby + (ry >> 3)) * w.nbz + (rz >> 3)]);
    if (b == 0u) return 0;
    return __ldg(&w.pool[(size_t)(b - 1u) * 512u + (((rx & 7) << 6) | ((ry & 7) << 3) | (rz & 7))]);
}

// `pos >= chunk_min and pos <= chunk_max` on one axis (lib.py:228-244), chunk_max inclusive: the position may sit
// exactly on the far face and still count as inside.
__device__ __forceinline__ bool axis_inside(double p, int ip, int lo, int hi) {
    return ip >= lo && (ip < hi || (ip == hi && p == (double)ip));
}

// ------------------------------------------------------------------------------------------------------
struct Ray {
    double px, py, pz, vx, vy, vz;
    double step, life, bounces, energy;
    double cr, cg, cb;              // rgb integers held in doubles (re-rounded after every mix, lib.py:393-395)
    int hits, fflag, fx, fy, fz;
    double first_step;
};

struct Counters {
    unsigned int in_chunk, in_void, hits, probes, rechunk, iters;
};

// init.py:43-45.  The lens euler vector is (0, -lens_x, +lens_y) degrees, so sin_x = 0 and cos_x = 1 in
// lib.py:322-338; the zero terms are kept so signed zeros come out as in the reference.  (sy, cy) / (sz, cz)
// are sin/cos of half the yaw / pitch angle.
__device__ __forceinline__ void ray_dir(const DevCamera &cam, double sy, double cy, double sz, double cz,
                                        double &dx, double &dy, double &dz) {
    const double sin_x = 0.0, cos_x = 1.0;
    double ox = sin_x * cy * cz - cos_x * sy * sz;
    double oy = cos_x * sy * cz - sin_x * cy * sz;
    double oz = cos_x * cy * sz + sin_x * sy * cz;
    double ow = cos_x * cy * cz + sin_x * sy * sz;
    double qx = cam.rot[0], qy = cam.rot[1], qz = cam.rot[2], qw = cam.rot[3];
    // quaternion.multiply, the reference's own sign layout (lib.py:353-358)
    double x = qw * ox + qz * oy - qy * oz + qx * ow;
    double y = qz * ox + qw * oy + qx * oz + qy * ow;
    double z = qy * ox - qx * oy + qw * oz + qz * ow;
    double w = qx * ox - qy * oy - qz * oz + qw * ow;
    // quaternion.vec_forward (lib.py:372-376)
    dx = 2.0 * (z * x + w * y);
    dy = 2.0 * (y * x - w * z);
    dz = 1.0 - 2.0 * (z * z + y * y);
}

__device__ __forceinline__ double mixd(double a, double b, double bias1) {
    double bias2 = 1.0 - bias1;
    return a * bias2 + b * bias1;
}

// lib.material (lib.py:448-460); returns the bounce weight (mat.absorption)
__device__ __forceinline__ double shade_material(Ray &r, const double *__restrict__ m, const vt_settings &s, Rng &g) {
    double alb_r = __ldg(m + 0), alb_g = __ldg(m + 1), alb_b = __ldg(m + 2);
    double rough = __ldg(m + 3), absorb = __ldg(m + 4), emit = __ldg(m + 6);
    double a = fmin(1.0, absorb / pow(1.0 + r.bounces, 1.0 + s.falloff));
    double b2 = 1.0 - a;
    r.cr = rint(r.cr * b2 + alb_r * a);
    r.cg = rint(r.cg * b2 + alb_g * a);
    r.cb = rint(r.cb * b2 + alb_b * a);
    r.energy = r.energy * b2 + emit * a;
    r.life *= 1.0 - (rough * a);
    double jx = rng_rand(g, rough), jy = rng_rand(g, rough), jz = rng_rand(g, rough);
    r.vx += jx; r.vy += jy; r.vz += jz;
    return absorb;
}

// lib.material_background (lib.py:463-476)
__device__ __forceinline__ void shade_background(Ray &r, const vt_settings &s) {
    double a = fmin(1.0, 1.0 / pow(1.0 + r.bounces, 1.0 + s.falloff));
    double up = fmax(0.0, r.vy);
    double b2 = 1.0 - a;
    r.cr = rint(r.cr * b2 + 127.0 * a);
    r.cg = rint(r.cg * b2 + (127.0 + up * 64.0) * a);
    r.cb = rint(r.cb * b2 + (127.0 + up * 128.0) * a);
    r.energy = mixd(r.energy, 1.0 + up, a);
    r.cr = fmin(255.0, rint(r.cr * r.energy));
    r.cg = fmin(255.0, rint(r.cg * r.energy));
    r.cb = fmin(255.0, rint(r.cb * r.energy));
}

__device__ __forceinline__ void mark_visited(uint32_t *visited, int slot) {
    if (!visited) return;
    uint32_t bit = 1u << (slot & 31);
    uint32_t *word = visited + (slot >> 5);
    if (!VT_IN(slot >= 0, 10)) return;
    if (!(*(volatile uint32_t *)word & bit)) atomicOr(word, bit);
}

// Brick-table entry cache: consecutive samples of a ray mostly stay inside one 8^3 brick, so the table is
// re-read only when the brick coordinate changes (one dependent load per sample instead of two).
struct BrickCache {
    int key;            // packed brick coordinate, -1 = nothing cached
    uint32_t slot1;     // table entry: 0 = empty brick, else pool slot + 1
};

__device__ __forceinline__ int get_voxel_cached(const DevWorld &w, int cs, int cx, int cy, int cz, int res, int px,
                                                int py, int pz, BrickCache &bc) {
    int qx = px, qy = py, qz = pz;
    if (res > 1) { qx = ifloordiv(px, res) * res; qy = ifloordiv(py, res) * res; qz = ifloordiv(pz, res) * res; }
    if (qx < cx || qy < cy || qz < cz || qx >= cx + cs || qy >= cy + cs || qz >= cz + cs) return 0;
    int rx = qx - w.ox, ry = qy - w.oy, rz = qz - w.oz;
    int key = ((rx >> 3) * w.nby + (ry >> 3)) * w.nbz + (rz >> 3);
    if (key != bc.key) { bc.key = key; bc.slot1 = __ldg(&w.brick_tab[key]); }
    if (bc.slot1 == 0u) return 0;
    return __ldg(&w.pool[(size_t)(bc.slot1 - 1u) * 512u + (((rx & 7) << 6) | ((ry & 7) << 3) | (rz & 7))]);
}

// void step of init.py:114: 1 + |radius - (min(pos) + radius) % chunk_size| with Python's float modulo.
// For a power-of-two chunk size a/cs, floor and cs*k are exact, so a - cs*floor(a/cs) is the same single rounding
// Python performs (fmod is exact, then `mod += cs` rounds once for negatives); otherwise use fmod itself.
__device__ __forceinline__ double void_step(double px, double py, double pz, int radius, int cs, int cs_shift) {
    double a = fmin(px, fmin(py, pz)) + (double)radius;
    double m;
    if (cs_shift >= 0) { double dcs = (double)cs; m = a - dcs * floor(a / dcs); }
    else m = py_mod_pos(a, (double)cs);
    return 1.0 + fabs((double)radius - m);
}

// Camera.trace (init.py:37-121), executed by a whole warp in lock step: every lane owns one ray.
//
// The reference's loop body mixes a cheap common case (advance one sample) with a rare expensive one (a material
// hit: pow, divisions, three neighbour probes, RNG).  Run naively, the lanes of a warp serialise on the rare
// branch.  Here each round has two phases: (1) every live lane marches until it finds a solid cell or dies;
// (2) the lanes that found one shade together.  Sample positions and their order per ray are unchanged.
//
// FULL_TAIL = false lets a ray stop once it is outside the world box and moving away from it on that axis:
// nothing it could still touch changes colour, energy, first hit or the visited set of existing chunks (only
// ray.step / ray.pos, which no caller of tile() sees).
template <bool FULL_TAIL>
__device__ __forceinline__ void trace_warp(const DevParams &P, bool valid, double dir_x, double dir_y, double detail,
                                           Rng &g, Ray &r, Counters &c, int px_x, int px_y) {
    const vt_settings &s = P.s;
    const DevWorld &w = P.w;
    const int cs = s.chunk_size, cs_shift = P.cs_shift;
    r.cr = r.cg = r.cb = 0.0; r.energy = 0.0; r.step = 0.0; r.bounces = 0.0; r.life = 0.0;
    r.px = r.py = r.pz = 0.0; r.vx = r.vy = r.vz = 0.0;
    r.hits = 0; r.fflag = 0; r.fx = r.fy = r.fz = 0; r.first_step = -1.0;
    if (valid) {
        double sy, cy, sz, cz;
        if (s.dof == 0.0 && P.cam.col_sc) {
            sy = __ldg(P.cam.col_sc + 2 * px_x); cy = __ldg(P.cam.col_sc + 2 * px_x + 1);
            sz = __ldg(P.cam.row_sc + 2 * px_y); cz = __ldg(P.cam.row_sc + 2 * px_y + 1);
        } else {
            const double deg2rad = 3.14159265358979323846 / 180.0;
            double lens_x = (dir_x / s.proportions) * P.cam.lens + rng_rand(g, s.dof);
            double lens_y = (dir_y * s.proportions) * P.cam.lens + rng_rand(g, s.dof);
            double rad_y = (-lens_x) * deg2rad, rad_z = (+lens_y) * deg2rad;
            sincos(rad_y / 2.0, &sy, &cy);
            sincos(rad_z / 2.0, &sz, &cz);
        }
        double dx, dy, dz;
        ray_dir(P.cam, sy, cy, sz, cz, dx, dy, dz);
        r.px = P.cam.pos[0] + dx * s.dist_min; r.py = P.cam.pos[1] + dy * s.dist_min; r.pz = P.cam.pos[2] + dz * s.dist_min;
        r.vx = dx; r.vy = dy; r.vz = dz;
        r.life = (s.dist_max - s.dist_min) * detail;
    }

    int cx = 0, cy_ = 0, cz_ = 0;        // chunk_min
    int hx = 0, hy = 0, hz = 0;          // chunk_max (= chunk_min + cs except for the initial (0,0,0)/(0,0,0))
    int res = 0;                         // 0 == `chunk is None`
    int ix = 0, iy = 0, iz = 0;          // floor(pos) of the sample being examined
    int hit = 0;                         // material id found by phase 1, consumed by phase 2
    BrickCache bc; bc.key = -1; bc.slot1 = 0u;
    const int wx1 = w.ox + w.nx, wy1 = w.oy + w.ny, wz1 = w.oz + w.nz;
    bool alive = valid && (r.step < r.life);

    for (;;) {
        // ---- phase 1: march to the next solid cell (init.py:66-77, 114-116) -------------------------------
        while (alive && !hit) {
            c.iters++;
            ix = __double2int_rd(r.px); iy = __double2int_rd(r.py); iz = __double2int_rd(r.pz);
            if (!(axis_inside(r.px, ix, cx, hx) && axis_inside(r.py, iy, cy_, hy) && axis_inside(r.pz, iz, cz_, hz))) {
                cx = snap_chunk(ix, cs, cs_shift); cy_ = snap_chunk(iy, cs, cs_shift); cz_ = snap_chunk(iz, cs, cs_shift);
                hx = cx + cs; hy = cy_ + cs; hz = cz_ + cs;
                int slot = chunk_slot(w, cs, cs_shift, cx, cy_, cz_);
                res = 0;
                if (slot >= 0) { res = __ldg(&w.chunk_res[slot]); mark_visited(P.visited, slot); }
                c.rechunk++;
                if (!FULL_TAIL && slot < 0) {
                    bool gone = (ix >= wx1 && r.vx >= 0.0) || (ix < w.ox && r.vx <= 0.0) ||
                                (iy >= wy1 && r.vy >= 0.0) || (iy < w.oy && r.vy <= 0.0) ||
                                (iz >= wz1 && r.vz >= 0.0) || (iz < w.oz && r.vz <= 0.0);
                    if (gone) { alive = false; break; }
                }
            }
            double step;
            if (res) {
                c.in_chunk++;
                hit = get_voxel_cached(w, cs, cx, cy_, cz_, res, ix, iy, iz, bc);
                if (hit) break;                       // shade in phase 2, then advance there
                step = (double)res;
            } else {
                c.in_void++;
                step = void_step(r.px, r.py, r.pz, s.chunk_radius, cs, cs_shift);
            }
            r.step += step;
            r.px += r.vx * step; r.py += r.vy * step; r.pz += r.vz * step;
            alive = r.step < r.life;
        }
        if (!__any_sync(0xffffffffu, hit != 0)) break;
        // ---- phase 2: material interaction (init.py:78-111), all hitting lanes together ----------------------
        if (hit) {
            const double *m = w.mats + 8 * hit;
            if (!r.fflag) { r.fflag = 1; r.fx = ix; r.fy = iy; r.fz = iz; r.first_step = r.step; }
            r.hits++; c.hits++;
            double bounce = shade_material(r, m, s, g);                           // init.py:81
            r.bounces += bounce;
            r.life /= (double)res + bounce * s.lod_bounces;                       // init.py:83
            double ref = fmax(fabs(r.vx), fmax(fabs(r.vy), fabs(r.vz)));         // vec3.normalize, lib.py:310-314
            if (ref != 0.0 && ref != 1.0) { r.vx /= ref; r.vy /= ref; r.vz /= ref; }
            if (r.step >= r.life || r.energy >= s.max_light || r.bounces >= s.max_bounces + 1.0) {
                alive = false;                                                    // `break`, init.py:85-86
            } else {
                double ior = __ldg(m + 5);
                if (ior != 0.0) {                                                 // init.py:92-111
                    double direction = (ior - 0.5) * 2.0;
                    bool flip[3];
#pragma unroll
                    for (int a = 0; a < 3; a++) {
                        double va = a == 0 ? r.vx : a == 1 ? r.vy : r.vz;
                        double qa = a == 0 ? r.px : a == 1 ? r.py : r.pz;
                        qa = (va < direction) ? qa + 1.0 : qa - 1.0;
                        int ja = __double2int_rd(qa);
                        int jx = a == 0 ? ja : ix, jy = a == 1 ? ja : iy, jz = a == 2 ? ja : iz;
                        int lo = a == 0 ? cx : a == 1 ? cy_ : cz_, hi = a == 0 ? hx : a == 1 ? hy : hz;
                        int nid = 0;
                        if (axis_inside(qa, ja, lo, hi)) {
                            nid = get_voxel_cached(w, cs, cx, cy_, cz_, res, jx, jy, jz, bc); c.probes++;
                        } else {                                                  // Camera.chunk_get, init.py:28-33
                            int nx = snap_chunk(jx, cs, cs_shift), ny = snap_chunk(jy, cs, cs_shift), nz = snap_chunk(jz, cs, cs_shift);
                            int nslot = chunk_slot(w, cs, cs_shift, nx, ny, nz);
                            int nres = nslot >= 0 ? (int)__ldg(&w.chunk_res[nslot]) : 0;
                            if (nres) { nid = get_voxel_cached(w, cs, nx, ny, nz, nres, jx, jy, jz, bc); c.probes++; }
                        }
                        flip[a] = (nid == 0) || (__ldg(w.mats + 8 * nid + 5) != ior);
                    }
                    if (flip[0]) r.vx -= r.vx * ior * 2.0;
                    if (flip[1]) r.vy -= r.vy * ior * 2.0;
                    if (flip[2]) r.vz -= r.vz * ior * 2.0;
                }
                double step = (double)res;                                        // init.py:114-116
                r.step += step;
                r.px += r.vx * step; r.py += r.vy * step; r.pz += r.vz * step;
                alive = r.step < r.life;
            }
            hit = 0;
        }
    }
    if (valid && s.background) shade_background(r, s);                             // init.py:119-120
}

}  // namespace vt
