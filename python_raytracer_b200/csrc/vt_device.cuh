// vt_device.cuh — parameter blocks and arithmetic helpers of the render kernel (vt_march.cuh) for sm_100a.
//
// The kernel restates, as compiled device functions, what the reference does per ray in Python:
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

#define VT_LOD_TAB 16                                 // Frame.resolution values served by the multiply-high snap table
#define VT_POW_BITS 16
#define VT_POW_SLOTS (1 << VT_POW_BITS)               // entries of the memoised falloff-power table

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
    const double *lens_x;           // [width]  (dir_x / proportions) * lens  (init.py:41), host arithmetic
    const double *lens_y;           // [height] (dir_y * proportions) * lens  (init.py:42)
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
    double inv_samples;             // 1 / n when every pixel has the same power-of-two sample count n (exact mean), else 0
    int brick_runs;                 // 1: chunk_size % 8 == 0, empty-brick runs allowed (VT_RUNS=0 disables, A/B)
    uint32_t lod_magic[VT_LOD_TAB + 1];   // floor(n / r) == umulhi(n, magic[r]) for n < 2^24 / r ... (r >= 2)
    int32_t lod_bias[VT_LOD_TAB + 1][3];  // world origin mod r per axis
    // multi-GPU frame target: full [height][width] RGBA8 frame (possibly in a peer GPU's memory) + arrival flag
    uint8_t *frame_target;
    unsigned int *arrive_flag;
    unsigned int arrive_value;
    // multi-GPU culling: the launch's last warp ORs this rank's visited words into a peer's mask before the flag goes up
    uint32_t *visited_peer;
    int visited_words;
    // trace mode (vt_trace_rays, REC instantiations only): rays given as (dir_x, dir_y, detail) triples, not pixels
    const double *dirs;             // [n][3]
    const double *dir_sc;           // [n][4] sin, cos of the half yaw, sin, cos of the half pitch (host libm; dof == 0) or null
    double *trace_out;              // [n][VT_TRACE_RECORD]
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

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#ifndef VT_PHILOX_UNROLL
#define VT_PHILOX_UNROLL 10
#endif
#define VT_STR2(x) #x
#define VT_STR(x) VT_STR2(x)
    _Pragma(VT_STR(unroll VT_PHILOX_UNROLL))
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
// sm_100); the first kernel saturated it (profiles/r1_v2a: xu "realtime" 136 %).  Adding 2^52 + 2^51 puts the
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

// ------------------------------------------------------------------------------------------------------
// integer helpers (Python // on ints, data.py:137; vec3.snapped on floats reduces to it because the unit is a
// positive integer: floor(p / u) == floor(floor(p) / u))
__device__ __forceinline__ int ifloordiv(int a, int b) {
    int q = a / b, r = a - q * b;
    return (r < 0) ? q - 1 : q;
}
// Python float % with a positive integer modulus (init.py:114): fmod, then wrap negatives up
__device__ __forceinline__ double py_mod_pos(double a, double m) {
    double r = fmod(a, m);
    if (r < 0.0) r += m;
    return r;
}

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

__device__ __forceinline__ void mark_visited(uint32_t *visited, int slot) {
    if (!visited) return;
    uint32_t bit = 1u << (slot & 31);
    uint32_t *word = visited + (slot >> 5);
    if (!VT_IN(slot >= 0, 10)) return;
    if (!(*(volatile uint32_t *)word & bit)) atomicOr(word, bit);
}



}  // namespace vt
