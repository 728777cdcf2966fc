// vt_march.cuh — persistent ray-march kernel for sm_100a: Camera.tile / Camera.trace (reference init.py:37-150),
// lib.material / material_background (lib.py:448-476) and Frame.get_voxel (data.py:136-145) for a whole frame.
// Every ray produces the reference's sample positions, hits and colours; what is specific to the GPU is the schedule:
//
//   * One persistent grid per frame.  Every lane owns one ray at a time; a warp draws a PACKET of 32 ray slots (one
//     sample index of an 8 x 4 pixel tile) from a global queue when all its rays have ended.  (Refilling single lanes
//     as their rays end is implemented -- `refill_min` < 32 -- and measured slower: young primary rays mixed with
//     bounced ones march with fewer lanes per trip)  Pixel means are accumulated with one 64-bit
//     atomic per ray: the channel sums are sums of integers 0..255, so the order of arrival cannot change the result;
//     the sample that completes a pixel writes it and clears the accumulator.
//   * A warp is a small state machine.  Lanes are MARCH (stepping through cells), HIT (found a solid cell,
//     material interaction pending) or idle (ray ended / no ray).  Each round the warp runs the phase most lanes
//     wait for, so the expensive rare work (pow, divisions, Philox, sincos, atomics) runs with many lanes active
//     instead of serialising a warp on one lane.
//   * Per-sample work is integer-only apart from the position update: cells are addressed relative to the world
//     origin, the chunk test is one unsigned max, LOD snapping is a multiply-high.
#pragma once
#include "vt_device.cuh"
#include "vt_pow_cr.h"

// Compile-time switches (tools/build_variants.py); they must precede their first use.
#ifndef VT_MARCH_KEEP
#define VT_MARCH_KEEP 1             // the march phase goes on while >= KEEP/8 of the lanes that entered it still march
#endif
#ifndef VT_TRIPS           // dev instrumentation (tools/kbench.py KB_COUNTERS): phase trip counts in counters[8..13]
#define VT_TRIPS 0
#endif
#ifndef VT_SINCOS_SMALL
#define VT_SINCOS_SMALL 1            // sin / cos of the half lens angles without argument reduction when |a| <= pi/4
#endif
#ifndef VT_INVN
#define VT_INVN 1                    // pixel mean by an exact reciprocal multiply when the sample count is a power of two
#endif
#ifndef VT_COLD_SMEM
#define VT_COLD_SMEM 1               // per-ray state the march loops never touch lives in shared memory, not in registers
#endif
#ifndef VT_MINB
#define VT_MINB 10                   // resident 128-thread blocks per SM the register budget is sized for (48 registers)
#endif

namespace vt {

// chunk_min that no cell lies in (forces a chunk re-evaluation): cells relative to the world origin stay below 2.1e9 in
// magnitude (|camera|, |origin| <= 1e9, dist_max < 1e8; checked by the C ABI), and [VT_NO_BOX, VT_NO_BOX + 4096) lies above
#define VT_NO_BOX 0x7fffe000
enum : uint32_t { ST_NONE = 0u, ST_MARCH = 1u, ST_HIT = 2u, ST_DONE = 3u };

// Python's min(a, b) / max(a, b): the first argument unless the second is strictly smaller / larger.  No value on this
// path is ever NaN, so a compare and a select do (fmin / fmax carry 8 instructions of NaN handling each on sm_100).
__device__ __forceinline__ double pmin(double a, double b) { return b < a ? b : a; }
__device__ __forceinline__ double pmax(double a, double b) { return b > a ? b : a; }

// st: bits 0-1 state, bits 2-9 material id of a pending hit, bits 12-31 sample index of the ray
// Hot state (touched by every sample: 8 doubles + 5 words = 21 registers) and cold state (touched when a ray starts, is
// shaded or ends: 2 doubles + 6 words).  At 48 warps per SM the kernel has 40 registers per thread; with everything in
// registers ptxas spilled ~430 bytes per thread.  The cold fields therefore live in shared memory (structure of arrays,
// 40 bytes per thread, 60 KB per SM at 12 blocks): they are reference members, so the code below reads the same either
// way (-DVT_COLD_SMEM=0 puts them back into registers for an A/B: 1.760 vs 1.755 ms/frame, spill stores 406 -> 278 B).
struct ColdStore {
    double bounces[128], energy[128];
    int col[128];
    uint32_t xy[128], out_i[128], draw[128], w2[128], w3[128], nsmp[128];
};
struct Lane {
    double px, py, pz, vx, vy, vz, step, life;
    int cx, cy, cz;                 // chunk_min relative to the world origin (chunk_max = chunk_min + chunk_size)
    int res;                        // Frame.resolution of the current chunk, 0 = `chunk is None`
    uint32_t st;
#if VT_COLD_SMEM
    double &bounces, &energy;
    int &col;                       // ray.color as r | g << 8 | b << 16 (integers, lib.py:393-395)
    uint32_t &xy;                   // pixel x | y << 16
    uint32_t &out_i;                // output index of the pixel
    uint32_t &draw, &w2, &w3;       // Philox draw counter and the cached second uniform of the last block
    uint32_t &nsmp;                 // samples traced for the ray's pixel (init.py:134)
    __device__ __forceinline__ Lane(ColdStore &s, int t)
        : bounces(s.bounces[t]), energy(s.energy[t]), col(s.col[t]), xy(s.xy[t]), out_i(s.out_i[t]), draw(s.draw[t]), w2(s.w2[t]), w3(s.w3[t]),
          nsmp(s.nsmp[t]) {}
#else
    double bounces, energy;
    int col;
    uint32_t xy, out_i, draw, w2, w3, nsmp;
    __device__ __forceinline__ Lane(ColdStore &, int) {}
#endif
};

struct FirstHit {                   // parity probe state, only live in the REC instantiation
    int hits, fflag, fx, fy, fz;
    double first_step;
};

__device__ __forceinline__ uint32_t lane_state(const Lane &L) { return L.st & 3u; }
__device__ __forceinline__ uint32_t lane_sample(const Lane &L) { return L.st >> 12; }
__device__ __forceinline__ void lane_set_state(Lane &L, uint32_t s) { L.st = (L.st & ~0xffcu & ~3u) | s; }

// Out-of-line helpers: each is called from several places and is long (Philox 80 instructions, sincos and pow with
// their argument-reduction paths); inlined copies made the kernel 25 % larger and it is sensitive to instruction-cache
// misses (profiles/r1_v2b: no_instruction is the largest stall reason).
__device__ __noinline__ uint4 philox_block(uint32_t block, uint32_t sample, uint32_t seed_lo, uint32_t seed_hi, uint32_t pix) {
    uint32_t w[4];
    philox4x32_10(block, sample, seed_lo, seed_hi, pix, 0u, w);
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __noinline__ double2 sincos_out(double a) {
#if VT_SINCOS_SMALL
    // Half lens angles are small (fov 90: |a| < 0.41 rad): below pi/4 no argument reduction and no quadrant logic are
    // needed, sin and cos are two Horner chains of explicit FMAs (Taylor to a^17 / a^16, truncation < 1e-19 at pi/4;
    // within 1 ulp of libm like libdevice's own sincos, which parity tolerates -- DESIGN.md section 4, Numerics).
    if (fabs(a) <= 0.78539816339744830962) {
        const double z = a * a;
        double p = 2.81145725434552076320e-15;                                  //  1 / 17!
        p = __fma_rn(p, z, -7.64716373181981647590e-13);                       // -1 / 15!
        p = __fma_rn(p, z, 1.60590438368216145994e-10);                        //  1 / 13!
        p = __fma_rn(p, z, -2.50521083854417187751e-08);                       // -1 / 11!
        p = __fma_rn(p, z, 2.75573192239858906526e-06);                        //  1 / 9!
        p = __fma_rn(p, z, -1.98412698412698412698e-04);                       // -1 / 7!
        p = __fma_rn(p, z, 8.33333333333333333333e-03);                        //  1 / 5!
        p = __fma_rn(p, z, -1.66666666666666666667e-01);                       // -1 / 3!
        const double s = __fma_rn(a * z, p, a);
        double q = 4.77947733238738529744e-14;                                  //  1 / 16!
        q = __fma_rn(q, z, -1.14707455977297247139e-11);                       // -1 / 14!
        q = __fma_rn(q, z, 2.08767569878680989792e-09);                        //  1 / 12!
        q = __fma_rn(q, z, -2.75573192239858906526e-07);                       // -1 / 10!
        q = __fma_rn(q, z, 2.48015873015873015873e-05);                        //  1 / 8!
        q = __fma_rn(q, z, -1.38888888888888888889e-03);                       // -1 / 6!
        q = __fma_rn(q, z, 4.16666666666666666667e-02);                        //  1 / 4!
        const double hz = 0.5 * z, w = 1.0 - hz;                               // 1 - z/2 with its rounding error fed back
        const double c = w + (((1.0 - w) - hz) + (z * z) * q);
        return make_double2(s, c);
    }
#endif
    double s, c;
    sincos(a, &s, &c);
    return make_double2(s, c);
}
// (1 + bounces) ** (1 + falloff): correctly rounded (vt_pow_cr.h), which is what glibc's pow returns for 99.9 % of arguments;
// libdevice's pow (<= 2 ulp) left the last bit of `energy` different from the reference's on every second ray
__device__ __noinline__ double pow_out(double a, double b) { return vtp::pow_cr(a, b); }

// (1 + bounces) ^ (1 + falloff) of lib.material / material_background (lib.py:450, 465).  `bounces` is a sum of a few
// material absorptions, so a frame asks for the same handful of values millions of times: they are kept in a small
// table in device memory, filled by whoever computes a value first.  Entries only go from empty to {x, pow(x)} (key
// by compare-and-swap, then the value; a reader that sees the key without the value computes it itself) and the
// value stored is the one libdevice's pow returns for that x, so the table can never change a result.  The host
// clears it when `falloff` changes.
__device__ __noinline__ double pow_cached(unsigned long long *tab, double x, double y) {
    const unsigned long long key = (unsigned long long)__double_as_longlong(x);
    const uint32_t h = ((uint32_t)(key >> 32) ^ (uint32_t)key * 0x9E3779B1u) * 0x85EBCA6Bu;
    unsigned long long *e = tab + 2 * (h >> (32 - VT_POW_BITS));
    ulonglong2 kv = __ldcg(reinterpret_cast<const ulonglong2 *>(e));
    if (kv.x != key && kv.x != 0ull) {                                  // taken by another value: second home
        e = tab + 2 * ((h >> (32 - VT_POW_BITS)) ^ 1u);
        kv = __ldcg(reinterpret_cast<const ulonglong2 *>(e));
    }
    if (kv.x == key && kv.y != 0ull) return __longlong_as_double((long long)kv.y);
    const double v = pow_out(x, y);
    const unsigned long long old = kv.x == key ? key : atomicCAS(e, 0ull, key);
    if (old == 0ull || old == key) atomicExch(e + 1, (unsigned long long)__double_as_longlong(v));
    return v;
}
__device__ __forceinline__ double bounce_pow(const DevParams &P, double bounces) {
    if (bounces == 0.0) return 1.0;                                     // pow(1, y) == 1 exactly
    return P.pow_tab ? pow_cached(P.pow_tab, 1.0 + bounces, 1.0 + P.s.falloff) : pow_out(1.0 + bounces, 1.0 + P.s.falloff);
}

// ---- Philox stream of one ray: key = global pixel index, counter = (block, sample, frame seed) --------------
__device__ __forceinline__ double lane_rng_next(const DevParams &P, Lane &L) {
    uint32_t d = L.draw++;
    if ((d & 1u) == 0u) {
        uint32_t pix = (L.xy >> 16) * (uint32_t)P.s.width + (L.xy & 0xffffu);
        uint4 w = philox_block(d >> 1, lane_sample(L), P.seed_lo, P.seed_hi, pix);
        L.w2 = w.z; L.w3 = w.w;
        return u53m(w.x, w.y);
    }
    return u53m(L.w2, L.w3);
}
__device__ __forceinline__ double lane_rng_rand(const DevParams &P, Lane &L, double amp) {   // lib.py:431-434
    if (amp == 0.0) return 0.0;
    return (-1.0 + lane_rng_next(P, L) * 2.0) * amp;
}

// Three consecutive rand(amp) draws (the velocity jitter of lib.material, lib.py:458): same stream positions as three
// lane_rng_rand calls, but the Philox blocks are fetched where all lanes need one together -- block (draw + 1) >> 1
// serves both parities -- so a warp whose lanes sit at odd and even draw counts runs Philox twice, not three times.
__device__ __forceinline__ void lane_rng_rand3(const DevParams &P, Lane &L, double amp, double &r0, double &r1, double &r2) {
    r0 = r1 = r2 = 0.0;
    if (amp == 0.0) return;                                                 // rand(0) draws nothing (lib.py:431-434)
    const uint32_t d = L.draw;
    const uint32_t pix = (L.xy >> 16) * (uint32_t)P.s.width + (L.xy & 0xffffu);
    const uint4 a = philox_block((d + 1u) >> 1, lane_sample(L), P.seed_lo, P.seed_hi, pix);
    double u0, u1, u2;
    if (d & 1u) { u0 = u53m(L.w2, L.w3); u1 = u53m(a.x, a.y); u2 = u53m(a.z, a.w); }
    else {
        u0 = u53m(a.x, a.y); u1 = u53m(a.z, a.w);
        const uint4 b = philox_block((d >> 1) + 1u, lane_sample(L), P.seed_lo, P.seed_hi, pix);
        u2 = u53m(b.x, b.y);
        L.w2 = b.z; L.w3 = b.w;
    }
    L.draw = d + 3u;
    r0 = (-1.0 + u0 * 2.0) * amp; r1 = (-1.0 + u1 * 2.0) * amp; r2 = (-1.0 + u2 * 2.0) * amp;
}

// Integer divisions by run-time divisors (chunk sizes that are not powers of two, resolutions above the snap table)
// live in functions that are never inlined: inlined, ptxas if-converts the rare branch and every sample pays for a
// 20-instruction division it does not use.
__device__ __noinline__ int floordiv_mul_slow(int a, int b) { return ifloordiv(a, b) * b; }
__device__ __noinline__ int div_slow(int a, int b) { return a / b; }

// ---- cell addressing, relative to the world origin ------------------------------------------------------------
// res * floor(abs / res) - origin for abs = r + origin, r >= 0 (data.py:137 with Python's floor division)
template <bool FAST>
__device__ __forceinline__ int lod_snap(const DevParams &P, int res, int r, int axis) {
    if (FAST || res <= VT_LOD_TAB) {
        int b = P.lod_bias[res][axis];
        uint32_t f = __umulhi((uint32_t)(r + b), P.lod_magic[res]);
        return (int)(f * (uint32_t)res) - b;
    }
    int o = axis == 0 ? P.w.ox : axis == 1 ? P.w.oy : P.w.oz;
    return floordiv_mul_slow(r + o, res) - o;
}

__device__ __forceinline__ int cell_at(const DevWorld &w, int qx, int qy, int qz) {
    if (!VT_IN((unsigned)(qx >> 3) < (unsigned)w.nbx && (unsigned)(qy >> 3) < (unsigned)w.nby && (unsigned)(qz >> 3) < (unsigned)w.nbz, 1)) return 0;
    uint32_t b = __ldg(&w.brick_tab[((qx >> 3) * w.nby + (qy >> 3)) * w.nbz + (qz >> 3)]);
    if (b == 0u || !VT_IN(b <= w.pool_slots, 2)) return 0;
    return __ldg(&w.pool[(size_t)(b - 1u) * 512u + (((qx & 7) << 6) | ((qy & 7) << 3) | (qz & 7))]);
}

// Frame.get_voxel (data.py:136-145) of the LOD-`res` frame of the chunk at (cx, cy, cz) for ANY cell (jx, jy, jz):
// the key floor(p) // res exists iff the cell res * key lies inside that chunk (init.py:417-424, data.py:165-169).
template <bool FAST>
__device__ __forceinline__ int lookup_any(const DevParams &P, int cx, int cy, int cz, int res, int jx, int jy, int jz) {
    const int cs = P.s.chunk_size;
    if ((jx | jy | jz) < 0) return 0;                  // below the world box: res * key <= j is outside too
    if (res > 1) { jx = lod_snap<FAST>(P, res, jx, 0); jy = lod_snap<FAST>(P, res, jy, 1); jz = lod_snap<FAST>(P, res, jz, 2); }
    if ((unsigned)(jx - cx) >= (unsigned)cs || (unsigned)(jy - cy) >= (unsigned)cs || (unsigned)(jz - cz) >= (unsigned)cs)
        return 0;
    return cell_at(P.w, jx, jy, jz);
}

template <bool FAST>
__device__ __forceinline__ int snap_rel(int r, int cs, int cs_shift) {   // origin is a multiple of chunk_size
    if (FAST || cs_shift >= 0) return (r >> cs_shift) << cs_shift;
    return floordiv_mul_slow(r, cs);
}

// chunk slot of the chunk at relative origin (cx, cy, cz), or -1 (`post in self.chunks` is false outside the box)
template <bool FAST>
__device__ __forceinline__ int chunk_slot_rel(const DevWorld &w, int cs, int cs_shift, int cx, int cy, int cz) {
    if ((unsigned)cx >= (unsigned)w.nx || (unsigned)cy >= (unsigned)w.ny || (unsigned)cz >= (unsigned)w.nz) return -1;
    int sx, sy, sz;
    if (FAST || cs_shift >= 0) { sx = cx >> cs_shift; sy = cy >> cs_shift; sz = cz >> cs_shift; }
    else { sx = div_slow(cx, cs); sy = div_slow(cy, cs); sz = div_slow(cz, cs); }
    return (sx * w.ncy + sy) * w.ncz + sz;
}

// init.py:67-73: re-snap the chunk box around the cell (ix, iy, iz), look the chunk up, record it as traversed.
// Returns false when the ray can stop (FULL_TAIL == false): outside the world box and moving away from it.
template <bool FAST, bool FULL_TAIL, bool COUNT>
__device__ __forceinline__ bool rechunk(const DevParams &P, Lane &L, int ix, int iy, int iz, Counters &c) {
    const DevWorld &w = P.w;
    const int cs = P.s.chunk_size, sh = P.cs_shift;
    L.cx = snap_rel<FAST>(ix, cs, sh); L.cy = snap_rel<FAST>(iy, cs, sh); L.cz = snap_rel<FAST>(iz, cs, sh);
    int slot = chunk_slot_rel<FAST>(w, cs, sh, L.cx, L.cy, L.cz);
    L.res = 0;
    if (COUNT) c.rechunk++;
    if (slot >= 0 && VT_IN(slot < w.ncx * w.ncy * w.ncz, 3)) { L.res = __ldg(&w.chunk_res[slot]); mark_visited(P.visited, slot); }
    else if (!FULL_TAIL) {
        bool gone = (ix >= w.nx && L.vx >= 0.0) || (ix < 0 && L.vx <= 0.0) || (iy >= w.ny && L.vy >= 0.0) ||
                    (iy < 0 && L.vy <= 0.0) || (iz >= w.nz && L.vz >= 0.0) || (iz < 0 && L.vz <= 0.0);
        if (gone) return false;
    }
    return true;
}

// void step of init.py:114 (see vt_device.cuh void_step); for a power-of-two chunk size a / cs == a * (1 / cs) exactly
template <bool FAST>
__device__ __forceinline__ double void_step2(double px, double py, double pz, int radius, int cs, int cs_shift) {
    // Python's min(): the first smallest argument; positions are never NaN, so two compares and selects do (fmin's
    // NaN handling costs 8 instructions each on this target)
    double a = pz < py ? pz : py;
    a = (a < px ? a : px) + i2d(radius);
    double m;
    if (FAST || cs_shift >= 0) {
        double dcs = i2d(cs), inv = __hiloint2double((1023 - cs_shift) << 20, 0);
        m = a - dcs * floor_d(a * inv);
    } else m = py_mod_pos(a, i2d(cs));
    return 1.0 + fabs(i2d(radius) - m);
}

__device__ __forceinline__ void advance(Lane &L, double step) {          // init.py:115-116
    L.step += step;
    L.px += L.vx * step; L.py += L.vy * step; L.pz += L.vz * step;
    if (!(L.step < L.life)) lane_set_state(L, ST_DONE);
}

// One iteration of the reference's while loop for a MARCH lane, up to but excluding the material call
// (init.py:66-77, 114-116), in two blocks the kernel separates with a __syncwarp so that the rare block (A: the
// ray left its chunk box) cannot drag a private copy of the common block (B: look at the cell, advance) behind it.
//
// Block A (init.py:67-73).  `um` >= chunk_size: not strictly inside the chunk box.  The reference's test is
// inclusive at chunk_max (lib.py:228-244): a position exactly on the far face(s) stays in the old chunk ("sticky");
// then the cell is looked up with the old chunk's frame right here and `mat` carries the answer to block B as
// mat + 1 (0 = not looked up yet).
template <bool FAST, bool FULL_TAIL, bool COUNT>
__device__ __forceinline__ int march_leave_chunk(const DevParams &P, Lane &L, Counters &c, int ix, int iy, int iz) {
    const DevWorld &w = P.w;
    const unsigned cs = (unsigned)P.s.chunk_size;
    const unsigned ux = (unsigned)ix - (unsigned)L.cx, uy = (unsigned)iy - (unsigned)L.cy, uz = (unsigned)iz - (unsigned)L.cz;
    bool sticky = false;
    if (max(ux, max(uy, uz)) == cs)
        sticky = (ux < cs || L.px == i2d(ix + w.ox)) && (uy < cs || L.py == i2d(iy + w.oy)) && (uz < cs || L.pz == i2d(iz + w.oz));
    if (sticky) return 1 + (L.res ? lookup_any<FAST>(P, L.cx, L.cy, L.cz, L.res, ix, iy, iz) : 0);
    if (!rechunk<FAST, FULL_TAIL, COUNT>(P, L, ix, iy, iz, c)) lane_set_state(L, ST_DONE);
    return 0;
}

// Block B (init.py:75-77, 114-116): examine the cell under the ray; a solid one turns the lane into HIT (the position
// is not advanced -- the material phase does that), an empty one advances the ray.
//
// Brick-local loop: in a full-resolution chunk the lane then stays in a tight loop for as long as its samples fall into
// the same 8^3 brick (see below) -- this is what shortens the long rays a packet waits for.
template <bool FAST, bool COUNT>
__device__ __forceinline__ void march_cell(const DevParams &P, Lane &L, Counters &c, int ix, int iy, int iz, int looked) {
    const DevWorld &w = P.w;
    if (L.res) {
        if (COUNT) c.in_chunk++;
        int mat = looked - 1;
        if (!looked) {
            mat = 0;
            int qx = ix, qy = iy, qz = iz;
            bool in = true;
            if (L.res > 1) {
                qx = lod_snap<FAST>(P, L.res, ix, 0); qy = lod_snap<FAST>(P, L.res, iy, 1); qz = lod_snap<FAST>(P, L.res, iz, 2);
                in = qx >= L.cx && qy >= L.cy && qz >= L.cz;            // q <= i, so only the near faces can fail
            }
            if (in && VT_IN((unsigned)(qx >> 3) < (unsigned)w.nbx && (unsigned)(qy >> 3) < (unsigned)w.nby && (unsigned)(qz >> 3) < (unsigned)w.nbz, 4)) {
                const uint32_t b = __ldg(&w.brick_tab[((qx >> 3) * w.nby + (qy >> 3)) * w.nbz + (qz >> 3)]);
                // Brick-local loop (full-resolution chunks): every further sample that stays in this 8^3 brick -- same
                // chunk (bricks never straddle chunks when chunk_size % 8 == 0), same table entry -- costs the position
                // update, three floors and, in a non-empty brick, one byte load, instead of a whole trip through the
                // state machine; the sample positions themselves are the reference's, one fp64 add per axis per sample.
                // The world origin is a multiple of 8 here, so brick membership is tested on absolute cells.
                if (L.res == 1 && P.brick_runs && VT_IN(b <= w.pool_slots, 5)) {
                    const int ax = ix + w.ox, ay = iy + w.oy, az = iz + w.oz;
                    // Two specialised loops.  Empty brick: nothing to load, the loop is the position update, the
                    // life test, three floors and the same-brick test.  Non-empty brick: the cell index is taken
                    // relative to the brick's corner -- jx * 64 + jy * 8 + jz minus the same expression of the corner,
                    // in wrapping 32-bit arithmetic (the true difference is 0..511) -- instead of three masks, two
                    // shifts and two ors per sample.
                    if (!b) {
                        for (;;) {
                            L.step += 1.0;
                            L.px += L.vx; L.py += L.vy; L.pz += L.vz;   // vel * 1.0 == vel
                            if (!(L.step < L.life)) { lane_set_state(L, ST_DONE); return; }
                            const int jx = floor_i(L.px), jy = floor_i(L.py), jz = floor_i(L.pz);
                            if ((((jx ^ ax) | (jy ^ ay) | (jz ^ az)) & ~7) != 0) return;      // next sample: another brick
                            if (COUNT) c.in_chunk++;
                        }
                    }
                    const uint8_t *pb = w.pool + (size_t)(b - 1u) * 512u;
                    const uint32_t corner = ((uint32_t)(ax & ~7) << 6) + ((uint32_t)(ay & ~7) << 3) + (uint32_t)(az & ~7);
                    mat = __ldg(pb + ((((uint32_t)ax << 6) + ((uint32_t)ay << 3) + (uint32_t)az) - corner));
                    while (!mat) {
                        L.step += 1.0;
                        L.px += L.vx; L.py += L.vy; L.pz += L.vz;
                        if (!(L.step < L.life)) { lane_set_state(L, ST_DONE); return; }
                        const int jx = floor_i(L.px), jy = floor_i(L.py), jz = floor_i(L.pz);
                        if ((((jx ^ ax) | (jy ^ ay) | (jz ^ az)) & ~7) != 0) return;
                        if (COUNT) c.in_chunk++;
                        mat = __ldg(pb + ((((uint32_t)jx << 6) + ((uint32_t)jy << 3) + (uint32_t)jz) - corner));
                    }
                } else if (b && VT_IN(b <= w.pool_slots, 5))
                    mat = __ldg(&w.pool[(size_t)(b - 1u) * 512u + (((qx & 7) << 6) | ((qy & 7) << 3) | (qz & 7))]);
            }
        }
        if (mat) { L.st = (L.st & ~0xfffu) | ST_HIT | ((uint32_t)mat << 2); return; }
        if (L.res == 1) {
            L.step += 1.0;
            L.px += L.vx; L.py += L.vy; L.pz += L.vz;
            if (!(L.step < L.life)) lane_set_state(L, ST_DONE);
            return;
        }
        advance(L, i2d(L.res));
    } else {
        if (COUNT) c.in_void++;
        advance(L, void_step2<FAST>(L.px, L.py, L.pz, P.s.chunk_radius, P.s.chunk_size, P.cs_shift));
    }
}

// lib.material (lib.py:448-460) + the post-hit part of the loop body (init.py:78-116) for a HIT lane.
template <bool FAST, bool REC, bool COUNT>
__device__ __forceinline__ void shade_one(const DevParams &P, Lane &L, FirstHit &fh, Counters &c) {
    const vt_settings &s = P.s;
    const DevWorld &w = P.w;
    const int cs = s.chunk_size, sh = P.cs_shift;
    const int mat = (int)((L.st >> 2) & 0xffu);
    const double *m = w.mats + 8 * mat;
    const int ix = floor_i(L.px) - w.ox, iy = floor_i(L.py) - w.oy, iz = floor_i(L.pz) - w.oz;
    if (REC) {
        if (!fh.fflag) { fh.fflag = 1; fh.fx = ix + w.ox; fh.fy = iy + w.oy; fh.fz = iz + w.oz; fh.first_step = L.step; }
        fh.hits++;
    }
    if (COUNT) c.hits++;
    // ---- lib.material ----
    double alb_r = __ldg(m + 0), alb_g = __ldg(m + 1), alb_b = __ldg(m + 2);
    double rough = __ldg(m + 3), absorb = __ldg(m + 4), emit = __ldg(m + 6);
    // pow(1, y) == 1 exactly: the first interaction of a ray (bounces == 0) needs no pow
    double pw = bounce_pow(P, L.bounces);
    double a = pmin(1.0, absorb / pw);
    double b2 = 1.0 - a;
    int cr = rint_i(i2d(L.col & 255) * b2 + alb_r * a);
    int cg = rint_i(i2d((L.col >> 8) & 255) * b2 + alb_g * a);
    int cb = rint_i(i2d((L.col >> 16) & 255) * b2 + alb_b * a);
    L.col = cr | (cg << 8) | (cb << 16);
    L.energy = L.energy * b2 + emit * a;
    L.life *= 1.0 - (rough * a);
    double jx, jy, jz;
    lane_rng_rand3(P, L, rough, jx, jy, jz);
    L.vx += jx; L.vy += jy; L.vz += jz;
    // ---- init.py:82-86 ----
    L.bounces += absorb;
    L.life /= i2d(L.res) + absorb * s.lod_bounces;
    double ref = pmax(fabs(L.vx), pmax(fabs(L.vy), fabs(L.vz)));             // vec3.normalize, lib.py:310-314
    if (ref != 0.0 && ref != 1.0) { L.vx /= ref; L.vy /= ref; L.vz /= ref; }
    if (L.step >= L.life || L.energy >= s.max_light || L.bounces >= s.max_bounces + 1.0) {
        lane_set_state(L, ST_DONE);                                          // `break`, init.py:85-86
        return;
    }
    double ior = __ldg(m + 5);
    if (ior != 0.0) {                                                        // init.py:92-111
        double direction = (ior - 0.5) * 2.0;
        unsigned flip = 0u;
#ifndef VT_PROBE_UNROLL
#define VT_PROBE_UNROLL 1            // the three neighbour probes as a loop: 312 instructions less, -2 % time (3: unrolled)
#endif
        _Pragma(VT_STR(unroll VT_PROBE_UNROLL))
        for (int k = 0; k < 3; k++) {
            double va = k == 0 ? L.vx : k == 1 ? L.vy : L.vz;
            double qa = k == 0 ? L.px : k == 1 ? L.py : L.pz;
            int oa = k == 0 ? w.ox : k == 1 ? w.oy : w.oz;
            qa = (va < direction) ? qa + 1.0 : qa - 1.0;
            int ja = floor_i(qa) - oa;
            int jx_ = k == 0 ? ja : ix, jy_ = k == 1 ? ja : iy, jz_ = k == 2 ? ja : iz;
            int lo = k == 0 ? L.cx : k == 1 ? L.cy : L.cz;
            unsigned u = (unsigned)(ja - lo);
            // `pos >= chunk_min and pos <= chunk_max` on the probed axis (the other two are the hit cell's own): the hit
            // chunk's frame answers; otherwise the frame of the chunk the probe falls into (Camera.chunk_get, init.py:28-33)
            int nx = L.cx, ny = L.cy, nz = L.cz, nres = L.res;
            if (!(u < (unsigned)cs || (u == (unsigned)cs && qa == i2d(ja + oa)))) {
                nx = snap_rel<FAST>(jx_, cs, sh); ny = snap_rel<FAST>(jy_, cs, sh); nz = snap_rel<FAST>(jz_, cs, sh);
                const int nslot = chunk_slot_rel<FAST>(w, cs, sh, nx, ny, nz);
                nres = (nslot >= 0 && VT_IN(nslot < w.ncx * w.ncy * w.ncz, 6)) ? (int)__ldg(&w.chunk_res[nslot]) : 0;
            }
            int nid = 0;
            if (nres) { nid = lookup_any<FAST>(P, nx, ny, nz, nres, jx_, jy_, jz_); if (COUNT) c.probes++; }
            if ((nid == 0) || (__ldg(w.mats + 8 * nid + 5) != ior)) flip |= 1u << k;
        }
        if (flip & 1u) L.vx -= L.vx * ior * 2.0;
        if (flip & 2u) L.vy -= L.vy * ior * 2.0;
        if (flip & 4u) L.vz -= L.vz * ior * 2.0;
    }
    lane_set_state(L, ST_MARCH);
    advance(L, i2d(L.res));                                               // init.py:114-116
}

// init.py:131-134: samples traced for pixel (x, y); dir_x / dir_y come from host tables (exact divisions)
__device__ __forceinline__ int pixel_samples(const DevParams &P, int x, int y, double &dir_x, double &dir_y, double &detail) {
    const vt_settings &s = P.s;
    if (!VT_IN((unsigned)x < (unsigned)s.width && (unsigned)y < (unsigned)s.height, 7)) { dir_x = dir_y = 0.0; detail = 1.0; return 1; }
    if (s.lod_edge == 0.0) { dir_x = dir_y = 0.0; detail = 1.0; return s.samples < 1 ? 1 : s.samples; }
    dir_x = __ldg(P.cam.dir_x + x);
    dir_y = __ldg(P.cam.dir_y + y);
    detail = 1.0 - fabs(dir_x * dir_y) * s.lod_edge;
    int n = rint_i(i2d(s.samples) * detail);
    return n < 1 ? 1 : n;
}

// End of a ray (init.py:119-121, 141-146): background, alpha, then add the sample to its pixel.  The sample that
// completes the pixel computes the mean (lib.average, lib.py:419-428), writes it and clears the accumulator.
template <bool FAST, bool REC>
__device__ __forceinline__ void finish_ray(const DevParams &P, Lane &L, const FirstHit &fh) {
    const vt_settings &s = P.s;
    double cr = i2d(L.col & 255), cg = i2d((L.col >> 8) & 255), cb = i2d((L.col >> 16) & 255);
    if (s.background) {                                                      // lib.material_background, lib.py:463-476
        double pw = bounce_pow(P, L.bounces);
        double a = pmin(1.0, 1.0 / pw);
        double up = pmax(0.0, L.vy);
        double b2 = 1.0 - a;
        cr = rint_d(cr * b2 + 127.0 * a);
        cg = rint_d(cg * b2 + (127.0 + up * 64.0) * a);
        cb = rint_d(cb * b2 + (127.0 + up * 128.0) * a);
        L.energy = mixd(L.energy, 1.0 + up, a);
        cr = pmin(255.0, rint_d(cr * L.energy));
        cg = pmin(255.0, rint_d(cg * L.energy));
        cb = pmin(255.0, rint_d(cb * L.energy));
    }
    if (REC && P.trace_out) {                                                // Camera.trace's return value (init.py:121)
        double *o = P.trace_out + (size_t)VT_TRACE_RECORD * L.out_i;
        o[0] = cr; o[1] = cg; o[2] = cb; o[3] = L.energy; o[4] = L.bounces; o[5] = L.step;
        o[6] = i2d(fh.hits); o[7] = i2d(fh.fx); o[8] = i2d(fh.fy); o[9] = i2d(fh.fz);
        o[10] = i2d(fh.fflag); o[11] = fh.first_step; o[12] = L.vx; o[13] = L.vy; o[14] = L.vz; o[15] = L.life;
        o[16] = L.px; o[17] = L.py; o[18] = L.pz; o[19] = 0.0;
        return;
    }
    double alpha = rint_d(pmin(1.0, L.energy + s.shutter) * 255.0);            // init.py:141
    const int x = (int)(L.xy & 0xffffu), y = (int)(L.xy >> 16);
    const int n = (int)L.nsmp;
    const size_t i = L.out_i;
    if (!VT_IN((long long)i < P.n_pixels, 8)) return;
    if (REC && lane_sample(L) == 0u) {
        double *o = P.ray0 + (size_t)VT_RAY_RECORD * i;
        o[0] = cr; o[1] = cg; o[2] = cb; o[3] = L.energy; o[4] = L.bounces; o[5] = L.step;
        o[6] = i2d(fh.hits); o[7] = i2d(fh.fx); o[8] = i2d(fh.fy); o[9] = i2d(fh.fz);
        o[10] = i2d(fh.fflag); o[11] = fh.first_step; o[12] = L.vx; o[13] = L.vy; o[14] = L.vz; o[15] = i2d(n);
    }
    // colour channels can be negative only through a negative albedo/energy; clamp like the u8 store would not --
    // the accumulator holds non-negative integers, so clamp at 0 here and keep the mean of the clamped values
    unsigned long long ur = (unsigned)rint_i(pmax(cr, 0.0)), ug = (unsigned)rint_i(pmax(cg, 0.0));
    unsigned long long ub = (unsigned)rint_i(pmax(cb, 0.0)), ua = (unsigned)rint_i(pmax(alpha, 0.0));
    unsigned long long sr, sg, sb, sa;
    bool last;
    if (FAST || !P.acc_wide) {
        unsigned long long add = ur | (ug << 14) | (ub << 28) | (ua << 42) | (1ull << 56);
        unsigned long long tot = atomicAdd(P.acc + i, add) + add;
        last = (int)(tot >> 56) == n;
        sr = tot & 0x3fffull; sg = (tot >> 14) & 0x3fffull; sb = (tot >> 28) & 0x3fffull; sa = (tot >> 42) & 0x3fffull;
        if (last) P.acc[i] = 0ull;
    } else {
        atomicAdd(P.acc + 2 * i, ur | (ug << 21) | (ub << 42));
        __threadfence();
        unsigned long long add = ua | (1ull << 32);
        unsigned long long tot = atomicAdd(P.acc + 2 * i + 1, add) + add;
        last = (int)(tot >> 32) == n;
        sa = tot & 0xffffffffull;
        sr = sg = sb = 0ull;
        if (last) {
            __threadfence();
            unsigned long long rgb = atomicExch(P.acc + 2 * i, 0ull);
            sr = rgb & 0x1fffffull; sg = (rgb >> 21) & 0x1fffffull; sb = (rgb >> 42) & 0x1fffffull;
            P.acc[2 * i + 1] = 0ull;
        }
    }
    if (last) {
        double mr, mg, mb, ma;
        if (VT_INVN && P.inv_samples != 0.0) {                               // n is a power of two: x / n == x * (1 / n) exactly
            mr = u2d((uint32_t)sr) * P.inv_samples; mg = u2d((uint32_t)sg) * P.inv_samples;
            mb = u2d((uint32_t)sb) * P.inv_samples; ma = u2d((uint32_t)sa) * P.inv_samples;
        } else {
            const double dn = i2d(n);
            mr = u2d((uint32_t)sr) / dn; mg = u2d((uint32_t)sg) / dn; mb = u2d((uint32_t)sb) / dn; ma = u2d((uint32_t)sa) / dn;
        }
        if (P.mean) reinterpret_cast<float4 *>(P.mean)[i] = make_float4((float)mr, (float)mg, (float)mb, (float)ma);
        if (P.rgba8 || P.frame_target) {
            uchar4 q;                                                        // truncation, like a C cast
            q.x = (unsigned char)floor_i(pmin(255.0, mr)); q.y = (unsigned char)floor_i(pmin(255.0, mg));
            q.z = (unsigned char)floor_i(pmin(255.0, mb)); q.w = (unsigned char)floor_i(pmin(255.0, ma));
            // frame target (multi-GPU): the pixel goes straight to its place in the full frame, which may live on a
            // peer GPU (NVLink store), instead of this rank's compact band
            const size_t o = P.frame_target ? (size_t)y * (size_t)s.width + (size_t)x : i;
            if (!VT_IN(!P.frame_target || o < (size_t)s.width * (size_t)s.height, 9)) return;
            reinterpret_cast<uchar4 *>(P.frame_target ? P.frame_target : P.rgba8)[o] = q;
        }
    }
}

// Start of a ray (init.py:131-139, 41-59) for ray slot `id`.  Slots enumerate (pixel tile, sample round, lane):
// 32 consecutive slots are the G samples of 32/G neighbouring pixels, so lanes refilled together start coherent.
template <bool FAST, bool REC, bool FULL_TAIL, bool COUNT>
__device__ __forceinline__ void start_ray(const DevParams &P, Lane &L, FirstHit &fh, Counters &c, unsigned long long id,
                                          const int32_t *__restrict__ row_y, unsigned &rays) {
    const vt_settings &s = P.s;
    L.st = ST_NONE;
    const uint32_t packet = (uint32_t)(id >> 5), slot = (uint32_t)id & 31u;
    uint32_t tile = packet, round = 0;
    if (P.rounds > 1) { tile = packet / (uint32_t)P.rounds; round = packet - tile * (uint32_t)P.rounds; }
    const uint32_t pi = slot >> P.group_shift, sample = (slot & (uint32_t)(P.group - 1)) + (uint32_t)P.group * round;
    int x = 0, y = 0;
    bool pv;
    const bool tracing = REC && P.dirs != nullptr;                           // vt_trace_rays: slot i is ray i
    if (tracing) {
        const long long i = (long long)tile * 32 + pi;
        pv = i < P.n_pixels;
        // the Philox key of ray i is i itself: y * width + x below
        if (pv) { y = (int)(i / s.width); x = (int)(i - (long long)y * s.width); L.out_i = (uint32_t)i; }
    } else if (P.pixels) {
        long long i = (long long)tile * (32 >> P.group_shift) + pi;
        pv = i < P.n_pixels;
        if (pv) { x = __ldg(P.pixels + 2 * i); y = __ldg(P.pixels + 2 * i + 1); L.out_i = (uint32_t)i; }
    } else {
        const int bw = P.x1 - P.x0;
        uint32_t ty = tile / (uint32_t)P.tiles_x, tx = tile - ty * (uint32_t)P.tiles_x;
        int lx = pi & (P.tile_w - 1), ly = pi >> P.tile_w_shift;
        int col = (int)tx * P.tile_w + lx, lr = (int)ty * P.tile_h + ly;
        pv = col < bw && lr < P.band_rows;
        if (pv && VT_IN(lr >= 0 && col >= 0, 11)) { x = P.x0 + col; y = __ldg(row_y + lr); L.out_i = (uint32_t)(lr * bw + col); }
    }
    if (!pv) return;
    double dir_x, dir_y, detail;
    int n = 1;
    if (tracing) { dir_x = __ldg(P.dirs + 3 * (size_t)L.out_i); dir_y = __ldg(P.dirs + 3 * (size_t)L.out_i + 1); detail = __ldg(P.dirs + 3 * (size_t)L.out_i + 2); }
    else n = pixel_samples(P, x, y, dir_x, dir_y, detail);
    if ((int)sample >= n) return;
    rays++;
    L.xy = (uint32_t)x | ((uint32_t)y << 16);
    L.st = ST_MARCH | (sample << 12);
    L.nsmp = (uint32_t)n;
    L.draw = 0; L.w2 = L.w3 = 0;
    double ray_detail = detail;                                              // trace() takes `detail` as given (init.py:56)
    double jit_x = 0.0, jit_y = 0.0;                                         // rand(dof), rand(dof) of init.py:41-42
    if (!tracing) {
        // Draws 0, 1, 2 of the ray's stream: lod_random (drawn even when it is 0, init.py:139), then the two lens jitters
        // (rand(0) draws nothing, lib.py:431-434).  Written out block by block -- the generic draw-by-draw helper inlined
        // three times cost 60 instructions more.
        const uint32_t pix = (uint32_t)y * (uint32_t)s.width + (uint32_t)x;
        const uint4 w0 = philox_block(0u, sample, P.seed_lo, P.seed_hi, pix);
        const double u = u53m(w0.x, w0.y);
        L.draw = 1; L.w2 = w0.z; L.w3 = w0.w;
        if (s.dof != 0.0) {
            jit_x = (-1.0 + u53m(w0.z, w0.w) * 2.0) * s.dof;
            const uint4 w1 = philox_block(1u, sample, P.seed_lo, P.seed_hi, pix);
            jit_y = (-1.0 + u53m(w1.x, w1.y) * 2.0) * s.dof;
            L.draw = 3; L.w2 = w1.z; L.w3 = w1.w;
        }
        ray_detail = detail / (1.0 + u2d(sample) * s.lod_samples) * (1.0 - s.lod_random * u);
    } else {
        jit_x = lane_rng_rand(P, L, s.dof); jit_y = lane_rng_rand(P, L, s.dof);
    }
    double sy, cy, sz, cz;
    if (tracing && s.dof == 0.0 && P.dir_sc) {
        const double *sc = P.dir_sc + 4 * (size_t)L.out_i;
        sy = __ldg(sc); cy = __ldg(sc + 1); sz = __ldg(sc + 2); cz = __ldg(sc + 3);
    } else if (!tracing && s.dof == 0.0 && P.cam.col_sc) {
        sy = __ldg(P.cam.col_sc + 2 * x); cy = __ldg(P.cam.col_sc + 2 * x + 1);
        sz = __ldg(P.cam.row_sc + 2 * y); cz = __ldg(P.cam.row_sc + 2 * y + 1);
    } else {
        const double deg2rad = 3.14159265358979323846 / 180.0;
        // (dir_x / proportions) * lens depends on the column only, (dir_y * proportions) * lens on the row only: host tables
        double lens_x = (tracing ? (dir_x / s.proportions) * P.cam.lens : __ldg(P.cam.lens_x + x)) + jit_x;
        double lens_y = (tracing ? (dir_y * s.proportions) * P.cam.lens : __ldg(P.cam.lens_y + y)) + jit_y;
        double rad_y = (-lens_x) * deg2rad, rad_z = (+lens_y) * deg2rad;
        double2 scy = sincos_out(rad_y / 2.0), scz = sincos_out(rad_z / 2.0);
        sy = scy.x; cy = scy.y; sz = scz.x; cz = scz.y;
    }
    double dx, dy, dz;
    ray_dir(P.cam, sy, cy, sz, cz, dx, dy, dz);
    L.px = P.cam.pos[0] + dx * s.dist_min; L.py = P.cam.pos[1] + dy * s.dist_min; L.pz = P.cam.pos[2] + dz * s.dist_min;
    L.vx = dx; L.vy = dy; L.vz = dz;
    L.life = (s.dist_max - s.dist_min) * ray_detail;
    L.step = 0.0; L.bounces = 0.0; L.energy = 0.0; L.col = 0;
    // The loop starts with chunk_min == chunk_max == (0, 0, 0) and no chunk (init.py:60), so its first iteration
    // re-evaluates the chunk: the box is set to one no cell can lie in and the first march trip does exactly that, for
    // the whole packet at once (a second inlined copy of the chunk lookup here cost 100 instructions of code).
    L.cx = L.cy = L.cz = VT_NO_BOX; L.res = 0;
    if (REC) { fh.hits = 0; fh.fflag = 0; fh.fx = fh.fy = fh.fz = 0; fh.first_step = -1.0; }
    if (!(L.step < L.life)) { lane_set_state(L, ST_DONE); return; }
    // A ray that starts exactly at the world's zero point is inside that initial box: one void step before its first
    // chunk lookup.
    if (L.px == 0.0 && L.py == 0.0 && L.pz == 0.0) {
        if (COUNT) c.in_void++;
        advance(L, void_step2<FAST>(L.px, L.py, L.pz, s.chunk_radius, s.chunk_size, P.cs_shift));
    }
}

// Camera.tile (reference init.py:126-150) for a whole region: one persistent launch.
template <bool FAST, bool REC, bool FULL_TAIL, bool COUNT>
__global__ void __launch_bounds__(128, VT_MINB) render_persistent(const __grid_constant__ DevParams P, const int32_t *__restrict__ row_y) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
#if VT_COLD_SMEM
    __shared__ ColdStore cold;
#else
    ColdStore &cold = *(ColdStore *)nullptr;
#endif
    Lane L(cold, threadIdx.x);
    L.st = ST_NONE; L.px = L.py = L.pz = L.vx = L.vy = L.vz = 0.0; L.step = L.life = L.bounces = L.energy = 0.0;
    L.col = 0; L.cx = L.cy = L.cz = 0; L.res = 0; L.xy = 0; L.out_i = 0; L.draw = L.w2 = L.w3 = 0; L.nsmp = 1;
    FirstHit fh; fh.hits = fh.fflag = fh.fx = fh.fy = fh.fz = 0; fh.first_step = -1.0;
    Counters c = {0, 0, 0, 0, 0, 0};
    unsigned rays = 0;
#if VT_TRIPS
    unsigned trips[6] = {0, 0, 0, 0, 0, 0};   // march trips, shade trips, shade lanes, refill trips, refill lanes, march phases
    long long cyc[4] = {0, 0, 0, 0}, tc0 = 0;  // cycles of this warp in refill, shade, march, block A of march
    long long bk[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // march trips / cycles by marching lanes: <= 4, <= 8, <= 16, <= 32
#define VT_T0() (tc0 = clock64())
#define VT_T1(k) (cyc[k] += clock64() - tc0)
#else
#define VT_T0()
#define VT_T1(k)
#endif
    bool exhausted = false;                                                  // warp-uniform: the queue is empty
    for (;;) {
        const uint32_t state = lane_state(L);
        const unsigned mM = __ballot_sync(full, state == ST_MARCH), mH = __ballot_sync(full, state == ST_HIT);
        const unsigned mD = __ballot_sync(full, state == ST_DONE);
        const int nM = __popc(mM), nH = __popc(mH), nD = __popc(mD);
        const int nI = exhausted ? nD : 32 - nM - nH;                        // lanes a refill round would serve
        if (nM + nH == 0 && nI == 0) break;
        if ((nI >= P.refill_min && nI > nM && nI > nH) || nM + nH == 0) {
            // ---- refill: finish ended rays, fetch new ray slots --------------------------------------------
#if VT_TRIPS
            if (COUNT && lane == 0) { trips[3]++; trips[4] += nI; }
#endif
            VT_T0();
            if (state == ST_DONE) { finish_ray<FAST, REC>(P, L, fh); L.st = ST_NONE; }
            if (!exhausted) {
                const unsigned need = __ballot_sync(full, lane_state(L) == ST_NONE);
                const int cnt = __popc(need), leader = __ffs(need) - 1;
                unsigned long long base = 0;
                if (lane == leader) base = atomicAdd(P.queue, (unsigned long long)cnt);
                base = __shfl_sync(full, base, leader);
                if (lane_state(L) == ST_NONE) {
                    unsigned long long id = base + (unsigned)__popc(need & ((1u << lane) - 1u));
                    if (id < P.total_rays) start_ray<FAST, REC, FULL_TAIL, COUNT>(P, L, fh, c, id, row_y, rays);
                }
                exhausted = base + (unsigned)cnt >= P.total_rays;
            }
            VT_T1(0);
        } else if (nH > nM) {
#if VT_TRIPS
            if (COUNT && lane == 0) { trips[1]++; trips[2] += nH; }
#endif
            VT_T0();
            if (state == ST_HIT) shade_one<FAST, REC, COUNT>(P, L, fh, c);
#if VT_TRIPS
            __syncwarp(full);
#endif
            VT_T1(1);
        } else {
            // ---- march: iterate until too few lanes are still marching -------------------------------------
            const int thr = max(1, (nM * VT_MARCH_KEEP) >> 3);
#if VT_TRIPS
            if (COUNT && lane == 0) trips[5]++;
            const long long tm0 = clock64();
#endif
            for (;;) {
#if VT_TRIPS
                if (COUNT && lane == 0) trips[0]++;
                const int nm_ = __popc(__ballot_sync(full, lane_state(L) == ST_MARCH));
                const int b_ = nm_ <= 4 ? 0 : nm_ <= 8 ? 1 : nm_ <= 16 ? 2 : 3;
                const long long tt0 = clock64();
#endif
                int ix = 0, iy = 0, iz = 0, looked = 0;
                bool leave = false;
                if (lane_state(L) == ST_MARCH) {
                    ix = floor_i(L.px) - P.w.ox; iy = floor_i(L.py) - P.w.oy; iz = floor_i(L.pz) - P.w.oz;
                    const unsigned ux = (unsigned)ix - (unsigned)L.cx, uy = (unsigned)iy - (unsigned)L.cy, uz = (unsigned)iz - (unsigned)L.cz;
                    leave = max(ux, max(uy, uz)) >= (unsigned)P.s.chunk_size;
                }
                VT_T0();
                if (leave) looked = march_leave_chunk<FAST, FULL_TAIL, COUNT>(P, L, c, ix, iy, iz);
                __syncwarp(full);
                VT_T1(3);
                if (lane_state(L) == ST_MARCH) march_cell<FAST, COUNT>(P, L, c, ix, iy, iz, looked);
#if VT_TRIPS
                __syncwarp(full);
                bk[b_]++; bk[4 + b_] += clock64() - tt0;
#endif
                if (__popc(__ballot_sync(full, lane_state(L) == ST_MARCH)) < thr) break;
            }
#if VT_TRIPS
            cyc[2] += clock64() - tm0;
#endif
        }
    }
    if (COUNT && P.counters) {
        unsigned v[7] = {rays, c.in_chunk, c.in_void, c.hits, c.probes, c.rechunk, c.in_chunk + c.in_void};
#pragma unroll
        for (int k = 0; k < 7; k++) {
            unsigned t = __reduce_add_sync(full, v[k]);
            if (lane == 0 && t) atomicAdd(P.counters + k, (unsigned long long)t);
        }
#if VT_TRIPS
        if (lane == 0) {
            for (int k = 0; k < 6; k++) atomicAdd(P.counters + 8 + k, (unsigned long long)trips[k]);
            for (int k = 0; k < 4; k++) atomicAdd(P.counters + 14 + k, (unsigned long long)cyc[k]);
            for (int k = 0; k < 8; k++) atomicAdd(P.counters + 18 + k, (unsigned long long)bk[k]);
        }
#endif
    }
    // The last warp to leave resets the queue for the next launch and, with a frame target, publishes the arrival
    // flag: every thread's pixel stores are fenced at system scope first, so the peer that sees the flag sees the band.
    if (P.arrive_flag) { __threadfence_system(); __syncwarp(full); }
    unsigned long long done = 0ull;
    if (lane == 0) {
        __threadfence();
        done = atomicAdd(P.queue + 1, 1ull);
    }
    done = __shfl_sync(full, done, 0);
    if (done == (unsigned long long)P.total_warps - 1ull) {
        // Multi-GPU culling (init.py:189, 393: the workers' traversed lists are merged every frame): this rank's visited
        // words -- complete now, every other warp has fenced and left -- are OR-ed into the peer's mask over NVLink,
        // and the arrival flag that publishes the pixels publishes them too.
        if (P.visited_peer && P.visited) {
            __threadfence();
            for (int w = lane; w < P.visited_words; w += 32) {
                const uint32_t v = __ldcg(P.visited + w);
                if (v) atomicOr_system(P.visited_peer + w, v);
            }
            __threadfence_system();
            __syncwarp(full);
        }
        if (lane == 0) {
            P.queue[0] = 0ull; P.queue[1] = 0ull;
            if (P.arrive_flag) { __threadfence_system(); atomicExch_system(P.arrive_flag, P.arrive_value); }
        }
    }
}

}  // namespace vt
