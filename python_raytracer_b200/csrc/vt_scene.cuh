// vt_scene.cuh — scene-building kernels for sm_100a: what Window.chunk_update (reference init.py:389-452) does before
// the trace loop, as streaming passes over u8 cells.  All of them are HBM-bound byte work: cells are
// moved eight at a time along z (the fastest axis of the dense grid), every global access of a warp is one contiguous
// 256- or 512-byte run, and the slot numbering of the brick pool is a device-side scan (no host round trip over the
// brick table).
//
//   build_cells_kernel    objects -> merged dense world (Sprite.pos_rotated, later objects override), whole world
//                          or one box of it; one thread per 8 cells, one 8-byte store
//   brick_mask_kernel     dense world -> one occupancy bit per 8^3 brick; a warp reads the 64 rows of 32 bricks
//   task_scan_kernel      exclusive scan of the per-task brick counts: pool slots in brick-index order
//   brick_pack_kernel     dense world -> brick pool (+ brick table), rows staged through shared memory so that
//                          every brick leaves as one 512-byte store of a warp
//   chunk_present_kernel  brick table -> "chunk holds a voxel" per chunk slot (chunk_size % 8 == 0)
//   region_flags_kernel / region_fill_kernel / scatter_bricks_kernel: the small in-place updates (moving objects,
//                          dirty bricks from the host); latency-bound, a few launches of a few blocks
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/voxtrace.h"

namespace vt {

// Eight cells z0 .. z0 + 7 of one row of a [..][..][nz] grid, packed little-endian (byte c = cell z0 + c); cells past
// nz read as 0.  `vec`: nz % 8 == 0 and the grid is 8-byte aligned, so the run is one aligned 8-byte load.
__device__ __forceinline__ uint2 load_run8(const uint8_t *__restrict__ row, int z0, int nz, bool vec) {
    if (vec) return __ldg(reinterpret_cast<const uint2 *>(row + z0));
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) {
        if (z0 + c < nz) lo |= (uint32_t)__ldg(row + z0 + c) << (8 * c);
        if (z0 + 4 + c < nz) hi |= (uint32_t)__ldg(row + z0 + 4 + c) << (8 * c);
    }
    return make_uint2(lo, hi);
}
__device__ __forceinline__ void store_run8(uint8_t *__restrict__ row, int z0, int nz, bool vec, uint2 v) {
    if (vec) { *reinterpret_cast<uint2 *>(row + z0) = v; return; }
#pragma unroll
    for (int c = 0; c < 4; c++) {
        if (z0 + c < nz) row[z0 + c] = (uint8_t)(v.x >> (8 * c));
        if (z0 + 4 + c < nz) row[z0 + 4 + c] = (uint8_t)(v.y >> (8 * c));
    }
}

// Sprite.pos_rotated (reference data.py:338-371): 90-degree steps, each axis only when the two other sizes match
__device__ __forceinline__ void pos_rotated(const vt_object &o, int &x, int &y, int &z) {
    const int ex = o.size[0] - 1, ey = o.size[1] - 1, ez = o.size[2] - 1;
    int ax = o.angle[0], ay = o.angle[1], az = o.angle[2], nx, ny, nz;
    if (ax && o.size[1] == o.size[2]) {
        if (ax == 1) { ny = ez - z; nz = y; } else if (ax == 2) { ny = ey - y; nz = ez - z; } else { ny = z; nz = ey - y; }
        y = ny; z = nz;
    }
    if (ay && o.size[0] == o.size[2]) {
        if (ay == 1) { nx = z; nz = ex - x; } else if (ay == 2) { nx = ex - x; nz = ez - z; } else { nx = ez - z; nz = x; }
        x = nx; z = nz;
    }
    if (az && o.size[0] == o.size[1]) {
        if (az == 1) { nx = ey - y; ny = x; } else if (az == 2) { nx = ex - x; ny = ey - y; } else { nx = y; ny = ex - x; }
        x = nx; y = ny;
    }
}

// 0xff in every byte of the result whose byte in v is not zero
__device__ __forceinline__ unsigned long long nonzero_bytes(unsigned long long v) {
    const unsigned long long lo7 = 0x7f7f7f7f7f7f7f7full;
    unsigned long long t = (((v & lo7) + lo7) | v) & ~lo7;               // 0x80 where the byte is non-zero
    return t | (t - (t >> 7));
}

// Window.chunk_update's voxel gathering (init.py:413-424, 437-439) for the eight world cells (wx, wy, wz0 .. wz0 + 7):
// the last object in list order that has a voxel at a cell wins.  Objects are tested against the whole run first
// (point test in x and y, interval overlap in z); only the cells still empty are taken from the next object.
// `need` holds 0xff for every cell that is still empty.  An unrotated object whose box covers the whole run gives
// its eight cells with two aligned 8-byte loads and a funnel shift (sprite rows start at any byte); everything else
// -- rotations, runs that straddle the box, the last bytes of a sprite -- goes cell by cell.
__device__ __forceinline__ uint2 gather_run8(const vt_object *__restrict__ objs, int n, int wx, int wy, int wz0) {
    unsigned long long out = 0ull, need = ~0ull;
    for (int k = n - 1; k >= 0 && need; k--) {
        const vt_object &o = objs[k];
        if (wx < o.mins[0] || wx >= o.maxs[0] || wy < o.mins[1] || wy >= o.maxs[1]) continue;
        const int c0 = max(0, o.mins[2] - wz0), c1 = min(8, o.maxs[2] - wz0);
        if (c0 >= c1) continue;
        const bool plain = (o.angle[0] | o.angle[1] | o.angle[2]) == 0;
        const int bx = wx - o.mins[0], by = wy - o.mins[1], bz = wz0 - o.mins[2];
        const int s0 = o.size[0], s1 = o.size[1], s2 = o.size[2];
        if (plain && c0 == 0 && c1 == 8 && (unsigned)bx < (unsigned)s0 && (unsigned)by < (unsigned)s1 && bz >= 0 && bz + 8 <= s2 &&
            (reinterpret_cast<uintptr_t>(o.dev_sprite) & 7u) == 0) {
            const size_t off = ((size_t)bx * s1 + by) * s2 + bz;
            if (off + 16 <= (size_t)s0 * s1 * s2) {                       // both aligned words lie inside the sprite
                const unsigned long long *p = reinterpret_cast<const unsigned long long *>(o.dev_sprite + (off & ~(size_t)7));
                const int sh = (int)(off & 7) * 8;
                unsigned long long v = __ldg(p);
                if (sh) v = (v >> sh) | (__ldg(p + 1) << (64 - sh));
                const unsigned long long m = nonzero_bytes(v) & need;
                out |= v & m;
                need &= ~m;
                continue;
            }
        }
#pragma unroll
        for (int c = 0; c < 8; c++) {
            if (c < c0 || c >= c1 || !((need >> (8 * c)) & 1ull)) continue;
            int lx = bx, ly = by, lz = bz + c;
            if (!plain) pos_rotated(o, lx, ly, lz);
            if ((unsigned)lx >= (unsigned)s0 || (unsigned)ly >= (unsigned)s1 || (unsigned)lz >= (unsigned)s2) continue;
            const unsigned long long m = __ldg(o.dev_sprite + ((size_t)lx * s1 + ly) * s2 + lz);
            if (m) { out |= m << (8 * c); need &= ~(0xffull << (8 * c)); }
        }
    }
    return make_uint2((uint32_t)out, (uint32_t)(out >> 32));
}

// Cells of the box [l, l + e) of the dense grid [..][ny][nz] (the whole world: l = 0, e = dims), one thread per run
// of eight cells along z; l[2] is a multiple of 8.  (ox, oy, oz) = world cell of grid cell (0, 0, 0).
__global__ void __launch_bounds__(256) build_cells_kernel(const vt_object *__restrict__ objs, int n, int ox, int oy, int oz, int ny, int nz,
                                                          int lx0, int ly0, int lz0, int ex, int ey, int ez, bool vec,
                                                          uint8_t *__restrict__ cells) {
    const int runs = (ez + 7) >> 3;
    const long long total = (long long)ex * ey * runs;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        int x, y, r;
        if (total <= 0x7fffffffLL) {                                      // 32-bit divisions where they suffice
            const unsigned u = (unsigned)i, row = u / (unsigned)runs;
            r = (int)(u - row * (unsigned)runs); x = (int)(row / (unsigned)ey); y = (int)(row - (unsigned)x * (unsigned)ey);
        } else {
            const long long row = i / runs;
            r = (int)(i - row * runs); x = (int)(row / ey); y = (int)(row - (long long)x * ey);
        }
        x += lx0; y += ly0;
        const int z0 = lz0 + 8 * r;
        uint2 v = gather_run8(objs, n, x + ox, y + oy, z0 + oz);
        const int zend = lz0 + ez;                                        // cells of the run past the box keep their value
        uint8_t *row_p = cells + ((size_t)x * ny + y) * nz;
        if (z0 + 8 <= zend) store_run8(row_p, z0, nz, vec, v);
        else store_run8(row_p, z0, zend < nz ? zend : nz, false, v);
    }
}

// ---- the same world build, wide variant -------------------------------------------------------------------------------
// build_cells_kernel spends ~200 warp instructions per 32 runs of 8 cells: a fifth in the two integer divisions of the index
// decode, the rest in re-reading the object record and in one short, latency-exposed load chain per thread (ncu, round 2:
// issue slots 70 % busy, long_scoreboard 5 of 12 stall cycles, 1.24 TB/s).  This variant serves worlds whose rows are
// multiples of 16 cells and 16-byte aligned: one thread makes 16 cells of ROWS consecutive rows (same z, y counting up), so
// that the decode (two multiply-high divisions with host-made multipliers) and every object record (four 16-byte loads) are
// paid once per ROWS x 16 cells, the sprite words of all ROWS rows are requested before any is used, and rows leave as
// 16-byte stores.  Cell values are exactly build_cells_kernel's (same object order, same pos_rotated path for the rest).
struct ObjRegs {
    const uint8_t *sp;
    int s0, s1, s2, mn0, mn1, mn2, mx0, mx1, mx2, a0, a1, a2;
};
__device__ __forceinline__ ObjRegs load_object(const vt_object *__restrict__ o) {      // 64-byte records, 16-byte aligned
    const int4 q0 = __ldg(reinterpret_cast<const int4 *>(o)), q1 = __ldg(reinterpret_cast<const int4 *>(o) + 1);
    const int4 q2 = __ldg(reinterpret_cast<const int4 *>(o) + 2), q3 = __ldg(reinterpret_cast<const int4 *>(o) + 3);
    ObjRegs r;
    r.sp = reinterpret_cast<const uint8_t *>(((unsigned long long)(unsigned)q0.y << 32) | (unsigned)q0.x);
    r.s0 = q0.z; r.s1 = q0.w; r.s2 = q1.x; r.mn0 = q1.y; r.mn1 = q1.z; r.mn2 = q1.w;
    r.mx0 = q2.x; r.mx1 = q2.y; r.mx2 = q2.z; r.a0 = q2.w; r.a1 = q3.x; r.a2 = q3.y;
    return r;
}
// floor(u / d) for u < 2^32 with m = floor(2^64 / d) + 1 (d >= 2); exact: the error term u * (m d - 2^64) stays below 2^64
__device__ __forceinline__ unsigned div_magic(unsigned u, unsigned long long m, unsigned d) {
    return d == 1u ? u : (unsigned)__umul64hi((unsigned long long)u, m);
}

// the 8-cell run at (wx, wy, wz0) from one object, cell by cell (rotations, box edges, sprite ends): the rare path, one
// out-of-line copy (inlined into the unrolled row loop it cost 117 registers and 6 300 instructions).  -> {out, need}
__device__ __noinline__ ulonglong2 object_cells8(const vt_object *__restrict__ orec, int wx, int wy, int wz0, unsigned long long out,
                                                 unsigned long long need) {
    const vt_object &o = *orec;
    const int c0 = max(0, o.mins[2] - wz0), c1 = min(8, o.maxs[2] - wz0);
    if (c0 >= c1) return make_ulonglong2(out, need);
    const bool plain = (o.angle[0] | o.angle[1] | o.angle[2]) == 0;
    const int bx = wx - o.mins[0], by = wy - o.mins[1], bz = wz0 - o.mins[2];
    const int s0 = o.size[0], s1 = o.size[1], s2 = o.size[2];
    for (int c = c0; c < c1; c++) {
        if (!((need >> (8 * c)) & 1ull)) continue;
        int lx = bx, ly = by, lz = bz + c;
        if (!plain) pos_rotated(o, lx, ly, lz);
        if ((unsigned)lx >= (unsigned)s0 || (unsigned)ly >= (unsigned)s1 || (unsigned)lz >= (unsigned)s2) continue;
        const unsigned long long m = __ldg(o.dev_sprite + ((size_t)lx * s1 + ly) * s2 + lz);
        if (m) { out |= m << (8 * c); need &= ~(0xffull << (8 * c)); }
    }
    return make_ulonglong2(out, need);
}

#ifndef VT_BC_MINB
#define VT_BC_MINB 2                 // resident 256-thread blocks per SM the wide world build is sized for
#endif
template <int ROWS>
__global__ void __launch_bounds__(256, VT_BC_MINB) build_cells16_kernel(const vt_object *__restrict__ objs, int n, int ox, int oy, int oz, int ny, int nz,
                                                            int lx0, int ly0, int lz0, int ex, int ey, int chunks, unsigned long long m_chunks,
                                                            unsigned long long m_ey, uint8_t *__restrict__ cells) {
    const unsigned n_rows = (unsigned)ex * (unsigned)ey, groups = (n_rows + ROWS - 1) / ROWS;
    const unsigned total = groups * (unsigned)chunks;                     // < 2^31 (host)
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const unsigned g = div_magic(i, m_chunks, (unsigned)chunks), c = i - g * (unsigned)chunks;
        const unsigned row0 = g * ROWS;
        const unsigned x0 = div_magic(row0, m_ey, (unsigned)ey);
        int x = (int)x0, y = (int)(row0 - x0 * (unsigned)ey);
        const int z0 = lz0 + 16 * (int)c, wz0 = z0 + oz;
        int wx[ROWS], wy[ROWS];
        bool valid[ROWS];
        unsigned long long out0[ROWS], out1[ROWS], need0[ROWS], need1[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; r++) {
            valid[r] = row0 + r < n_rows;
            wx[r] = x + lx0 + ox; wy[r] = y + ly0 + oy;
            out0[r] = out1[r] = 0ull;
            need0[r] = need1[r] = valid[r] ? ~0ull : 0ull;
            if (++y == ey) { y = 0; x++; }
        }
        for (int k = n - 1; k >= 0; k--) {
            unsigned long long any = 0ull;
#pragma unroll
            for (int r = 0; r < ROWS; r++) any |= need0[r] | need1[r];
            if (!any) break;
            // quick reject on the box alone (two of the record's four words): most objects of a many-object world miss the
            // thread's 16 cells in z, its x (rows of a thread span at most two x) or its y range
            {
                const int4 q1 = __ldg(reinterpret_cast<const int4 *>(objs + k) + 1), q2 = __ldg(reinterpret_cast<const int4 *>(objs + k) + 2);
                const int mn0 = q1.y, mn1 = q1.z, mn2 = q1.w, mx0 = q2.x, mx1 = q2.y, mx2 = q2.z;
                if (mn2 >= wz0 + 16 || mx2 <= wz0) continue;
                if ((wx[0] < mn0 || wx[0] >= mx0) && (wx[ROWS - 1] < mn0 || wx[ROWS - 1] >= mx0)) continue;
                if (wx[0] == wx[ROWS - 1] && (wy[ROWS - 1] < mn1 || wy[0] >= mx1)) continue;
            }
            const ObjRegs o = load_object(objs + k);
            const int c0 = max(0, o.mn2 - wz0), c1 = min(16, o.mx2 - wz0);
            if (c0 >= c1) continue;                                       // the 16 cells miss the object in z (same for every row)
            // Word path for unrotated objects: the cells c0 <= c < c1 the object covers are sprite bytes zlo .. zlo + cnt - 1 of
            // the row, fetched as up to three aligned 8-byte words, funnel-shifted to byte 0, moved up to cell c0 and masked
            // to [c0, c1) -- whole chunks and the partial ones at the object's z faces alike (c0, c1 are the same for every
            // row of the thread).
            const int bz = wz0 - o.mn2, zlo = bz + c0, cnt = c1 - c0;
            const bool words = (o.a0 | o.a1 | o.a2) == 0 && (reinterpret_cast<uintptr_t>(o.sp) & 7u) == 0;
            const size_t sprite_bytes = (size_t)o.s0 * o.s1 * o.s2;
            const unsigned long long lo_c0 = c0 >= 8 ? ~0ull : (1ull << (8 * c0)) - 1ull, lo_c1 = c1 >= 8 ? ~0ull : (1ull << (8 * c1)) - 1ull;
            const unsigned long long hi_c0 = c0 <= 8 ? 0ull : (1ull << (8 * (c0 - 8))) - 1ull;
            const unsigned long long hi_c1 = c1 <= 8 ? 0ull : (c1 >= 16 ? ~0ull : (1ull << (8 * (c1 - 8))) - 1ull);
            const unsigned long long keep0 = lo_c1 & ~lo_c0, keep1 = hi_c1 & ~hi_c0;
            // pass 1: which rows take the word path, and where; pass 2: all their loads; pass 3: merge / cell-by-cell rest
            const unsigned long long *p[ROWS];
            int sh[ROWS], nw[ROWS];
            bool in_xy[ROWS], fast[ROWS];
#pragma unroll
            for (int r = 0; r < ROWS; r++) {
                const int bx = wx[r] - o.mn0, by = wy[r] - o.mn1;
                in_xy[r] = (need0[r] | need1[r]) != 0ull && wx[r] >= o.mn0 && wx[r] < o.mx0 && wy[r] >= o.mn1 && wy[r] < o.mx1;
                const size_t off = ((size_t)bx * o.s1 + by) * o.s2 + zlo;
                const size_t w0 = off >> 3, wl = (off + cnt - 1) >> 3;        // first and last aligned word touched
                fast[r] = in_xy[r] && words && (unsigned)bx < (unsigned)o.s0 && (unsigned)by < (unsigned)o.s1 && zlo >= 0 && zlo + cnt <= o.s2 &&
                          (wl + 1) * 8 <= sprite_bytes;
                p[r] = reinterpret_cast<const unsigned long long *>(o.sp) + w0;
                sh[r] = (int)(off & 7) * 8;
                nw[r] = (int)(wl - w0);
            }
            unsigned long long a[ROWS], b[ROWS], cw[ROWS];
#pragma unroll
            for (int r = 0; r < ROWS; r++) {
                a[r] = fast[r] ? __ldg(p[r]) : 0ull;
                b[r] = fast[r] && nw[r] >= 1 ? __ldg(p[r] + 1) : 0ull;
                cw[r] = fast[r] && nw[r] >= 2 ? __ldg(p[r] + 2) : 0ull;
            }
#pragma unroll
            for (int r = 0; r < ROWS; r++) {
                if (fast[r]) {
                    const unsigned long long u0 = sh[r] ? (a[r] >> sh[r]) | (b[r] << (64 - sh[r])) : a[r];      // sprite bytes zlo .. zlo + 7
                    const unsigned long long u1 = sh[r] ? (b[r] >> sh[r]) | (cw[r] << (64 - sh[r])) : b[r];     // zlo + 8 .. zlo + 15
                    unsigned long long v0, v1;                                                              // the same, moved up to cell c0
                    if (c0 == 0) { v0 = u0; v1 = u1; }
                    else if (c0 < 8) { v0 = u0 << (8 * c0); v1 = (u1 << (8 * c0)) | (u0 >> (64 - 8 * c0)); }
                    else if (c0 == 8) { v0 = 0ull; v1 = u0; }
                    else { v0 = 0ull; v1 = u0 << (8 * (c0 - 8)); }
                    const unsigned long long m0 = nonzero_bytes(v0) & need0[r] & keep0, m1 = nonzero_bytes(v1) & need1[r] & keep1;
                    out0[r] |= v0 & m0; need0[r] &= ~m0;
                    out1[r] |= v1 & m1; need1[r] &= ~m1;
                } else if (in_xy[r]) {
                    const ulonglong2 h0 = object_cells8(objs + k, wx[r], wy[r], wz0, out0[r], need0[r]);
                    const ulonglong2 h1 = object_cells8(objs + k, wx[r], wy[r], wz0 + 8, out1[r], need1[r]);
                    out0[r] = h0.x; need0[r] = h0.y; out1[r] = h1.x; need1[r] = h1.y;
                }
            }
        }
        x = (int)x0; y = (int)(row0 - x0 * (unsigned)ey);
#pragma unroll
        for (int r = 0; r < ROWS; r++) {
            if (valid[r]) {
                uint8_t *row_p = cells + ((size_t)(x + lx0) * ny + (y + ly0)) * nz;
                *reinterpret_cast<uint4 *>(row_p + z0) = make_uint4((uint32_t)out0[r], (uint32_t)(out0[r] >> 32), (uint32_t)out1[r], (uint32_t)(out1[r] >> 32));
            }
            if (++y == ey) { y = 0; x++; }
        }
    }
}

// ---- dense world -> bricks ------------------------------------------------------------------------------------
// A *task* is 32 bricks that follow each other along z: bz in [32 g, 32 g + 32) of brick column (bx, by).  Task
// index t = (bx * nby + by) * G + g, G = ceil(nbz / 32): tasks, and the bricks inside them, are in brick-index order,
// so the exclusive scan of the per-task counts numbers the non-empty bricks in brick-index order.

__device__ __forceinline__ void task_coords(long long t, int G, int nby, int &bx, int &by, int &g) {
    const long long col = t / G;
    g = (int)(t - col * G);
    bx = (int)(col / nby);
    by = (int)(col - (long long)bx * nby);
}

// One warp per task: lane j reads the 64 eight-byte rows of brick 32 g + j (a warp load is 256 contiguous bytes).
__global__ void __launch_bounds__(256) brick_mask_kernel(const uint8_t *__restrict__ dense, int nx, int ny, int nz, int nby, int nbz, int G,
                                                         long long n_tasks, bool vec, uint32_t *__restrict__ task_mask,
                                                         uint32_t *__restrict__ task_count) {
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long t = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < n_tasks; t += warps) {
        int bx, by, g;
        task_coords(t, G, nby, bx, by, g);
        const int bz = g * 32 + lane, z0 = bz * 8;
        uint32_t acc = 0;
        if (bz < nbz) {
            if (vec && bx * 8 + 8 <= nx && by * 8 + 8 <= ny) {            // interior column: eight independent loads in flight
                const uint2 *p = reinterpret_cast<const uint2 *>(dense + ((size_t)(bx * 8) * ny + by * 8) * nz + z0);
                const size_t sy = (size_t)nz >> 3, sx = sy * ny;
#pragma unroll 2
                for (int lx = 0; lx < 8; lx++) {
                    uint2 v[8];
#pragma unroll
                    for (int ly = 0; ly < 8; ly++) v[ly] = __ldg(p + lx * sx + ly * sy);
#pragma unroll
                    for (int ly = 0; ly < 8; ly++) acc |= v[ly].x | v[ly].y;
                }
            } else {
                for (int r = 0; r < 64; r++) {
                    const int x = bx * 8 + (r >> 3), y = by * 8 + (r & 7);
                    if (x < nx && y < ny) {
                        const uint2 v = load_run8(dense + ((size_t)x * ny + y) * nz, z0, nz, vec);
                        acc |= v.x | v.y;
                    }
                }
            }
        }
        const uint32_t mask = __ballot_sync(0xffffffffu, acc != 0u);
        if (lane == 0) { task_mask[t] = mask; task_count[t] = (uint32_t)__popc(mask); }
    }
}

// Exclusive scan of count[0 .. n) by one block of 1024 threads, eight consecutive entries per thread; *total = sum.
// (n = bricks / 32: 8 192 for a 512^3 world, i.e. one pass of the loop.)
__global__ void __launch_bounds__(1024) task_scan_kernel(const uint32_t *__restrict__ count, long long n, uint32_t *__restrict__ base,
                                                         unsigned long long *__restrict__ total) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t chunk_sum;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long carry = 0;
    for (long long i0 = 0; i0 < n; i0 += 8192) {
        const long long i = i0 + 8 * threadIdx.x;
        uint32_t v[8], mine = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) { v[k] = i + k < n ? count[i + k] : 0u; }
#pragma unroll
        for (int k = 0; k < 8; k++) mine += v[k];
        uint32_t s = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(full, s, d); if (lane >= d) s += u; }
        if (lane == 31) wsum[w] = s;
        __syncthreads();
        if (w == 0) {
            const uint32_t ws = wsum[lane];
            uint32_t q = ws;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(full, q, d); if (lane >= d) q += u; }
            wsum[lane] = q - ws;
            if (lane == 31) chunk_sum = q;
        }
        __syncthreads();
        uint32_t run = (uint32_t)carry + wsum[w] + s - mine;
#pragma unroll
        for (int k = 0; k < 8; k++) { if (i + k < n) base[i + k] = run; run += v[k]; }
        carry += chunk_sum;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

// One block of 256 threads per task.  The 64 rows x 256 bytes of the task's bricks are staged in shared memory
// (rows padded by 8 bytes against bank conflicts), then every non-empty brick is written to its pool slot as
// 32 lanes x 16 bytes = 512 contiguous bytes: pool index (lx * 8 + ly) * 8 + lz = row * 8 + lz, so 16 bytes are the
// runs of rows 2k and 2k + 1.  Warp 0 also writes the task's 32 brick-table entries (0 = empty, else slot + 1).
// The host launches this pass before it knows the brick count when a pool already exists (no host round trip in the
// middle of the pipeline): bricks whose slot lies beyond `pool_cap` are skipped and the pass is repeated after the
// pool has grown.
#define VT_PACK_ROW 264
__global__ void __launch_bounds__(256) brick_pack_kernel(const uint8_t *__restrict__ dense, int nx, int ny, int nz, int nby, int nbz, int G,
                                                         long long n_tasks, bool vec, const uint32_t *__restrict__ task_mask,
                                                         const uint32_t *__restrict__ task_base, uint32_t pool_cap,
                                                         uint8_t *__restrict__ pool, uint32_t *__restrict__ brick_tab) {
    __shared__ __align__(16) uint8_t tile[64 * VT_PACK_ROW];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (long long t = blockIdx.x; t < n_tasks; t += gridDim.x) {
        const uint32_t mask = task_mask[t], base = task_base[t];
        int bx, by, g;
        task_coords(t, G, nby, bx, by, g);
        if (w == 0 && g * 32 + lane < nbz) {
            const long long b = ((long long)bx * nby + by) * nbz + g * 32 + lane;
            brick_tab[b] = ((mask >> lane) & 1u) ? base + (uint32_t)__popc(mask & ((1u << lane) - 1u)) + 1u : 0u;
        }
        if (mask == 0u) continue;                                         // the same for every thread of the block
        const int z0 = (g * 32 + lane) * 8;
        const bool mine = (mask >> lane) & 1u;                            // implies 32 g + lane < nbz
#pragma unroll
        for (int p = 0; p < 8; p++) {
            const int r = w + 8 * p;
            const int x = bx * 8 + (r >> 3), y = by * 8 + (r & 7);
            uint2 v = make_uint2(0u, 0u);
            if (mine && x < nx && y < ny) v = load_run8(dense + ((size_t)x * ny + y) * nz, z0, nz, vec);
            *reinterpret_cast<uint2 *>(tile + r * VT_PACK_ROW + 8 * lane) = v;
        }
        __syncthreads();
        for (int j = w; j < 32; j += 8) {
            if (!((mask >> j) & 1u)) continue;                            // warp-uniform
            const uint32_t slot = base + (uint32_t)__popc(mask & ((1u << j) - 1u));
            if (slot >= pool_cap) continue;                               // optimistic launch into a pool that is too small: redone by the host
            const uint2 a = *reinterpret_cast<const uint2 *>(tile + (2 * lane) * VT_PACK_ROW + 8 * j);
            const uint2 b = *reinterpret_cast<const uint2 *>(tile + (2 * lane + 1) * VT_PACK_ROW + 8 * j);
            reinterpret_cast<uint4 *>(pool + (size_t)slot * 512u)[lane] = make_uint4(a.x, a.y, b.x, b.y);
        }
        __syncthreads();
    }
}

// Chunk slot (cx, cy, cz) holds a voxel iff one of its (chunk_size / 8)^3 bricks has a pool slot.
__global__ void chunk_present_kernel(const uint32_t *__restrict__ brick_tab, int nby, int nbz, int ncy, int ncz, long long n_chunks, int bpc,
                                     uint8_t *__restrict__ present) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_chunks; i += (long long)gridDim.x * blockDim.x) {
        const int cz = (int)(i % ncz), cy = (int)((i / ncz) % ncy), cx = (int)(i / ((long long)ncz * ncy));
        uint32_t any = 0;
        for (int a = 0; a < bpc; a++)
            for (int b = 0; b < bpc; b++)
                for (int c = 0; c < bpc; c++)
                    any |= __ldg(&brick_tab[((size_t)(cx * bpc + a) * nby + (cy * bpc + b)) * nbz + (cz * bpc + c)]);
        present[i] = any ? 1 : 0;
    }
}

// Inverse of the pack pass: bricks -> dense [nx][ny][nz] grid (read-back for checks and for callers that switch from
// the chunk_set path to the dense path); one thread per run of eight cells = one 8-byte row of a brick.
__global__ void __launch_bounds__(256) unpack_cells_kernel(const uint32_t *__restrict__ brick_tab, const uint8_t *__restrict__ pool, int nx, int ny,
                                                           int nz, int nby, int nbz, bool vec, uint8_t *__restrict__ dense) {
    const int runs = (nz + 7) >> 3;
    const long long total = (long long)nx * ny * runs;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / runs;
        const int r = (int)(i - row * runs), x = (int)(row / ny), y = (int)(row - (long long)x * ny);
        const uint32_t b = __ldg(&brick_tab[((size_t)(x >> 3) * nby + (y >> 3)) * nbz + r]);
        uint2 v = make_uint2(0u, 0u);
        if (b) v = __ldg(reinterpret_cast<const uint2 *>(pool + (size_t)(b - 1u) * 512u + (((x & 7) << 6) | ((y & 7) << 3))));
        store_run8(dense + ((size_t)x * ny + y) * nz, 8 * r, nz, vec, v);
    }
}

// ---- small in-place updates -----------------------------------------------------------------------------------

// scatter m staged bricks into their pool slots and patch the brick table
__global__ void scatter_bricks_kernel(const uint8_t *__restrict__ staged, const uint32_t *__restrict__ slots,
                                      const uint32_t *__restrict__ tab_index, const uint32_t *__restrict__ tab_value,
                                      uint8_t *pool, uint32_t *brick_tab, int m) {
    int b = blockIdx.x;
    if (b >= m) return;
    uint32_t slot = slots[b];
    if (slot != 0xffffffffu) {
        const uint4 *src = reinterpret_cast<const uint4 *>(staged + (size_t)b * 512);
        uint4 *dst = reinterpret_cast<uint4 *>(pool + (size_t)slot * 512);
        dst[threadIdx.x] = src[threadIdx.x];          // 32 threads x 16 B
    }
    if (threadIdx.x == 0) brick_tab[tab_index[b]] = tab_value[b];
}

// one warp per brick of the box of bricks [b0, b0 + nbr): flags[i] = brick holds a voxel
__global__ void region_flags_kernel(const uint8_t *__restrict__ dense, int nx, int ny, int nz, int bx0, int by0, int bz0,
                                    int rbx, int rby, int rbz, uint8_t *flags) {
    const int warps = (gridDim.x * blockDim.x) >> 5, lane = threadIdx.x & 31;
    const int total = rbx * rby * rbz;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < total; i += warps) {
        const int bz = bz0 + i % rbz, by = by0 + (i / rbz) % rby, bx = bx0 + i / (rbz * rby);
        unsigned any = 0;
        for (int t = lane; t < 512; t += 32) {
            const int x = bx * 8 + (t >> 6), y = by * 8 + ((t >> 3) & 7), z = bz * 8 + (t & 7);
            if (x < nx && y < ny && z < nz) any |= dense[((size_t)x * ny + y) * nz + z];
        }
        any = __any_sync(0xffffffffu, any != 0);
        if (lane == 0) flags[i] = (uint8_t)any;
    }
}

// copy the region's non-empty bricks from the dense grid into their pool slots and patch the brick table
// (one block of 512 threads per brick; slot 0xffffffff = brick became empty)
__global__ void region_fill_kernel(const uint8_t *__restrict__ dense, int nx, int ny, int nz, int nby, int nbz, int bx0, int by0,
                                   int bz0, int rby, int rbz, const uint32_t *__restrict__ slots, int m, uint8_t *pool,
                                   uint32_t *brick_tab) {
    for (int i = blockIdx.x; i < m; i += gridDim.x) {
        const int bz = bz0 + i % rbz, by = by0 + (i / rbz) % rby, bx = bx0 + i / (rbz * rby);
        const uint32_t slot = slots[i];
        const int t = threadIdx.x;
        if (slot != 0xffffffffu) {
            const int x = bx * 8 + (t >> 6), y = by * 8 + ((t >> 3) & 7), z = bz * 8 + (t & 7);
            pool[(size_t)slot * 512 + t] = (x < nx && y < ny && z < nz) ? dense[((size_t)x * ny + y) * nz + z] : 0;
        }
        if (t == 0) brick_tab[((size_t)bx * nby + by) * nbz + bz] = slot == 0xffffffffu ? 0u : slot + 1u;
    }
}

}  // namespace vt
