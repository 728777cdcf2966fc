// vt_physics.cuh — the voxel slice test of Object.update_physics (reference data.py:520-542) for sm_100a.
//
// The reference moves a physical object one cell at a time; before each move it walks the one-cell-thick slice of cells
// the object would enter and, for every other object whose box touches the slice, asks cell by cell (x, then y, then z):
// does the other object have a solid voxel here, and does the moving object have a solid voxel at the cell that would
// land here (pos - self.mins - vel_dir)?  Every such pair blocks the move and adds to the friction / elasticity sums.
// Here the cells of the slice are tested in parallel (two sprite lookups each, through Sprite.pos_rotated like the world
// builder), and the sums are then taken by one warp in the reference's cell order, so the fp64 results are the
// interpreter's bit for bit.  `solidity > random.random()` (data.py:537, 539) is exact for solidity 0 and 1 -- every
// shipped material -- and otherwise drawn from a Philox stream keyed by the cell (statistical parity, like the renderer).
#pragma once
#include "vt_device.cuh"
#include "vt_scene.cuh"

namespace vtphys {

struct SliceArgs {
    vt_phys_object self, other;
    int lo[3], n[3];                 // slice corner and cells per axis (post6 is inclusive at both ends, data.py:533-535)
    int dir[3];                      // vel_dir, one of the six axis directions
    double scale;                    // settings.friction (data.py:540-541)
    uint32_t seed_lo, seed_hi;
};

// Sprite.get_voxel(None, pos - obj.mins, obj.rot) (data.py:417-419): material id or 0.  The voxel dictionary of a sprite
// may hold keys outside [0, size) -- Sprite.load maps file column x = 0 to x = size.x (data.py:285), which the renderer
// never draws but this test does ask for -- so the grid covers the keys' bounding box [grid_lo, grid_lo + grid_dims).
__device__ __forceinline__ int sprite_voxel(const vt_phys_object &o, int x, int y, int z) {
    x -= o.mins[0]; y -= o.mins[1]; z -= o.mins[2];
    vt_object r;                                           // pos_rotated reads size and angle only
    r.size[0] = o.size[0]; r.size[1] = o.size[1]; r.size[2] = o.size[2];
    r.angle[0] = o.angle[0]; r.angle[1] = o.angle[1]; r.angle[2] = o.angle[2];
    vt::pos_rotated(r, x, y, z);
    x -= o.grid_lo[0]; y -= o.grid_lo[1]; z -= o.grid_lo[2];
    if ((unsigned)x >= (unsigned)o.grid_dims[0] || (unsigned)y >= (unsigned)o.grid_dims[1] || (unsigned)z >= (unsigned)o.grid_dims[2]) return 0;
    return __ldg(o.dev_grid + ((size_t)x * o.grid_dims[1] + y) * o.grid_dims[2] + z);
}

__device__ __forceinline__ bool solid(double solidity, uint32_t cell, uint32_t which, uint32_t seed_lo, uint32_t seed_hi) {
    if (solidity >= 1.0) return true;                                   // random.random() < 1
    if (!(solidity > 0.0)) return false;
    uint32_t w[4];
    vt::philox4x32_10(which, 0u, seed_lo, seed_hi, cell, 0x50485953u, w);
    return solidity > vt::u53m(w[0], w[1]);
}

// props: [n_mat + 1][4] solidity, friction, elasticity, 0.   flag / contrib: one entry per slice cell, x-major, z fastest.
__global__ void slice_cells_kernel(const SliceArgs a, const double *__restrict__ props, long long n_cells,
                                   uint8_t *__restrict__ flag, double2 *__restrict__ contrib) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_cells; i += (long long)gridDim.x * blockDim.x) {
        const int z = a.lo[2] + (int)(i % a.n[2]);
        const long long t = i / a.n[2];
        const int y = a.lo[1] + (int)(t % a.n[1]), x = a.lo[0] + (int)(t / a.n[1]);
        uint8_t f = 0;
        double2 c = make_double2(0.0, 0.0);
        const int om = sprite_voxel(a.other, x, y, z);
        if (om && solid(__ldg(props + 4 * om), (uint32_t)i, 0u, a.seed_lo, a.seed_hi)) {
            const int sm = sprite_voxel(a.self, x - a.dir[0], y - a.dir[1], z - a.dir[2]);
            if (sm && solid(__ldg(props + 4 * sm), (uint32_t)i, 1u, a.seed_lo, a.seed_hi)) {
                f = 1;
                c.x = __ldg(props + 4 * om + 1) * __ldg(props + 4 * sm + 1) * a.scale;      // (a * b) * c like the interpreter
                c.y = __ldg(props + 4 * om + 2) * __ldg(props + 4 * sm + 2) * a.scale;
            }
        }
        flag[i] = f;
        contrib[i] = c;
    }
}

// The contributions are added in cell order (fp64 sums are order-sensitive and the reference's order is x, y, z): one
// block first compacts the flagged cells in place, in order (ballot + prefix over the 32 warps, 1024 cells per round),
// then thread 0 adds the compacted list sequentially; out = {friction, elasticity, blocked} continues from the sums passed
// in.  Contact surfaces are small next to the slice, so the sequential part is short.
__global__ void __launch_bounds__(1024) slice_sum_kernel(const uint8_t *__restrict__ flag, double2 *__restrict__ contrib, long long n_cells, double *__restrict__ out) {
    __shared__ int warp_cnt[32];
    __shared__ long long total;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    for (long long base = 0; base < n_cells; base += 1024) {
        const long long i = base + threadIdx.x;
        const bool f = i < n_cells && flag[i];
        const double2 c = f ? contrib[i] : make_double2(0.0, 0.0);
        const unsigned m = __ballot_sync(full, f);
        if (lane == 0) warp_cnt[warp] = __popc(m);
        __syncthreads();                                    // every entry of this round is in registers: in-place writes are safe
        int before = 0, all = 0;
        for (int w = 0; w < 32; w++) { const int k = warp_cnt[w]; all += k; if (w < warp) before += k; }
        if (f) contrib[total + before + __popc(m & ((1u << lane) - 1u))] = c;
        __syncthreads();
        if (threadIdx.x == 0) total += all;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        double fr = out[0], el = out[1];
        for (long long k = 0; k < total; k++) { fr += contrib[k].x; el += contrib[k].y; }
        out[0] = fr; out[1] = el; out[2] = total > 0 ? 1.0 : 0.0;
    }
}

}  // namespace vtphys
