/* voxtrace.h — C ABI of the B200-native voxel ray-march path (libvoxtrace.so).
 *
 * Drop-in boundary for python_raytracer's CPU trace loop.  The reference has no FFI; its interface for this
 * path is the Python object protocol of `Camera` (reference init.py:13-150).  Each entry point below names the
 * reference interface it stands in for (paths relative to the reference root).  The Python mirror of that
 * protocol (`python_raytracer_b200.GpuCamera`) is a thin ctypes layer over these calls; INTEGRATION.md shows
 * the binding a maintainer of the reference would add.
 *
 * Conventions: every call returns 0 on success, non-zero on error; vt_last_error() then returns a host
 * string owned by the context.  A context is bound to ONE CUDA device and ONE stream and is not
 * thread-safe (the reference's Camera is likewise used by one caller at a time, init.py:205).
 * Pointers named dev_* are device pointers on the context's device; host_* are host pointers.
 * There is no CPU fallback: vt_create fails if no CUDA device is usable.
 */
#ifndef VOXTRACE_H
#define VOXTRACE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VT_ABI_VERSION 1
#define VT_BRICK 8                 /* bricks are 8x8x8 cells = 512 bytes */
#define VT_MAX_MATERIALS 255       /* material id 0 is "empty" */
#define VT_RAY_RECORD 16           /* doubles per ray record (see vt_render) */
#define VT_TRACE_RECORD 20         /* doubles per record of vt_trace_rays */

typedef struct vt_ctx vt_ctx;

/* Subset of `data.settings` the path reads (reference data.py:18-50) plus the derived fields of
 * data.py:64-68.  Same names and meaning; `static_noise` is `settings.static`. */
typedef struct {
    int32_t width, height;         /* WINDOW width/height: the canvas the lens is defined on (init.py:131-132) */
    int32_t samples;               /* RENDER samples (init.py:134) */
    int32_t static_noise;          /* RENDER static: per-pixel streams do not change between frames (init.py:136) */
    int32_t chunk_size;            /* RENDER chunk_size (init.py:68) */
    int32_t chunk_radius;          /* round(chunk_size / 2) (data.py:68) */
    int32_t background;            /* 0: data.background is None; 1: lib.material_background (init.py:119-120) */
    int32_t reserved;
    double shutter, falloff, dof, dist_min, dist_max, max_light, max_bounces;
    double lod_bounces, lod_samples, lod_random, lod_edge;
    double proportions;            /* ((w + h) / 2) / max(w, h) (data.py:66) */
} vt_settings;

/* One row of the structure-of-arrays material table: the optical fields `lib.material` reads
 * (reference lib.py:448-460; mods/default/init.py:6-17).  Row 0 is unused (id 0 = empty cell). */
typedef struct {
    double albedo[3];              /* rgb 0..255 */
    double roughness, absorption, ior, energy;
    double reserved;
} vt_material;

/* ---- life cycle -------------------------------------------------------------------------------- */
/* Camera.__init__ (init.py:14-18).  `device` is a CUDA ordinal. */
int vt_create(int device, vt_ctx **out);
void vt_destroy(vt_ctx *ctx);
const char *vt_last_error(const vt_ctx *ctx);   /* ctx may be NULL: error of the last failed vt_create */
int vt_abi_version(void);
/* Build variant of the loaded library: bit 0 = bounds-checked kernels (-DVT_CHECKED=1, tests/check_bounds.sh),
 * bit 1 = phase-trip instrumentation (-DVT_TRIPS=1, tools/kbench.py).  0 for the product build. */
int vt_build_flags(void);
/* Scheduling knobs of the render kernel for A/B measurements (tools/kbench.py): "runs" 0/1 brick-local loops, "powtab"
 * 0/1 memoised falloff power, "refill" 1..32 idle lanes that trigger a refill, "group" 1..32 lanes per pixel of a packet, "buildrows" 0 / 2 / 4 rows per thread of the wide world-build kernel (0: off),
 * "debug" 0/1.  None changes a result and the product never calls it (nothing is read from the environment). */
int vt_set_tuning(vt_ctx *ctx, const char *key, int value);
/* Use an existing CUDA stream (cudaStream_t, e.g. torch's current stream) instead of the context's own. */
int vt_set_stream(vt_ctx *ctx, void *cuda_stream);
int vt_sync(vt_ctx *ctx);

/* ---- settings, materials, camera ----------------------------------------------------------------- */
/* `data.settings` as read inside Camera.trace / Camera.tile (init.py:41-42, 53, 56, 85, 114, 131-141). */
int vt_set_settings(vt_ctx *ctx, const vt_settings *host_settings);
/* Material attribute bag (data.py:85-93) for ids 1..n; host_materials has n+1 rows. */
int vt_set_materials(vt_ctx *ctx, const vt_material *host_materials, int n);
/* cam.pos / cam.rot / cam.lens attribute writes (init.py:464-465, init.py:282). */
int vt_set_camera(vt_ctx *ctx, const double pos[3], const double rot_xyzw[4], double lens);

/* ---- world: what Camera.chunk_set / chunk_get expose (init.py:21-33) ------------------------------- */
/* Declares the cell box covered by the camera's chunk dictionary: origin (world cell of the minimum corner)
 * and dims (cells), both multiples of chunk_size.  Drops any previous world (its brick-pool memory is kept for the next
 * one and released by vt_destroy; the tables are reused when the box has the same shape).  Needs vt_set_settings first. */
int vt_set_world(vt_ctx *ctx, const int32_t origin[3], const int32_t dims[3]);
/* Drops the world and releases its memory; afterwards vt_set_settings accepts another chunk_size (an engine that edits
 * data.settings.chunk_size in place, data.py:41). */
int vt_clear_world(vt_ctx *ctx);
/* Per-chunk-slot byte, [dims0/cs][dims1/cs][dims2/cs] (z fastest): 0 = no frame for that chunk (absent or
 * culled, init.py:447-452), else Frame.resolution of the frame the camera holds (init.py:448-450). */
int vt_set_chunk_table(vt_ctx *ctx, const uint8_t *host_res, int64_t n_slots);
/* Upload m dirty bricks.  brick_coords[i] = brick index per axis relative to the world origin
 * ((cell - origin) / 8); cells[i] = 512 material ids, index (lx*8 + ly)*8 + lz.  An all-zero brick releases
 * its storage.  Stands in for chunk_set replacing one chunk's Frame (init.py:21-25, 450-452). */
int vt_upload_bricks(vt_ctx *ctx, const int32_t *host_brick_coords, const uint8_t *host_cells, int m);
/* Bulk path for synthetic / imported worlds: a dense [dims0][dims1][dims2] (z fastest) u8 array in DEVICE
 * memory is bricked on the device; empty bricks get no storage.  Three streaming launches (occupancy bits, slot
 * scan, pack); pool slots follow brick-index order; the only host step is reading the brick count (8 bytes). */
int vt_upload_dense(vt_ctx *ctx, const uint8_t *dev_cells);
/* The inverse: the world held in bricks written out as a dense [dims0][dims1][dims2] u8 array in DEVICE memory. */
int vt_download_dense(vt_ctx *ctx, uint8_t *dev_cells);
/* Which chunk slots hold a voxel -- what `post_chunk in self.chunks` answers once chunk_update has run
 * (init.py:441-452), before LOD and culling: [dims0/cs][dims1/cs][dims2/cs] bytes (z fastest, 0 / 1) into HOST
 * memory, derived from the brick table on the device.  chunk_size must be a multiple of 8. */
int vt_chunk_presence(vt_ctx *ctx, uint8_t *host_present, int64_t n_slots);
/* Bytes of device memory held by bricks / number of non-empty bricks (reporting). */
int vt_world_stats(vt_ctx *ctx, int64_t *n_bricks, int64_t *bytes);

/* ---- scene building: what Window.chunk_update does before the path (init.py:389-452) ------------------------- */
/* One world object as chunk_update sees it: the active frame of its sprite as a dense grid of material ids
 * (Sprite.get_voxel -> Frame.get_voxel, data.py:417-419, 136-145; sprites of resolution 1), its integer box
 * (Object.mins / maxs, data.py:480-484, 592-594) and its 90-degree rotation steps round(rot / 90) % 4
 * (Sprite.pos_rotated, data.py:338-371). */
typedef struct {
    const uint8_t *dev_sprite;     /* [size0][size1][size2] u8, z fastest; 0 = no voxel */
    int32_t size[3];               /* Sprite.size */
    int32_t mins[3], maxs[3];      /* cells mins <= p < maxs are drawn (init.py:413-419) */
    int32_t angle[3];              /* 0..3 per axis */
    int32_t reserved;
} vt_object;
/* Draws n objects into a dense [dims0][dims1][dims2] u8 grid in DEVICE memory covering the box of vt_set_world:
 * cell = material of the LAST object in the list that has a voxel there (chunk frames are merged in object order,
 * later ones override, init.py:437-439).  Follow with vt_upload_dense(ctx, dev_cells).  One kernel launch. */
int vt_build_world(vt_ctx *ctx, const vt_object *host_objects, int n, uint8_t *dev_cells);
/* Objects whose `redraw` flag is set (init.py:397-404: their old and new chunks are rebuilt): redraws the cells of the
 * box lo <= p < hi (absolute cells; clipped to the world, widened to whole bricks) of the same dense grid from the
 * FULL object list in its current order, then updates the bricks of that box in place (new bricks take free slots,
 * bricks that became empty release theirs).  The caller passes the union of the old and new chunk boxes of the
 * objects that moved.  Three small launches and one host round trip, independent of the world size. */
int vt_update_region(vt_ctx *ctx, const vt_object *host_objects, int n, uint8_t *dev_cells, const int32_t lo[3], const int32_t hi[3]);

/* ---- physics: the voxel slice test of Object.update_physics (data.py:520-542) -------------------------------------
 * Physical fields of a Material (mods/default/init.py:6-17): rows for ids 0..n, row 0 unused. */
typedef struct { double solidity, friction, elasticity, reserved; } vt_phys_material;
int vt_set_phys_materials(vt_ctx *ctx, const vt_phys_material *host_materials, int n);
/* An object as the physics step sees it: the voxel dictionary of its sprite's active frame as a dense grid of material ids
 * over the bounding box of the dictionary's keys (keys may lie outside [0, size): Sprite.load maps file column x = 0 to
 * x = size.x, data.py:285), Sprite.size, Object.mins (data.py:480-484) and the 90-degree rotation steps. */
typedef struct {
    const uint8_t *dev_grid;       /* [grid_dims0][grid_dims1][grid_dims2] u8, z fastest; 0 = no voxel */
    int32_t grid_lo[3];            /* sprite-local coordinate of grid[0][0][0] */
    int32_t grid_dims[3];
    int32_t size[3];               /* Sprite.size */
    int32_t mins[3];               /* Object.mins */
    int32_t angle[3];              /* round(rot / 90) % 4 per axis */
    int32_t reserved;
} vt_phys_object;
/* One moving object against one other object whose box touches the slice the mover would enter.  lo / hi: the
 * inclusive corners post6 of that one-cell-thick slice (data.py:506-517); vel_dir: the unit axis step.  For every cell of
 * the slice in the reference's order (x, then y, then z): other's voxel at the cell, the mover's voxel at cell - vel_dir
 * (both through Sprite.pos_rotated); where both exist and pass `solidity > random.random()` the move is blocked and
 * friction += of * sf * friction_scale, elasticity += oe * se * friction_scale (data.py:540-541) are added to
 * inout_sums[0..1] in that order -- the sums a caller carries from object to object like the reference's loop.
 * Solidity 0 and 1 are exact; fractional solidity draws from a Philox stream keyed by (seed, cell).  Synchronises. */
int vt_slice_collide(vt_ctx *ctx, const vt_phys_object *self, const vt_phys_object *other, const int32_t lo[3], const int32_t hi[3],
                     const int32_t vel_dir[3], double friction_scale, uint64_t seed, double inout_sums[2], int32_t *out_blocked);

/* Allocation-free variant of vt_slice_collide.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same test, same kernels, same results; what differs:
 *   - the per-cell scratch lives in a file-scope `static __device__` array of the library: at most
 *     VT_STATIC_SLICE_MAX_CELLS slice cells per call (larger slices are rejected with an error -- use vt_slice_collide),
 *     1 114 112 bytes that every GPU which loads the library holds until the process exits;
 *   - inputs and outputs are the caller's: dev_materials = DEVICE pointer to vt_phys_material rows 0..n (what
 *     vt_set_phys_materials would have uploaded), dev_inout = DEVICE pointer to 3 doubles {friction sum, elasticity sum,
 *     blocked}: the two sums are read, continued and written back, blocked is written as 1.0 / 0.0;
 *   - the call only enqueues two kernels on the context's stream (it does not synchronise and copies nothing): the
 *     result is in dev_inout once the stream has reached that point.  No cudaMalloc / cudaFree / allocating library call
 *     is made; the first call on a device looks the scratch address up with cudaGetSymbolAddress (do that call before
 *     starting a graph capture).
 * Every call on a GPU shares the one array, which keeps the last call's contents (nothing in it needs clearing: every
 * cell a call reads it has written first).  Calls that overlap on the GPU -- two streams, two contexts on one device --
 * race on it: serialise them. */
#define VT_STATIC_SLICE_MAX_CELLS 65536
int vt_slice_collide_static(vt_ctx *ctx, const vt_phys_object *self, const vt_phys_object *other, const int32_t lo[3],
                            const int32_t hi[3], const int32_t vel_dir[3], double friction_scale, uint64_t seed,
                            const vt_phys_material *dev_materials, double *dev_inout);

/* Goxel text model -> voxel list, the parser rules of Sprite.load (data.py:271-289): single-space separated
 * "x y z rrggbb" lines, non-numeric lines skipped, colours looked up in `keys` (n_keys strings; unknown ones are
 * dropped), position (size_x - x, z, y).  Returns malloc'ed arrays in FILE ORDER (later entries override earlier
 * ones): out_xyz [n][3] int32, out_mat [n] u8 = index into keys + 1.  Release both with vt_free.  `.txt` is
 * mmap'ed, `.gz` goes through zlib.  Pure host code (no context, no GPU). */
int vt_goxel_parse(const char *path, int32_t size_x, const char *const *keys, int n_keys, int32_t **out_xyz,
                   uint8_t **out_mat, int64_t *out_n, char *err, int err_len);
void vt_free(void *p);

/* ---- render: Camera.tile (init.py:126-150) --------------------------------------------------------- */
/* Pixel set of one call.  The reference hands each worker `settings.pixels[thread]` (data.py:70-77); here a
 * call covers either a rectangle of interleaved row stripes or an explicit pixel list:
 *   rect mode  (dev_pixels == NULL): pixels x0 <= x < x1, y0 <= y < y1 whose stripe (y / stripe_rows) %
 *              stripe_count == stripe_index.  Output index i = row_local * (x1 - x0) + (x - x0), where
 *              row_local counts the selected rows in increasing y (a compact band image).
 *   list mode  (dev_pixels != NULL): n_pixels (x, y) int32 pairs; output index i = list position.       */
typedef struct {
    int32_t x0, y0, x1, y1;
    int32_t stripe_rows, stripe_count, stripe_index;   /* 1, 1, 0 = every row */
    int32_t reserved;
    const int32_t *dev_pixels;
    int64_t n_pixels;
} vt_region;

/* Outputs (device pointers; any may be NULL):
 *   dev_rgba8    [n][4] u8: per-pixel mean of r, g, b and alpha = round(min(1, energy + shutter) * 255)
 *                (init.py:141-146), truncated to u8 like a C cast
 *   dev_mean     [n][4] f32: the same means before quantisation
 *   dev_ray0     [n][VT_RAY_RECORD] f64, sample 0 of each pixel (parity probe): r, g, b, energy, bounces, step,
 *                hits, first-hit x, y, z, first-hit flag, first-hit step, vel x, y, z, samples traced
 *   dev_visited  u32 bitmask over chunk slots, OR-ed in (caller clears): ray.traversed (init.py:72-73, 143)
 *   dev_counters [8] u64, added to (caller clears): rays, in-chunk samples, void samples, hits,
 *                neighbour probes, chunk re-evaluations, loop iterations, 0
 * frame_seed perturbs the Philox streams when settings.static is 0; ignored when it is 1. */
int vt_render(vt_ctx *ctx, const vt_region *region, uint8_t *dev_rgba8, float *dev_mean, double *dev_ray0,
              uint32_t *dev_visited, uint64_t *dev_counters, uint64_t frame_seed);
/* Camera.trace(dir_x, dir_y, detail) (init.py:37-121) for n caller-given rays: dev_dirs [n][3] f64 = dir_x, dir_y, detail
 * (detail scales the ray's life as given, init.py:56; no lod_random draw -- that happens in tile, init.py:139).  Ray i
 * draws its dof / roughness jitter from the Philox stream keyed i (n <= 65535 * width).  dev_sincos (nullable, used
 * when dof == 0): [n][4] f64 sin, cos of the half yaw angle -lens_x / 2 and sin, cos of the half pitch angle lens_y / 2
 * in radians (init.py:41-43, lib.py:322-332) computed by the caller with the libm the reference's interpreter uses;
 * without it the device evaluates them (<= 1 ulp).  dev_records [n][VT_TRACE_RECORD] f64 = the returned ray: r, g, b,
 * energy, bounces, step, hits, first-hit x, y, z, first-hit flag, first-hit step, vel x, y, z, life, pos x, y, z, 0.
 * dev_visited / dev_counters as in vt_render (ray.traversed of all rays, OR-ed). */
int vt_trace_rays(vt_ctx *ctx, const double *dev_dirs, const double *dev_sincos, int64_t n, double *dev_records,
                  uint32_t *dev_visited, uint64_t *dev_counters, uint64_t frame_seed);

/* ---- multi-GPU frame assembly over peer memory (replaces the pool's result pipes, init.py:185-190, 205) -------
 * One process per GPU.  Rank 0 owns the full frame(s) and an array of arrival flags in memory from vt_shared_alloc;
 * the other ranks open them with CUDA IPC and render with a FRAME TARGET: every finished pixel is stored at
 * frame[(y * width + x) * 4] -- a store over NVLink when the frame is a peer's -- and when a rank's launch is
 * complete its last warp publishes `arrive_value` in its arrival flag (system-scope fence first; a rank whose region
 * is empty publishes it too).  Rank 0 queues
 * vt_stream_wait_flags on its stream instead of a collective.  Flag values only grow (frame numbers).
 * vt_stream_signal_flags is the reverse direction (rank 0 hands a frame buffer back to the renderers). */
int vt_shared_alloc(vt_ctx *ctx, int64_t bytes, void **out_dev_ptr);     /* zero-filled, plain cudaMalloc */
int vt_shared_free(vt_ctx *ctx, void *dev_ptr);
int vt_ipc_export(vt_ctx *ctx, void *dev_ptr, uint8_t handle[64]);
int vt_ipc_open(vt_ctx *ctx, const uint8_t handle[64], void **out_dev_ptr);
int vt_ipc_close(vt_ctx *ctx, void *dev_ptr);
/* dev_frame NULL switches the target off.  Rectangle regions only; dev_rgba8 of vt_render may then be NULL. */
int vt_set_frame_target(vt_ctx *ctx, uint8_t *dev_frame, uint32_t *dev_arrive_flag, uint32_t arrive_value);
/* Multi-GPU culling -- the union of the workers' traversed lists (init.py:189, 393).  With a frame target set, the last
 * warp of every vt_render launch that is given dev_visited ORs the launch's visited words into dev_peer_mask (n_words
 * u32 in a peer's memory, e.g. rank 0's, opened with vt_ipc_open) before it raises the arrival flag, so whoever has
 * seen all arrival flags holds the union.  NULL switches it off. */
int vt_set_visited_peer(vt_ctx *ctx, uint32_t *dev_peer_mask, int64_t n_words);
/* Stream-ordered: continue once flags[0..n) have all reached `value` (n <= 32).  A wait longer than timeout_s
 * (default 2 s) gives up and records 1 + flag index for vt_wait_status instead of hanging the GPU. */
int vt_stream_wait_flags(vt_ctx *ctx, const uint32_t *dev_flags, int n, uint32_t value, double timeout_s);
/* Stream-ordered: store `value` into n flags (n <= 16) given by a HOST array of device pointers (peers' memory). */
int vt_stream_signal_flags(vt_ctx *ctx, uint32_t *const *host_flag_ptrs, int n, uint32_t value);
/* Same, on another CUDA stream of the context's device (e.g. the copy stream that reads the frame out). */
int vt_stream_signal_flags_on(vt_ctx *ctx, void *cuda_stream, uint32_t *const *host_flag_ptrs, int n, uint32_t value);
int vt_wait_status(vt_ctx *ctx, int *status);                            /* synchronises; 0 = no wait timed out; resets */
/* The same word without synchronising (it lives in pinned host memory the wait kernel writes): for callers that have
 * already waited on an event recorded after the wait.  Non-zero is reported once, then reset. */
int vt_wait_status_peek(vt_ctx *ctx, int *status);
int vt_copy_to_host(vt_ctx *ctx, void *host_dst, const void *dev_src, int64_t bytes);   /* async on the ctx stream */

/* ---- image-space steps after the path: Window.draw_tile / Window.draw (init.py:185-190, 207-253) ----------------
 * pygame's own rounding is unpinned; the formulas are stated in csrc/vt_postfx.cuh. */
typedef struct {
    int32_t width, height;         /* canvas = settings.window */
    int32_t out_width, out_height; /* settings.window_scaled (data.py:65) */
    int32_t iris_on;               /* settings.iris and settings.iris_time (init.py:219) */
    int32_t reserved;
    double spill;                  /* settings.spill, 0 = off (init.py:213) */
    double iris;                   /* Window.iris, the adaptation state (init.py:220-222, 462) */
    double bloom, bloom_blur;      /* settings.bloom, settings.bloom_blur, 0 = off (init.py:232) */
    double subsamples;             /* settings.subsamples (init.py:242-244) */
    double smooth;                 /* settings.smooth: 0 nearest, 1 smooth, else sharp-then-smooth (init.py:247-254) */
} vt_fx;
/* draw_tile: alpha-blit one tile (RGBA8, alpha 0 where the tile has no pixel) onto the canvas, n_pixels each. */
int vt_fx_compose(vt_ctx *ctx, uint8_t *dev_canvas, const uint8_t *dev_tile, int64_t n_pixels);
/* draw: spill, iris, bloom, subsampling and scaling of a COPY of the canvas into dev_out [out_height][out_width][4].
 * host_avg receives transform.average_color(canvas, consider_alpha=True), which the caller needs for
 * iris_target = 1 - max(r, g, b) / 255 * 2 (init.py:229).  Synchronises the stream. */
int vt_fx_draw(vt_ctx *ctx, const vt_fx *fx, const uint8_t *dev_canvas, uint8_t *dev_out, int32_t host_avg[4]);

/* Self-test hooks.  (1 + bounces) ** (1 + falloff) of lib.material / material_background (lib.py:450, 465) is evaluated by a
 * correctly rounded double-double pow (csrc/vt_pow_cr.h), the same code on the host and on the device: vt_pow_host runs it on
 * the host (no GPU needed), vt_selftest_pow on the device for n argument pairs. */
double vt_pow_host(double x, double y);
int vt_selftest_pow(vt_ctx *ctx, const double *dev_x, const double *dev_y, int64_t n, double *dev_out);

/* Number of output pixels a region produces (host arithmetic only). */
int64_t vt_region_pixels(const vt_region *region);
/* Number of kernel launches issued by this context since creation (bench.py's gpu_launches). */
int64_t vt_launch_count(const vt_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif
