// voxtrace.cu — C ABI (include/voxtrace.h) + kernels of the B200-native voxel ray-march path.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -shared -Xcompiler -fPIC
// (see python_raytracer_b200/build.py).  No torch types cross this boundary.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "vt_device.cuh"
#include "vt_march.cuh"
#include "vt_scene.cuh"

using namespace vt;

// ------------------------------------------------------------------------------------------------------
// kernels


// Peer-memory handshakes of the multi-GPU frame path.  Flags are u32 words in cudaMalloc'ed memory
// that peers reach through CUDA IPC; values only grow (frame sequence numbers), so a waiter never misses one.
__global__ void wait_flags_kernel(const unsigned int *flags, int n, unsigned int value, long long timeout_cycles, int *status) {
    const int t = threadIdx.x;
    if (t >= n) return;
    const long long t0 = clock64();
    while ((int)(*(volatile const unsigned int *)(flags + t) - value) < 0) {
        if (clock64() - t0 > timeout_cycles) { *(volatile int *)status = 1 + t; __threadfence_system(); return; }
        __nanosleep(200);
    }
    __threadfence_system();
}

struct FlagTargets { unsigned int *p[16]; };
__global__ void signal_flags_kernel(FlagTargets tg, int n, unsigned int value) {
    const int t = threadIdx.x;
    if (t >= n || !tg.p[t]) return;
    __threadfence_system();
    atomicExch_system(tg.p[t], value);
}

// ------------------------------------------------------------------------------------------------------
// host side: context + C ABI

#define VT_STAGES 8
struct vt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    long long launches = 0;

    bool have_settings = false;
    vt_settings s{};
    int cs_shift = -1;

    double cam_pos[3] = {0, 0, 0}, cam_rot[4] = {0, 0, 0, 0}, cam_lens = 0;
    // lens tables (dof == 0): rebuilt when (lens, width, height, proportions) change
    double *d_col_sc = nullptr, *d_row_sc = nullptr;
    double *d_dir_x = nullptr, *d_dir_y = nullptr; int dir_w = 0, dir_h = 0;   // init.py:131-132 per column / row
    double *d_lens_x = nullptr, *d_lens_y = nullptr; double dir_lens = -1, dir_prop = -1;   // init.py:41-42 without the jitter
    int tab_w = 0, tab_h = 0; double tab_lens = -1, tab_prop = -1;

    // world
    bool have_world = false;
    int origin[3] = {0, 0, 0}, dims[3] = {0, 0, 0}, nb[3] = {0, 0, 0}, nc[3] = {0, 0, 0};
    uint8_t *d_chunk_res = nullptr;
    uint32_t *d_brick_tab = nullptr;
    std::vector<uint32_t> h_brick_tab;          // host mirror: 0 = empty else slot + 1
    int world_cs = 0;                           // chunk_size the chunk table was sized for
    bool mirror_stale = false;                  // vt_upload_dense numbered the slots on the device: mirror not fetched yet
    // scene-building scratch (grow-only device buffers) and one pinned word for the device-side brick count
    void *scene_buf[3] = {nullptr, nullptr, nullptr}; size_t scene_cap[3] = {0, 0, 0};
    unsigned long long *h_word = nullptr;
    uint8_t *d_pool = nullptr;
    long long pool_cap = 0, pool_used = 0;      // in bricks
    std::vector<uint32_t> free_slots;
    double *d_mats = nullptr;                   // [256][8]
    vt_phys_material *d_phys = nullptr; double *d_phys_out = nullptr;     // physics slice test: material table, result

    // persistent kernel state: per-pixel accumulators (zero between launches), ray queue, launch geometry
    unsigned long long *d_acc = nullptr; long long acc_cap = 0;     // in u64 words
    unsigned long long *d_queue = nullptr;
    unsigned long long *d_pow_tab = nullptr; double pow_tab_falloff = 0.0; bool pow_tab_on = true;   // VT_POWTAB=0 disables
    int persist_blocks[8] = {0, 0, 0, 0, 0, 0, 0, 0};                          // resident blocks per SM, per instantiation
    int sm_count = 0;
    int max_res = 0;                                                // largest Frame.resolution in the chunk table
    int brick_runs = 1;                                             // VT_RUNS=0 disables empty-brick runs (A/B)
    int refill_min = 32;                                            // vt_set_tuning("refill")
    int group_max = 1;                                              // vt_set_tuning("group"): lanes per pixel of a packet
    int build_rows = 4;                                             // vt_set_tuning("buildrows"): rows per thread of the wide world build (0: off)
    int debug = 0;
    uint32_t *visited_peer = nullptr; int visited_words = 0;        // vt_set_visited_peer
    // multi-GPU frame target (vt_set_frame_target) and the status word of the wait kernel
    uint8_t *frame_target = nullptr; unsigned int *arrive_flag = nullptr; unsigned int arrive_value = 0;
    int *d_wait_status = nullptr; int clock_khz = 0;
    struct Stage { void *h = nullptr; size_t cap = 0; cudaEvent_t ev = nullptr; };
    Stage stage[8]; int stage_next = 0;
    // post-FX scratch (vt_fx_draw): grow-only device buffers
    uint8_t *fx_buf[4] = {nullptr, nullptr, nullptr, nullptr}; long long fx_cap[4] = {0, 0, 0, 0};
    unsigned long long *d_fx_sums = nullptr; int *d_fx_avg = nullptr;
    // region row map
    int32_t *d_row_y = nullptr; int row_cap = 0;
    std::vector<int32_t> h_row_y; int row_key[7] = {-1, -1, -1, -1, -1, -1, -1};
};

static std::string g_create_err;

#define VT_CHECK(ctx, call)                                                                         \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                        \
            return 2;                                                                               \
        }                                                                                           \
    } while (0)

static int fail(vt_ctx *ctx, const char *msg) { ctx->err = msg; return 1; }

extern "C" int vt_abi_version(void) { return VT_ABI_VERSION; }

extern "C" int vt_create(int device, vt_ctx **out) {
    if (!out) { g_create_err = "vt_create: out is NULL"; return 1; }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_create_err = std::string("vt_create: no usable CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback";
        return 2;
    }
    if (device < 0 || device >= n) { g_create_err = "vt_create: device ordinal out of range"; return 1; }
    vt_ctx *ctx = new vt_ctx();
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->d_mats, 256 * 8 * sizeof(double))) != cudaSuccess ||
        (e = cudaMemset(ctx->d_mats, 0, 256 * 8 * sizeof(double))) != cudaSuccess) {
        g_create_err = std::string("vt_create: ") + cudaGetErrorString(e);
        delete ctx;
        return 2;
    }
    ctx->own_stream = true;
    if ((e = cudaMalloc(&ctx->d_queue, 2 * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaMemset(ctx->d_queue, 0, 2 * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) {
        g_create_err = std::string("vt_create: ") + cudaGetErrorString(e);
        delete ctx;
        return 2;
    }
    if ((e = cudaMalloc(&ctx->d_pow_tab, VT_POW_SLOTS * 2 * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaMemset(ctx->d_pow_tab, 0, VT_POW_SLOTS * 2 * sizeof(unsigned long long))) != cudaSuccess) {
        g_create_err = std::string("vt_create: ") + cudaGetErrorString(e);
        delete ctx;
        return 2;
    }
    *out = ctx;
    return 0;
}

// Scheduling knobs of the render kernel for A/B measurements (tools/kbench.py).  None of them changes a result; the
// product never calls this (the defaults are the measured optimum).
extern "C" int vt_set_tuning(vt_ctx *ctx, const char *key, int value) {
    if (!ctx || !key) return 1;
    const std::string k(key);
    if (k == "runs") ctx->brick_runs = value != 0;                  // 0: no brick-local loops
    else if (k == "powtab") ctx->pow_tab_on = value != 0;           // 0: no memoised falloff power
    else if (k == "refill") { if (value < 1 || value > 32) return fail(ctx, "vt_set_tuning: refill must be 1..32"); ctx->refill_min = value; }
    else if (k == "group") { if (value < 1 || value > 32) return fail(ctx, "vt_set_tuning: group must be 1..32"); ctx->group_max = value; }
    else if (k == "buildrows") { if (value != 0 && value != 2 && value != 4) return fail(ctx, "vt_set_tuning: buildrows must be 0, 2 or 4"); ctx->build_rows = value; }
    else if (k == "debug") ctx->debug = value;
    else return fail(ctx, "vt_set_tuning: unknown key");
    return 0;
}

static void free_world(vt_ctx *ctx) {
    cudaFree(ctx->d_chunk_res); cudaFree(ctx->d_brick_tab); cudaFree(ctx->d_pool);
    ctx->d_chunk_res = nullptr; ctx->d_brick_tab = nullptr; ctx->d_pool = nullptr;
    ctx->pool_cap = ctx->pool_used = 0; ctx->free_slots.clear(); ctx->h_brick_tab.clear();
    ctx->mirror_stale = false;
    ctx->have_world = false;
}

extern "C" void vt_destroy(vt_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    free_world(ctx);
    cudaFree(ctx->d_mats); cudaFree(ctx->d_col_sc); cudaFree(ctx->d_row_sc); cudaFree(ctx->d_row_y);
    cudaFree(ctx->d_acc); cudaFree(ctx->d_queue); cudaFree(ctx->d_pow_tab);
    if (ctx->d_wait_status) cudaFreeHost(ctx->d_wait_status);
    for (int k = 0; k < 4; k++) cudaFree(ctx->fx_buf[k]);
    for (int k = 0; k < 3; k++) cudaFree(ctx->scene_buf[k]);
    cudaFree(ctx->d_phys); cudaFree(ctx->d_phys_out);
    if (ctx->h_word) cudaFreeHost(ctx->h_word);
    for (int k = 0; k < VT_STAGES; k++) { if (ctx->stage[k].h) cudaFreeHost(ctx->stage[k].h); if (ctx->stage[k].ev) cudaEventDestroy(ctx->stage[k].ev); }
    cudaFree(ctx->d_fx_sums); cudaFree(ctx->d_fx_avg); cudaFree(ctx->d_dir_x); cudaFree(ctx->d_dir_y);
    cudaFree(ctx->d_lens_x); cudaFree(ctx->d_lens_y);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int vt_build_flags(void) { return (VT_CHECKED ? 1 : 0) | (VT_TRIPS ? 2 : 0); }
extern "C" const char *vt_last_error(const vt_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }
extern "C" int64_t vt_launch_count(const vt_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int vt_set_stream(vt_ctx *ctx, void *cuda_stream) {
    if (!ctx) return 1;
    if (ctx->stream == (cudaStream_t)cuda_stream) return 0;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->own_stream) { cudaStreamDestroy(ctx->stream); ctx->own_stream = false; }
    ctx->stream = (cudaStream_t)cuda_stream;
    return 0;
}

extern "C" int vt_sync(vt_ctx *ctx) {
    if (!ctx) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int vt_set_settings(vt_ctx *ctx, const vt_settings *hs) {
    if (!ctx || !hs) return 1;
    if (hs->width <= 0 || hs->height <= 0) return fail(ctx, "vt_set_settings: width/height must be positive");
    if (hs->chunk_size <= 0 || hs->chunk_size > 4096) return fail(ctx, "vt_set_settings: chunk_size out of range");
    if (hs->samples < 0) return fail(ctx, "vt_set_settings: samples must be >= 0");
    if (!(hs->dist_max < 1e8) || !(hs->dist_min > -1e8)) return fail(ctx, "vt_set_settings: dist_min/dist_max out of range");
    if (ctx->have_world && ctx->have_settings && hs->chunk_size != ctx->s.chunk_size)
        return fail(ctx, "vt_set_settings: chunk_size changed while a world is loaded; call vt_set_world again");
    ctx->s = *hs;
    ctx->have_settings = true;
    ctx->cs_shift = -1;
    for (int k = 0; k < 13; k++) if ((1 << k) == hs->chunk_size) ctx->cs_shift = k;
    return 0;
}

// Small per-frame tables (materials, chunk resolutions) go through a ring of pinned staging slots: the caller's
// buffer is free on return and the host never waits for the frame in flight.
static int staged_upload(vt_ctx *ctx, void *dst, const void *src, size_t bytes) {
    vt_ctx::Stage &st = ctx->stage[ctx->stage_next];
    ctx->stage_next = (ctx->stage_next + 1) % VT_STAGES;
    if (st.ev) VT_CHECK(ctx, cudaEventSynchronize(st.ev));
    else VT_CHECK(ctx, cudaEventCreateWithFlags(&st.ev, cudaEventDisableTiming));
    if (st.cap < bytes) {
        if (st.h) cudaFreeHost(st.h);
        st.h = nullptr; st.cap = 0;
        size_t cap = bytes < 65536 ? 65536 : bytes;
        VT_CHECK(ctx, cudaMallocHost(&st.h, cap));
        st.cap = cap;
    }
    memcpy(st.h, src, bytes);
    VT_CHECK(ctx, cudaMemcpyAsync(dst, st.h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaEventRecord(st.ev, ctx->stream));
    return 0;
}

extern "C" int vt_set_materials(vt_ctx *ctx, const vt_material *hm, int n) {
    if (!ctx || !hm) return 1;
    if (n < 0 || n > VT_MAX_MATERIALS) return fail(ctx, "vt_set_materials: at most 255 materials (id 0 is empty)");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    return staged_upload(ctx, ctx->d_mats, hm, (size_t)(n + 1) * sizeof(vt_material));
}

extern "C" int vt_set_camera(vt_ctx *ctx, const double pos[3], const double rot[4], double lens) {
    if (!ctx || !pos || !rot) return 1;
    for (int a = 0; a < 3; a++)
        if (!(fabs(pos[a]) < 1e9)) return fail(ctx, "vt_set_camera: position out of range (|p| < 1e9)");
    memcpy(ctx->cam_pos, pos, sizeof ctx->cam_pos);
    memcpy(ctx->cam_rot, rot, sizeof ctx->cam_rot);
    ctx->cam_lens = lens;
    return 0;
}

// ---- world ------------------------------------------------------------------------------------------
extern "C" int vt_set_world(vt_ctx *ctx, const int32_t origin[3], const int32_t dims[3]) {
    if (!ctx || !origin || !dims) return 1;
    if (!ctx->have_settings) return fail(ctx, "vt_set_world: call vt_set_settings first (chunk_size)");
    const int cs = ctx->s.chunk_size;
    long long slots = 1, bricks = 1;
    for (int a = 0; a < 3; a++) {
        if (dims[a] <= 0 || dims[a] % cs != 0) return fail(ctx, "vt_set_world: dims must be positive multiples of chunk_size");
        int r = origin[a] % cs;
        if (r != 0) return fail(ctx, "vt_set_world: origin must be a multiple of chunk_size");
        if (fabs((double)origin[a]) > 1e9 || dims[a] > (1 << 20)) return fail(ctx, "vt_set_world: box out of range");
        slots *= dims[a] / cs; bricks *= (dims[a] + VT_BRICK - 1) / VT_BRICK;
    }
    if (bricks > (1LL << 31) || slots > (1LL << 31)) return fail(ctx, "vt_set_world: world too large");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    // The brick pool is kept across worlds (its contents are dropped, its memory is reused: a rebuild packs into it
    // without waiting for the brick count) and so are the tables when the box has the same shape.
    bool same = ctx->have_world && ctx->world_cs == cs;
    for (int a = 0; a < 3; a++) same = same && ctx->dims[a] == dims[a];
    if (!same) {
        VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_chunk_res); cudaFree(ctx->d_brick_tab);
        ctx->d_chunk_res = nullptr; ctx->d_brick_tab = nullptr;
        ctx->have_world = false;
        VT_CHECK(ctx, cudaMalloc(&ctx->d_chunk_res, (size_t)slots));
        VT_CHECK(ctx, cudaMalloc(&ctx->d_brick_tab, (size_t)bricks * sizeof(uint32_t)));
    }
    for (int a = 0; a < 3; a++) {
        ctx->origin[a] = origin[a]; ctx->dims[a] = dims[a];
        ctx->nb[a] = (dims[a] + VT_BRICK - 1) / VT_BRICK; ctx->nc[a] = dims[a] / cs;
    }
    VT_CHECK(ctx, cudaMemsetAsync(ctx->d_chunk_res, 0, (size_t)slots, ctx->stream));
    VT_CHECK(ctx, cudaMemsetAsync(ctx->d_brick_tab, 0, (size_t)bricks * sizeof(uint32_t), ctx->stream));
    ctx->h_brick_tab.assign((size_t)bricks, 0u);
    ctx->pool_used = 0; ctx->free_slots.clear(); ctx->mirror_stale = false;
    ctx->max_res = 0; ctx->world_cs = cs;
    ctx->have_world = true;
    return 0;
}

// Drops the world (tables and brick pool released): afterwards vt_set_settings accepts another chunk_size.
extern "C" int vt_clear_world(vt_ctx *ctx) {
    if (!ctx) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    free_world(ctx);
    for (int a = 0; a < 3; a++) { ctx->dims[a] = 0; ctx->nb[a] = 0; ctx->nc[a] = 0; }
    ctx->world_cs = 0; ctx->max_res = 0;
    return 0;
}

extern "C" int vt_set_chunk_table(vt_ctx *ctx, const uint8_t *host_res, int64_t n_slots) {
    if (!ctx || !host_res) return 1;
    if (!ctx->have_world) return fail(ctx, "vt_set_chunk_table: no world (vt_set_world)");
    long long slots = (long long)ctx->nc[0] * ctx->nc[1] * ctx->nc[2];
    if (n_slots != slots) return fail(ctx, "vt_set_chunk_table: n_slots does not match the world's chunk grid");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    int mx = 0;
    for (long long k = 0; k < slots; k++) mx = host_res[k] > mx ? host_res[k] : mx;
    ctx->max_res = mx;
    return staged_upload(ctx, ctx->d_chunk_res, host_res, (size_t)slots);
}

static int grow_pool(vt_ctx *ctx, long long need) {
    if (need <= ctx->pool_cap) return 0;
    long long cap = ctx->pool_cap ? ctx->pool_cap : 1024;
    while (cap < need) cap *= 2;
    uint8_t *np = nullptr;
    VT_CHECK(ctx, cudaMalloc(&np, (size_t)cap * 512));
    if (ctx->d_pool && ctx->pool_used)
        VT_CHECK(ctx, cudaMemcpyAsync(np, ctx->d_pool, (size_t)ctx->pool_used * 512, cudaMemcpyDeviceToDevice, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->d_pool);
    ctx->d_pool = np; ctx->pool_cap = cap;
    return 0;
}

// Grow-only device scratch of the scene-building calls (object list, task tables, flags): no cudaMalloc / cudaFree --
// each of which synchronises the device -- on the per-update path.
static int scene_scratch(vt_ctx *ctx, int which, size_t bytes, void **out) {
    if (bytes == 0) bytes = 16;
    if (ctx->scene_cap[which] < bytes) {
        VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->scene_buf[which]);
        ctx->scene_buf[which] = nullptr; ctx->scene_cap[which] = 0;
        size_t cap = 4096;
        while (cap < bytes) cap *= 2;
        VT_CHECK(ctx, cudaMalloc(&ctx->scene_buf[which], cap));
        ctx->scene_cap[which] = cap;
    }
    *out = ctx->scene_buf[which];
    return 0;
}

// The host mirror of the brick table serves the in-place updates (slot reuse, free list).  vt_upload_dense numbers the
// slots on the device and leaves the mirror stale; the first in-place update after it fetches the table once.
static int ensure_mirror(vt_ctx *ctx) {
    if (!ctx->mirror_stale) return 0;
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->h_brick_tab.data(), ctx->d_brick_tab, ctx->h_brick_tab.size() * sizeof(uint32_t),
                                  cudaMemcpyDeviceToHost, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->mirror_stale = false;
    return 0;
}

static int check_objects(vt_ctx *ctx, const vt_object *objs, int n, const char *who) {
    for (int k = 0; k < n; k++) {
        if (!objs[k].dev_sprite) { ctx->err = std::string(who) + ": object without a sprite grid"; return 1; }
        for (int a = 0; a < 3; a++) {
            if (objs[k].size[a] <= 0) { ctx->err = std::string(who) + ": sprite size must be positive"; return 1; }
            if (objs[k].angle[a] < 0 || objs[k].angle[a] > 3) { ctx->err = std::string(who) + ": angle must be 0..3"; return 1; }
        }
    }
    return 0;
}

// object list -> device (scratch 0), through the pinned staging ring
static int stage_objects(vt_ctx *ctx, const vt_object *objs, int n, vt_object **d_objs) {
    void *p = nullptr;
    if (int rc = scene_scratch(ctx, 0, (size_t)n * sizeof(vt_object), &p)) return rc;
    *d_objs = static_cast<vt_object *>(p);
    if (n) return staged_upload(ctx, p, objs, (size_t)n * sizeof(vt_object));
    return 0;
}

static bool dense_vec(const vt_ctx *ctx, const uint8_t *dev_cells) {
    return ctx->dims[2] % 8 == 0 && (reinterpret_cast<uintptr_t>(dev_cells) & 7u) == 0;
}

extern "C" int vt_upload_bricks(vt_ctx *ctx, const int32_t *coords, const uint8_t *cells, int m) {
    if (!ctx || (m > 0 && (!coords || !cells))) return 1;
    if (!ctx->have_world) return fail(ctx, "vt_upload_bricks: no world (vt_set_world)");
    if (m <= 0) return 0;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    if (int rc = ensure_mirror(ctx)) return rc;
    for (int b = 0; b < m; b++)
        for (int a = 0; a < 3; a++)
            if (coords[3 * b + a] < 0 || coords[3 * b + a] >= ctx->nb[a]) return fail(ctx, "vt_upload_bricks: brick coordinate outside the world");
    std::vector<uint32_t> meta((size_t)m * 3);       // slots | table index | table value
    uint32_t *slots = meta.data(), *tidx = slots + m, *tval = tidx + m;
    long long need = ctx->pool_used;
    for (int b = 0; b < m; b++) {
        const int32_t *c = coords + 3 * b;
        uint32_t ti = (uint32_t)(((long long)c[0] * ctx->nb[1] + c[1]) * ctx->nb[2] + c[2]);
        bool empty = true;
        const uint64_t *q = reinterpret_cast<const uint64_t *>(cells + (size_t)b * 512);
        uint64_t acc = 0;
        if ((reinterpret_cast<uintptr_t>(q) & 7) == 0) { for (int k = 0; k < 64; k++) acc |= q[k]; empty = acc == 0; }
        else { for (int k = 0; k < 512; k++) if (cells[(size_t)b * 512 + k]) { empty = false; break; } }
        uint32_t cur = ctx->h_brick_tab[ti];
        if (empty) {
            if (cur) ctx->free_slots.push_back(cur - 1);
            slots[b] = 0xffffffffu; tval[b] = 0;
        } else {
            uint32_t slot;
            if (cur) slot = cur - 1;
            else if (!ctx->free_slots.empty()) { slot = ctx->free_slots.back(); ctx->free_slots.pop_back(); }
            else slot = (uint32_t)need++;
            slots[b] = slot; tval[b] = slot + 1;
        }
        tidx[b] = ti;
        ctx->h_brick_tab[ti] = tval[b];
    }
    if (int rc = grow_pool(ctx, need)) return rc;
    ctx->pool_used = need;
    void *d_stage = nullptr, *d_meta_v = nullptr;
    if (int rc = scene_scratch(ctx, 1, (size_t)m * 512, &d_stage)) return rc;
    if (int rc = scene_scratch(ctx, 2, meta.size() * sizeof(uint32_t), &d_meta_v)) return rc;
    uint32_t *d_meta = static_cast<uint32_t *>(d_meta_v);
    VT_CHECK(ctx, cudaMemcpyAsync(d_stage, cells, (size_t)m * 512, cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaMemcpyAsync(d_meta, meta.data(), meta.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    scatter_bricks_kernel<<<m, 32, 0, ctx->stream>>>(static_cast<const uint8_t *>(d_stage), d_meta, d_meta + m, d_meta + 2 * m, ctx->d_pool,
                                                     ctx->d_brick_tab, m);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));       // `cells` and `meta` are free on return
    return 0;
}

// Dense world -> brick pool + brick table, entirely on the device: occupancy bits (one pass over the cells), an
// exclusive scan of the per-task counts (slots in brick-index order), then the pack pass (second read of the cells,
// partly from L2; one 512-byte store per non-empty brick).  The only host step is reading the brick COUNT back
// (8 bytes, pinned) -- after the pack pass when a pool exists already (the pass is repeated if the pool turns out too
// small).  A single-pass variant (occupancy, slot numbering by decoupled look-back and the stores in one kernel, the
// grid read once) was measured and lost: 123-129 us against 67 us for the three kernels on the 512^3 terrain -- the
// look-back chain over 8 192 tasks costs more than the second, L2-assisted read.
static int upload_dense_split(vt_ctx *ctx, const uint8_t *dev_cells, int G, long long n_tasks, bool vec) {
    void *p = nullptr;
    if (int rc = scene_scratch(ctx, 1, (size_t)n_tasks * 3 * sizeof(uint32_t) + 16, &p)) return rc;
    unsigned long long *d_total = static_cast<unsigned long long *>(p);
    uint32_t *d_mask = reinterpret_cast<uint32_t *>(d_total + 2), *d_count = d_mask + n_tasks, *d_base = d_count + n_tasks;
    const int sms = ctx->sm_count > 0 ? ctx->sm_count : 148;
    long long g1 = (n_tasks + 7) / 8;
    if (g1 > (long long)sms * 8) g1 = (long long)sms * 8;
    brick_mask_kernel<<<(int)g1, 256, 0, ctx->stream>>>(dev_cells, ctx->dims[0], ctx->dims[1], ctx->dims[2], ctx->nb[1], ctx->nb[2], G, n_tasks,
                                                        vec, d_mask, d_count);
    task_scan_kernel<<<1, 1024, 0, ctx->stream>>>(d_count, n_tasks, d_base, d_total);
    ctx->launches += 2;
    VT_CHECK(ctx, cudaGetLastError());
    const long long g2 = n_tasks < (long long)sms * 8 ? n_tasks : (long long)sms * 8;
    const bool optimistic = ctx->pool_cap > 0;
    if (optimistic) {
        brick_pack_kernel<<<(int)g2, 256, 0, ctx->stream>>>(dev_cells, ctx->dims[0], ctx->dims[1], ctx->dims[2], ctx->nb[1], ctx->nb[2], G,
                                                            n_tasks, vec, d_mask, d_base, (uint32_t)ctx->pool_cap, ctx->d_pool, ctx->d_brick_tab);
        ctx->launches++;
        VT_CHECK(ctx, cudaGetLastError());
    }
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->h_word, d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));       // also: the caller may release dev_cells on return
    const long long used = (long long)*ctx->h_word;
    if (!optimistic || used > ctx->pool_cap) {
        if (int rc = grow_pool(ctx, used)) return rc;
        brick_pack_kernel<<<(int)g2, 256, 0, ctx->stream>>>(dev_cells, ctx->dims[0], ctx->dims[1], ctx->dims[2], ctx->nb[1], ctx->nb[2], G,
                                                            n_tasks, vec, d_mask, d_base, (uint32_t)ctx->pool_cap, ctx->d_pool, ctx->d_brick_tab);
        ctx->launches++;
        VT_CHECK(ctx, cudaGetLastError());
        VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    }
    ctx->pool_used = used;
    return 0;
}

extern "C" int vt_upload_dense(vt_ctx *ctx, const uint8_t *dev_cells) {
    if (!ctx || !dev_cells) return 1;
    if (!ctx->have_world) return fail(ctx, "vt_upload_dense: no world (vt_set_world)");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    const int G = (ctx->nb[2] + 31) / 32;
    const long long n_tasks = (long long)ctx->nb[0] * ctx->nb[1] * G;
    const bool vec = dense_vec(ctx, dev_cells);
    if (!ctx->h_word) VT_CHECK(ctx, cudaMallocHost(&ctx->h_word, 64));
    ctx->free_slots.clear();
    ctx->pool_used = 0;                                      // nothing to preserve
    if (int rc = upload_dense_split(ctx, dev_cells, G, n_tasks, vec)) return rc;
    ctx->mirror_stale = true;
    return 0;
}

// World cells of the box [l, l + e) from the object list: the wide kernel (16 cells x 4 rows per thread, vt_scene.cuh) when
// rows are multiples of 16 cells and 16-byte aligned, build_cells_kernel otherwise.
static void launch_build_cells(vt_ctx *ctx, const vt_object *d_objs, const vt_object *h_objs, int n, const int l[3], const int e[3], uint8_t *dev_cells) {
    bool rotated = false;                                    // cells of rotated objects go cell by cell: build_cells_kernel does that inline
    for (int k = 0; k < n; k++) rotated = rotated || (h_objs[k].angle[0] | h_objs[k].angle[1] | h_objs[k].angle[2]) != 0;
    const int sms = ctx->sm_count > 0 ? ctx->sm_count : 148;
    const int ny = ctx->dims[1], nz = ctx->dims[2];
    const long long rows = (long long)e[0] * e[1];
    const bool wide = ctx->build_rows > 0 && !rotated && nz % 16 == 0 && l[2] % 16 == 0 && e[2] % 16 == 0 && e[2] > 0 && rows > 0 &&
                      (reinterpret_cast<uintptr_t>(dev_cells) & 15u) == 0 && rows * (e[2] / 16) < (1LL << 31);
    if (wide) {
        const int R = ctx->build_rows, chunks = e[2] / 16;
        const long long items = ((rows + R - 1) / R) * chunks, want = (items + 255) / 256;
        const int grid = (int)(want < (long long)sms * 16 ? want : (long long)sms * 16);
        const unsigned long long mc = chunks > 1 ? ~0ull / (unsigned)chunks + 1ull : 0ull, me = e[1] > 1 ? ~0ull / (unsigned)e[1] + 1ull : 0ull;
        if (R == 2)
            build_cells16_kernel<2><<<grid, 256, 0, ctx->stream>>>(d_objs, n, ctx->origin[0], ctx->origin[1], ctx->origin[2], ny, nz, l[0], l[1], l[2],
                                                                   e[0], e[1], chunks, mc, me, dev_cells);
        else
            build_cells16_kernel<4><<<grid, 256, 0, ctx->stream>>>(d_objs, n, ctx->origin[0], ctx->origin[1], ctx->origin[2], ny, nz, l[0], l[1], l[2],
                                                                   e[0], e[1], chunks, mc, me, dev_cells);
        return;
    }
    const long long total = rows * ((e[2] + 7) / 8);
    long long want = (total + 255) / 256;
    const int grid = (int)(want < (long long)sms * 32 ? (want > 0 ? want : 1) : (long long)sms * 32);
    build_cells_kernel<<<grid, 256, 0, ctx->stream>>>(d_objs, n, ctx->origin[0], ctx->origin[1], ctx->origin[2], ny, nz, l[0], l[1], l[2], e[0], e[1], e[2],
                                                      dense_vec(ctx, dev_cells), dev_cells);
}

extern "C" int vt_build_world(vt_ctx *ctx, const vt_object *objs, int n, uint8_t *dev_cells) {
    if (!ctx || !dev_cells || n < 0 || (n && !objs)) return 1;
    if (!ctx->have_world) return fail(ctx, "vt_build_world: no world (vt_set_world)");
    if (int rc = check_objects(ctx, objs, n, "vt_build_world")) return rc;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    vt_object *d_objs = nullptr;
    if (int rc = stage_objects(ctx, objs, n, &d_objs)) return rc;
    const int l0[3] = {0, 0, 0};
    launch_build_cells(ctx, d_objs, objs, n, l0, ctx->dims, dev_cells);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

// Objects that moved: redraw the cells of the chunk-aligned box [lo, hi) (absolute cells, clipped to the world) from
// the whole object list, then bring the bricks of that box up to date -- flags on the device, slot assignment on the
// host (free list), fill + table patch on the device.  Cost scales with the box, not with the world.
extern "C" int vt_update_region(vt_ctx *ctx, const vt_object *objs, int n, uint8_t *dev_cells, const int32_t lo[3], const int32_t hi[3]) {
    if (!ctx || !dev_cells || !lo || !hi || n < 0 || (n && !objs)) return 1;
    if (!ctx->have_world) return fail(ctx, "vt_update_region: no world (vt_set_world)");
    if (int rc = check_objects(ctx, objs, n, "vt_update_region")) return rc;
    int l[3], e[3], b0[3], rb[3];
    for (int a = 0; a < 3; a++) {
        long long p = (long long)lo[a] - ctx->origin[a], q = (long long)hi[a] - ctx->origin[a];
        if (p < 0) p = 0;
        if (q > ctx->dims[a]) q = ctx->dims[a];
        if (q <= p) return 0;                                  // nothing of the box lies inside the world
        p = (p / VT_BRICK) * VT_BRICK;                         // whole bricks
        q = ((q + VT_BRICK - 1) / VT_BRICK) * VT_BRICK;
        if (q > ctx->dims[a]) q = ctx->dims[a];
        l[a] = (int)p; e[a] = (int)(q - p);
        b0[a] = (int)(p / VT_BRICK); rb[a] = (int)((q - p + VT_BRICK - 1) / VT_BRICK);
    }
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    if (int rc = ensure_mirror(ctx)) return rc;
    const int m = rb[0] * rb[1] * rb[2];
    vt_object *d_objs = nullptr;
    void *p1 = nullptr, *p2 = nullptr;
    if (int rc = stage_objects(ctx, objs, n, &d_objs)) return rc;
    if (int rc = scene_scratch(ctx, 1, (size_t)m, &p1)) return rc;
    if (int rc = scene_scratch(ctx, 2, (size_t)m * sizeof(uint32_t), &p2)) return rc;
    uint8_t *d_flags = static_cast<uint8_t *>(p1);
    uint32_t *d_slots = static_cast<uint32_t *>(p2);
    const int sms = ctx->sm_count > 0 ? ctx->sm_count : 148;
    launch_build_cells(ctx, d_objs, objs, n, l, e, dev_cells);
    region_flags_kernel<<<(m + 3) / 4 < sms * 16 ? (m + 3) / 4 : sms * 16, 128, 0, ctx->stream>>>(
        dev_cells, ctx->dims[0], ctx->dims[1], ctx->dims[2], b0[0], b0[1], b0[2], rb[0], rb[1], rb[2], d_flags);
    ctx->launches += 2;
    VT_CHECK(ctx, cudaGetLastError());
    std::vector<uint8_t> flags((size_t)m);
    VT_CHECK(ctx, cudaMemcpyAsync(flags.data(), d_flags, (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    std::vector<uint32_t> slots((size_t)m);
    long long need = ctx->pool_used;
    for (int i = 0; i < m; i++) {
        const int bz = b0[2] + i % rb[2], by = b0[1] + (i / rb[2]) % rb[1], bx = b0[0] + i / (rb[2] * rb[1]);
        const size_t ti = ((size_t)bx * ctx->nb[1] + by) * ctx->nb[2] + bz;
        const uint32_t cur = ctx->h_brick_tab[ti];
        if (!flags[(size_t)i]) {
            if (cur) ctx->free_slots.push_back(cur - 1);
            slots[(size_t)i] = 0xffffffffu; ctx->h_brick_tab[ti] = 0;
        } else {
            uint32_t slot;
            if (cur) slot = cur - 1;
            else if (!ctx->free_slots.empty()) { slot = ctx->free_slots.back(); ctx->free_slots.pop_back(); }
            else slot = (uint32_t)need++;
            slots[(size_t)i] = slot; ctx->h_brick_tab[ti] = slot + 1;
        }
    }
    if (int rc = grow_pool(ctx, need)) return rc;
    ctx->pool_used = need;
    VT_CHECK(ctx, cudaMemcpyAsync(d_slots, slots.data(), (size_t)m * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    region_fill_kernel<<<m < sms * 8 ? m : sms * 8, 512, 0, ctx->stream>>>(dev_cells, ctx->dims[0], ctx->dims[1], ctx->dims[2], ctx->nb[1],
                                                                       ctx->nb[2], b0[0], b0[1], b0[2], rb[1], rb[2], d_slots, m,
                                                                       ctx->d_pool, ctx->d_brick_tab);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));       // `slots` is free on return
    return 0;
}

// Bricks -> dense grid in DEVICE memory (the inverse of vt_upload_dense).
extern "C" int vt_download_dense(vt_ctx *ctx, uint8_t *dev_cells) {
    if (!ctx || !dev_cells) return 1;
    if (!ctx->have_world) return fail(ctx, "vt_download_dense: no world (vt_set_world)");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    const long long total = (long long)ctx->dims[0] * ctx->dims[1] * ((ctx->dims[2] + 7) / 8);
    long long want = (total + 255) / 256;
    unpack_cells_kernel<<<(int)(want < 148LL * 32 ? want : 148LL * 32), 256, 0, ctx->stream>>>(ctx->d_brick_tab, ctx->d_pool, ctx->dims[0], ctx->dims[1],
                                                                                                ctx->dims[2], ctx->nb[1], ctx->nb[2],
                                                                                                dense_vec(ctx, dev_cells), dev_cells);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

// Chunk slots that hold a voxel (what `post_chunk in self.chunks` answers after chunk_update, init.py:441-452, before
// LOD / culling), from the brick table: [ncx][ncy][ncz] bytes, z fastest, to HOST memory.  chunk_size % 8 == 0 only.
extern "C" int vt_chunk_presence(vt_ctx *ctx, uint8_t *host_present, int64_t n_slots) {
    if (!ctx || !host_present) return 1;
    if (!ctx->have_world) return fail(ctx, "vt_chunk_presence: no world (vt_set_world)");
    const int cs = ctx->s.chunk_size;
    if (cs % VT_BRICK != 0) return fail(ctx, "vt_chunk_presence: chunk_size must be a multiple of 8");
    const long long slots = (long long)ctx->nc[0] * ctx->nc[1] * ctx->nc[2];
    if (n_slots != slots) return fail(ctx, "vt_chunk_presence: n_slots does not match the world's chunk grid");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    void *p = nullptr;
    if (int rc = scene_scratch(ctx, 2, (size_t)slots, &p)) return rc;
    long long want = (slots + 127) / 128;
    chunk_present_kernel<<<(int)(want < 148LL * 32 ? want : 148LL * 32), 128, 0, ctx->stream>>>(ctx->d_brick_tab, ctx->nb[1], ctx->nb[2], ctx->nc[1],
                                                                                                 ctx->nc[2], slots, cs / VT_BRICK,
                                                                                                 static_cast<uint8_t *>(p));
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    VT_CHECK(ctx, cudaMemcpyAsync(host_present, p, (size_t)slots, cudaMemcpyDeviceToHost, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---- multi-GPU: peer-visible frame + flags ---------------------------------------------------------------
extern "C" int vt_shared_alloc(vt_ctx *ctx, int64_t bytes, void **out) {
    if (!ctx || !out || bytes <= 0) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaMalloc(out, (size_t)bytes));            // plain cudaMalloc: base pointer == IPC handle base
    VT_CHECK(ctx, cudaMemset(*out, 0, (size_t)bytes));
    return 0;
}
extern "C" int vt_shared_free(vt_ctx *ctx, void *p) {
    if (!ctx) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    VT_CHECK(ctx, cudaFree(p));
    return 0;
}
extern "C" int vt_ipc_export(vt_ctx *ctx, void *dev_ptr, uint8_t handle[64]) {
    if (!ctx || !dev_ptr || !handle) return 1;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    VT_CHECK(ctx, cudaIpcGetMemHandle(&h, dev_ptr));
    memcpy(handle, &h, 64);
    return 0;
}
extern "C" int vt_ipc_open(vt_ctx *ctx, const uint8_t handle[64], void **out) {
    if (!ctx || !handle || !out) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    VT_CHECK(ctx, cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int vt_ipc_close(vt_ctx *ctx, void *p) {
    if (!ctx) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    VT_CHECK(ctx, cudaIpcCloseMemHandle(p));
    return 0;
}
extern "C" int vt_set_frame_target(vt_ctx *ctx, uint8_t *dev_frame, uint32_t *dev_arrive_flag, uint32_t arrive_value) {
    if (!ctx) return 1;
    ctx->frame_target = dev_frame; ctx->arrive_flag = dev_frame ? dev_arrive_flag : nullptr; ctx->arrive_value = arrive_value;
    return 0;
}
// Multi-GPU culling: with a frame target set, the last warp of every vt_render launch that has dev_visited ORs the
// launch's visited words into dev_peer_mask (n_words u32, typically in rank 0's memory, opened with vt_ipc_open) before
// it raises the arrival flag -- the union of the workers' traversed lists (init.py:189, 393).  NULL switches it off.
extern "C" int vt_set_visited_peer(vt_ctx *ctx, uint32_t *dev_peer_mask, int64_t n_words) {
    if (!ctx || n_words < 0 || n_words > (1LL << 30)) return 1;
    ctx->visited_peer = dev_peer_mask; ctx->visited_words = dev_peer_mask ? (int)n_words : 0;
    return 0;
}
extern "C" int vt_stream_wait_flags(vt_ctx *ctx, const uint32_t *dev_flags, int n, uint32_t value, double timeout_s) {
    if (!ctx || !dev_flags || n < 1 || n > 32) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    if (!ctx->d_wait_status) {                 // pinned host word the wait kernel writes on a timeout: the host reads it without a CUDA call
        VT_CHECK(ctx, cudaHostAlloc(&ctx->d_wait_status, sizeof(int), cudaHostAllocMapped));
        *ctx->d_wait_status = 0;
    }
    if (ctx->clock_khz == 0) {                 // queried once: this attribute is slow to read (~1 ms)
        int khz = 0;
        cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device);
        ctx->clock_khz = khz > 0 ? khz : 1500000;
    }
    long long cycles = (long long)((timeout_s > 0 ? timeout_s : 2.0) * 1e3 * (double)ctx->clock_khz);
    wait_flags_kernel<<<1, 32, 0, ctx->stream>>>(dev_flags, n, value, cycles, ctx->d_wait_status);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    return 0;
}
extern "C" int vt_stream_signal_flags_on(vt_ctx *ctx, void *cuda_stream, uint32_t *const *host_ptrs, int n, uint32_t value) {
    if (!ctx || !host_ptrs || n < 1 || n > 16) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    FlagTargets tg;
    memset(&tg, 0, sizeof tg);
    for (int k = 0; k < n; k++) tg.p[k] = host_ptrs[k];
    signal_flags_kernel<<<1, 32, 0, cuda_stream ? (cudaStream_t)cuda_stream : ctx->stream>>>(tg, n, value);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    return 0;
}
extern "C" int vt_stream_signal_flags(vt_ctx *ctx, uint32_t *const *host_ptrs, int n, uint32_t value) {
    return vt_stream_signal_flags_on(ctx, nullptr, host_ptrs, n, value);
}
extern "C" int vt_wait_status(vt_ctx *ctx, int *status) {
    if (!ctx || !status) return 1;
    *status = 0;
    if (!ctx->d_wait_status) return 0;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    *status = *(volatile int *)ctx->d_wait_status;
    *ctx->d_wait_status = 0;
    return 0;
}
extern "C" int vt_wait_status_peek(vt_ctx *ctx, int *status) {
    if (!ctx || !status) return 1;
    *status = ctx->d_wait_status ? *(volatile int *)ctx->d_wait_status : 0;
    if (*status) *ctx->d_wait_status = 0;
    return 0;
}
extern "C" int vt_copy_to_host(vt_ctx *ctx, void *host_dst, const void *dev_src, int64_t bytes) {
    if (!ctx || !host_dst || !dev_src || bytes < 0) return 1;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    VT_CHECK(ctx, cudaMemcpyAsync(host_dst, dev_src, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return 0;
}

extern "C" int vt_world_stats(vt_ctx *ctx, int64_t *n_bricks, int64_t *bytes) {
    if (!ctx) return 1;
    long long live = ctx->pool_used - (long long)ctx->free_slots.size();
    if (n_bricks) *n_bricks = live;
    if (bytes) *bytes = ctx->pool_cap * 512 + (long long)ctx->h_brick_tab.size() * 4 + (long long)ctx->nc[0] * ctx->nc[1] * ctx->nc[2];
    return 0;
}

// ---- render -----------------------------------------------------------------------------------------
static bool row_selected(const vt_region *r, int y) {
    int sr = r->stripe_rows > 0 ? r->stripe_rows : 1, sc = r->stripe_count > 0 ? r->stripe_count : 1;
    return ((y / sr) % sc) == r->stripe_index;
}

extern "C" int64_t vt_region_pixels(const vt_region *r) {
    if (!r) return 0;
    if (r->dev_pixels) return r->n_pixels;
    if (r->x1 <= r->x0 || r->y1 <= r->y0) return 0;
    int64_t rows = 0;
    for (int y = r->y0; y < r->y1; y++) rows += row_selected(r, y) ? 1 : 0;
    return rows * (int64_t)(r->x1 - r->x0);
}

// Lens tables for dof == 0: the lens angles are separable (init.py:41-42: lens_x depends on the column only,
// lens_y on the row only), so sin/cos of the half angles are evaluated once per column / row on the host with
// the same libm the reference's interpreter calls (math.sin / math.cos, lib.py:327-332).
static int update_lens_tables(vt_ctx *ctx) {
    const vt_settings &s = ctx->s;
    if (ctx->d_col_sc && ctx->tab_w == s.width && ctx->tab_h == s.height && ctx->tab_lens == ctx->cam_lens &&
        ctx->tab_prop == s.proportions)
        return 0;
    cudaFree(ctx->d_col_sc); cudaFree(ctx->d_row_sc); ctx->d_col_sc = ctx->d_row_sc = nullptr;
    std::vector<double> col((size_t)s.width * 2), row((size_t)s.height * 2);
    const double deg2rad = 3.14159265358979323846 / 180.0;
    for (int x = 0; x < s.width; x++) {
        double dir_x = -1.0 + ((double)x / (double)s.width) * 2.0;
        double lens_x = (dir_x / s.proportions) * ctx->cam_lens;
        double rad = (-lens_x) * deg2rad;
        col[2 * x] = sin(rad / 2.0); col[2 * x + 1] = cos(rad / 2.0);
    }
    for (int y = 0; y < s.height; y++) {
        double dir_y = -1.0 + ((double)y / (double)s.height) * 2.0;
        double lens_y = (dir_y * s.proportions) * ctx->cam_lens;
        double rad = (+lens_y) * deg2rad;
        row[2 * y] = sin(rad / 2.0); row[2 * y + 1] = cos(rad / 2.0);
    }
    VT_CHECK(ctx, cudaMalloc(&ctx->d_col_sc, col.size() * sizeof(double)));
    VT_CHECK(ctx, cudaMalloc(&ctx->d_row_sc, row.size() * sizeof(double)));
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_col_sc, col.data(), col.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_row_sc, row.data(), row.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->tab_w = s.width; ctx->tab_h = s.height; ctx->tab_lens = ctx->cam_lens; ctx->tab_prop = s.proportions;
    return 0;
}

// dir_x / dir_y of init.py:131-132 for every column / row: a correctly rounded division each, done once on the host
static int update_dir_tables(vt_ctx *ctx) {
    const vt_settings &s = ctx->s;
    if (ctx->d_dir_x && ctx->dir_w == s.width && ctx->dir_h == s.height && ctx->dir_lens == ctx->cam_lens && ctx->dir_prop == s.proportions) return 0;
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->d_dir_x); cudaFree(ctx->d_dir_y); cudaFree(ctx->d_lens_x); cudaFree(ctx->d_lens_y);
    ctx->d_dir_x = ctx->d_dir_y = ctx->d_lens_x = ctx->d_lens_y = nullptr;
    std::vector<double> dx((size_t)s.width), dy((size_t)s.height), lx((size_t)s.width), ly((size_t)s.height);
    for (int x = 0; x < s.width; x++) { dx[x] = -1.0 + ((double)x / (double)s.width) * 2.0; lx[x] = (dx[x] / s.proportions) * ctx->cam_lens; }
    for (int y = 0; y < s.height; y++) { dy[y] = -1.0 + ((double)y / (double)s.height) * 2.0; ly[y] = (dy[y] * s.proportions) * ctx->cam_lens; }
    VT_CHECK(ctx, cudaMalloc(&ctx->d_dir_x, dx.size() * sizeof(double)));
    VT_CHECK(ctx, cudaMalloc(&ctx->d_dir_y, dy.size() * sizeof(double)));
    VT_CHECK(ctx, cudaMalloc(&ctx->d_lens_x, lx.size() * sizeof(double)));
    VT_CHECK(ctx, cudaMalloc(&ctx->d_lens_y, ly.size() * sizeof(double)));
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_dir_x, dx.data(), dx.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_dir_y, dy.data(), dy.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_lens_x, lx.data(), lx.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_lens_y, ly.data(), ly.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->dir_w = s.width; ctx->dir_h = s.height; ctx->dir_lens = ctx->cam_lens; ctx->dir_prop = s.proportions;
    return 0;
}

static int update_row_map(vt_ctx *ctx, const vt_region *r, int *band_rows) {
    int key[7] = {r->x0, r->y0, r->x1, r->y1, r->stripe_rows, r->stripe_count, r->stripe_index};
    if (ctx->d_row_y && memcmp(key, ctx->row_key, sizeof key) == 0) { *band_rows = (int)ctx->h_row_y.size(); return 0; }
    ctx->h_row_y.clear();
    for (int y = r->y0; y < r->y1; y++) if (row_selected(r, y)) ctx->h_row_y.push_back(y);
    int n = (int)ctx->h_row_y.size();
    if (n > ctx->row_cap) {
        cudaFree(ctx->d_row_y); ctx->d_row_y = nullptr;
        VT_CHECK(ctx, cudaMalloc(&ctx->d_row_y, (size_t)n * sizeof(int32_t)));
        ctx->row_cap = n;
    }
    if (n) {
        VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_row_y, ctx->h_row_y.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
        VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    }
    memcpy(ctx->row_key, key, sizeof key);
    *band_rows = n;
    return 0;
}

// Launch geometry and per-launch tables of the persistent kernel (vt_march.cuh).
typedef void (*persistent_fn)(const DevParams, const int32_t *);
// index = FAST << 2 | REC << 1 | COUNT.  FAST = the common configuration (power-of-two chunk size, <= 64 samples,
// resolutions inside the snap table) compiled without the generic branches: 13 % less code, 8 % faster (the kernel
// is sensitive to instruction-cache misses).  REC (parity probe) always runs the full tail.
static persistent_fn persistent_kernel(int which) {
    switch (which) {
        case 0: return render_persistent<false, false, false, false>;
        case 1: return render_persistent<false, false, false, true>;
        case 2: return render_persistent<false, true, true, false>;
        case 3: return render_persistent<false, true, true, true>;
        case 4: return render_persistent<true, false, false, false>;
        case 5: return render_persistent<true, false, false, true>;
        case 6: return render_persistent<true, true, true, false>;
        default: return render_persistent<true, true, true, true>;
    }
}

static int launch_persistent(vt_ctx *ctx, DevParams &P, bool rec, bool count) {
    const vt_settings &s = ctx->s;
    if (s.width > 65535 || s.height > 65535) return fail(ctx, "vt_render: canvas larger than 65535 pixels per side");
    if (P.n_pixels > 0xffffffffLL) return fail(ctx, "vt_render: more than 2^32 pixels in one region");
    // most samples any pixel can ask for: detail <= 1 + max(0, -lod_edge) (init.py:133-134)
    double dmax = 1.0 + (s.lod_edge < 0.0 ? -s.lod_edge : 0.0);
    long long smax = (long long)nearbyint((double)s.samples * dmax);
    if (smax < 1) smax = 1;
    if (smax > (1 << 20) - 1) return fail(ctx, "vt_render: more than 2^20 samples per pixel");
    P.rounds = (int)((smax + P.group - 1) / P.group);
    P.total_rays = (unsigned long long)P.n_warps * (unsigned long long)P.rounds * 32ull;
    if ((P.total_rays >> 5) > 0xffffffffull) return fail(ctx, "vt_render: region too large for one launch");
    P.acc_wide = smax > 64 ? 1 : 0;
    // lod_edge == 0: every pixel has n = max(1, samples) samples; for a power of two the mean x / n is x * (1 / n) exactly
    {
        const long long n = s.samples < 1 ? 1 : s.samples;
        P.inv_samples = (s.lod_edge == 0.0 && (n & (n - 1)) == 0) ? 1.0 / (double)n : 0.0;
    }
    long long words = P.n_pixels * (P.acc_wide ? 2 : 1);
    if (words > ctx->acc_cap) {
        VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->d_acc); ctx->d_acc = nullptr; ctx->acc_cap = 0;
        VT_CHECK(ctx, cudaMalloc(&ctx->d_acc, (size_t)words * sizeof(unsigned long long)));
        VT_CHECK(ctx, cudaMemsetAsync(ctx->d_acc, 0, (size_t)words * sizeof(unsigned long long), ctx->stream));
        ctx->acc_cap = words;
    }
    P.acc = ctx->d_acc; P.queue = ctx->d_queue; P.refill_min = ctx->refill_min;
    if (ctx->pow_tab_falloff != s.falloff) {                 // the table holds pow(x, 1 + falloff) for one falloff
        VT_CHECK(ctx, cudaMemsetAsync(ctx->d_pow_tab, 0, VT_POW_SLOTS * 2 * sizeof(unsigned long long), ctx->stream));
        ctx->pow_tab_falloff = s.falloff;
    }
    P.pow_tab = ctx->pow_tab_on ? ctx->d_pow_tab : nullptr;
    P.brick_runs = (ctx->brick_runs && s.chunk_size % VT_BRICK == 0) ? 1 : 0;
    for (int r = 2; r <= VT_LOD_TAB; r++) {
        P.lod_magic[r] = (uint32_t)((1ull << 32) / (unsigned)r) + 1u;
        for (int a = 0; a < 3; a++) { int m = ctx->origin[a] % r; P.lod_bias[r][a] = m < 0 ? m + r : m; }
    }
    const bool fast = ctx->cs_shift >= 0 && !P.acc_wide && ctx->max_res <= VT_LOD_TAB;
    const int which = (fast ? 4 : 0) | (rec ? 2 : 0) | (count ? 1 : 0);
    persistent_fn fn = persistent_kernel(which);
    if (ctx->persist_blocks[which] == 0) {
        int nb = 0;
        VT_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, 128, 0));
        if (ctx->debug) fprintf(stderr, "voxtrace: persistent kernel %d: %d blocks/SM\n", which, nb);
        ctx->persist_blocks[which] = nb > 0 ? nb : 1;
    }
    long long want = (long long)((P.total_rays + 127) / 128);             // one ray slot per thread at least once
    long long resident = (long long)ctx->sm_count * ctx->persist_blocks[which];
    int grid = (int)(want < resident ? want : resident);
    if (grid < 1) grid = 1;
    P.total_warps = (unsigned)grid * 4u;
    fn<<<grid, 128, 0, ctx->stream>>>(P, ctx->d_row_y);
#if VT_CHECKED
    {
        unsigned code = 0;
        VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        VT_CHECK(ctx, cudaMemcpyFromSymbol(&code, vt_oob_code, sizeof code));
        if (code) {
            unsigned zero = 0;
            cudaMemcpyToSymbol(vt_oob_code, &zero, sizeof zero);
            char msg[96];
            snprintf(msg, sizeof msg, "vt_render: bounds check failed in the render kernel, site %u", code);
            return fail(ctx, msg);
        }
    }
#endif
    return 0;
}

struct TraceArgs { const double *dirs; const double *sincos; double *records; long long n; };

static int render_impl(vt_ctx *ctx, const vt_region *region, uint8_t *dev_rgba8, float *dev_mean, double *dev_ray0,
                       uint32_t *dev_visited, uint64_t *dev_counters, uint64_t frame_seed, const TraceArgs *trace) {
    if (!ctx || !region) return 1;
    if (!ctx->have_settings) return fail(ctx, "vt_render: no settings (vt_set_settings)");
    if (!ctx->have_world) return fail(ctx, "vt_render: no world (vt_set_world)");
    const vt_settings &s = ctx->s;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    DevParams P;
    memset(&P, 0, sizeof P);
    P.s = s;
    P.cs_shift = ctx->cs_shift;
    if (!s.static_noise) { P.seed_lo = (uint32_t)frame_seed; P.seed_hi = (uint32_t)(frame_seed >> 32); }
    DevWorld &w = P.w;
    w.ox = ctx->origin[0]; w.oy = ctx->origin[1]; w.oz = ctx->origin[2];
    w.nx = ctx->dims[0]; w.ny = ctx->dims[1]; w.nz = ctx->dims[2];
    w.nbx = ctx->nb[0]; w.nby = ctx->nb[1]; w.nbz = ctx->nb[2];
    w.ncx = ctx->nc[0]; w.ncy = ctx->nc[1]; w.ncz = ctx->nc[2];
    w.brick_tab = ctx->d_brick_tab; w.pool = ctx->d_pool; w.chunk_res = ctx->d_chunk_res; w.mats = ctx->d_mats;
    w.pool_slots = (unsigned)ctx->pool_used;
    memcpy(P.cam.pos, ctx->cam_pos, sizeof P.cam.pos);
    memcpy(P.cam.rot, ctx->cam_rot, sizeof P.cam.rot);
    P.cam.lens = ctx->cam_lens;
    if (int rc = update_dir_tables(ctx)) return rc;
    P.cam.dir_x = ctx->d_dir_x; P.cam.dir_y = ctx->d_dir_y; P.cam.lens_x = ctx->d_lens_x; P.cam.lens_y = ctx->d_lens_y;
    if (s.dof == 0.0 && !trace) {
        if (int rc = update_lens_tables(ctx)) return rc;
        P.cam.col_sc = ctx->d_col_sc; P.cam.row_sc = ctx->d_row_sc;
    }
    if (trace) {
        // ray i draws from the Philox stream keyed i = y * width + x with x, y < 65536
        if (trace->n < 0 || trace->n > 65535LL * s.width) return fail(ctx, "vt_trace_rays: too many rays for one call");
        P.dirs = trace->dirs; P.dir_sc = trace->sincos; P.trace_out = trace->records; P.n_pixels = trace->n;
    } else if (region->dev_pixels) {
        if (region->n_pixels < 0) return fail(ctx, "vt_render: negative n_pixels");
        P.pixels = region->dev_pixels; P.n_pixels = region->n_pixels;
    } else {
        if (region->x0 < 0 || region->y0 < 0 || region->x1 > s.width || region->y1 > s.height || region->x1 < region->x0 ||
            region->y1 < region->y0)
            return fail(ctx, "vt_render: rectangle outside the canvas");
        if (region->stripe_rows < 1 || region->stripe_count < 1 || region->stripe_index < 0 || region->stripe_index >= region->stripe_count)
            return fail(ctx, "vt_render: bad stripe spec");
        int band_rows = 0;
        if (int rc = update_row_map(ctx, region, &band_rows)) return rc;
        P.x0 = region->x0; P.y0 = region->y0; P.x1 = region->x1; P.y1 = region->y1;
        P.stripe_rows = region->stripe_rows; P.stripe_count = region->stripe_count; P.stripe_index = region->stripe_index;
        P.band_rows = band_rows;
        P.n_pixels = (long long)band_rows * (region->x1 - region->x0);
    }
    if (ctx->frame_target && !trace && region->dev_pixels)
        return fail(ctx, "vt_render: a frame target needs a rectangle region");
    if (P.n_pixels == 0) {
        // nothing to render (a rank whose stripe set is empty, e.g. a 48-row canvas split eight ways): the arrival flag
        // must still go up, the consumer waits for one flag per rank
        if (ctx->frame_target && ctx->arrive_flag && !trace) {
            FlagTargets tg;
            memset(&tg, 0, sizeof tg);
            tg.p[0] = ctx->arrive_flag;
            signal_flags_kernel<<<1, 32, 0, ctx->stream>>>(tg, 1, ctx->arrive_value);
            ctx->launches++;
            VT_CHECK(ctx, cudaGetLastError());
        }
        return 0;
    }
    if (!trace) {
        P.frame_target = ctx->frame_target; P.arrive_flag = ctx->arrive_flag; P.arrive_value = ctx->arrive_value;
        if (ctx->frame_target && ctx->visited_peer && dev_visited) {
            const long long words = ((long long)ctx->nc[0] * ctx->nc[1] * ctx->nc[2] + 31) / 32;
            if (words != ctx->visited_words) return fail(ctx, "vt_render: the peer visited mask does not match the world's chunk grid");
            P.visited_peer = ctx->visited_peer; P.visited_words = ctx->visited_words;
        }
    }
    P.rgba8 = dev_rgba8; P.mean = dev_mean; P.ray0 = dev_ray0; P.visited = dev_visited;
    P.counters = reinterpret_cast<unsigned long long *>(dev_counters);
    // A packet of 32 ray slots is one sample index of an 8 x 4 pixel tile; the other samples of those pixels are
    // further packets ("rounds").  Neighbouring pixels at one sample index are as coherent as the samples of one pixel,
    // and a pixel's accumulator is then touched by different warps at different times (VT_GROUP=n: up to n lanes per
    // pixel instead; measured 1.7 % slower at 4 spp, equal at 8 and 16 spp).
    int G = 1, gs = 0;
    const int gmax = trace ? 1 : ctx->group_max;
    while (G < s.samples && G < gmax) { G <<= 1; gs++; }
    const int ppw = 32 / G;
    static const int TW[6] = {8, 4, 4, 2, 2, 1};      // tile width for 32,16,8,4,2,1 pixels per warp
    int tw = TW[gs], tws = 0;
    while ((1 << tws) < tw) tws++;
    P.group = G; P.group_shift = gs; P.tile_w = tw; P.tile_w_shift = tws; P.tile_h = ppw / tw;
    if (P.pixels || trace) P.n_warps = (P.n_pixels + ppw - 1) / ppw;
    else {
        P.tiles_x = (P.x1 - P.x0 + tw - 1) / tw;
        P.n_warps = (long long)P.tiles_x * ((P.band_rows + P.tile_h - 1) / P.tile_h);
    }
    if (int rc = launch_persistent(ctx, P, dev_ray0 != nullptr || trace, dev_counters != nullptr)) return rc;
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    return 0;
}

extern "C" int vt_render(vt_ctx *ctx, const vt_region *region, uint8_t *dev_rgba8, float *dev_mean, double *dev_ray0,
                         uint32_t *dev_visited, uint64_t *dev_counters, uint64_t frame_seed) {
    return render_impl(ctx, region, dev_rgba8, dev_mean, dev_ray0, dev_visited, dev_counters, frame_seed, nullptr);
}

// Camera.trace (init.py:37-121) for a batch of caller-given rays.
extern "C" int vt_trace_rays(vt_ctx *ctx, const double *dev_dirs, const double *dev_sincos, int64_t n, double *dev_records,
                             uint32_t *dev_visited, uint64_t *dev_counters, uint64_t frame_seed) {
    if (!ctx || n < 0 || (n > 0 && (!dev_dirs || !dev_records))) return 1;
    if (n == 0) return 0;
    vt_region r;
    memset(&r, 0, sizeof r);
    r.stripe_rows = 1; r.stripe_count = 1;
    TraceArgs t = {dev_dirs, dev_sincos, dev_records, (long long)n};
    return render_impl(ctx, &r, nullptr, nullptr, nullptr, dev_visited, dev_counters, frame_seed, &t);
}

// ---- self-test hooks: the correctly rounded power (vt_pow_cr.h) on the host and on the device -------------------
extern "C" double vt_pow_host(double x, double y) { return vtp::pow_cr(x, y); }
__global__ void pow_selftest_kernel(const double *x, const double *y, long long n, double *out) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < n) out[i] = vt::pow_out(x[i], y[i]);
}
extern "C" int vt_selftest_pow(vt_ctx *ctx, const double *dev_x, const double *dev_y, int64_t n, double *dev_out) {
    if (!ctx || n < 0 || (n > 0 && (!dev_x || !dev_y || !dev_out))) return 1;
    if (n == 0) return 0;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    pow_selftest_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(dev_x, dev_y, n, dev_out);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    return 0;
}

// ---- physics slice test (vt_physics.cuh) ---------------------------------------------------------------------
#include "vt_physics.cuh"

extern "C" int vt_set_phys_materials(vt_ctx *ctx, const vt_phys_material *hm, int n) {
    if (!ctx || !hm) return 1;
    if (n < 0 || n > VT_MAX_MATERIALS) return fail(ctx, "vt_set_phys_materials: at most 255 materials (id 0 is empty)");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    if (!ctx->d_phys) VT_CHECK(ctx, cudaMalloc(&ctx->d_phys, 256 * sizeof(vt_phys_material)));
    return staged_upload(ctx, ctx->d_phys, hm, (size_t)(n + 1) * sizeof(vt_phys_material));
}

extern "C" int vt_slice_collide(vt_ctx *ctx, const vt_phys_object *self, const vt_phys_object *other, const int32_t lo[3], const int32_t hi[3],
                                const int32_t vel_dir[3], double friction_scale, uint64_t seed, double inout_sums[2], int32_t *out_blocked) {
    if (!ctx || !self || !other || !lo || !hi || !vel_dir || !inout_sums || !out_blocked) return 1;
    if (!ctx->d_phys) return fail(ctx, "vt_slice_collide: no physical material table (vt_set_phys_materials)");
    vtphys::SliceArgs a;
    memset(&a, 0, sizeof a);
    a.self = *self; a.other = *other;
    long long n_cells = 1;
    for (int k = 0; k < 3; k++) {
        if (hi[k] < lo[k]) { *out_blocked = 0; return 0; }
        a.lo[k] = lo[k]; a.n[k] = hi[k] - lo[k] + 1; a.dir[k] = vel_dir[k];
        n_cells *= a.n[k];
    }
    if (n_cells > (1LL << 31)) return fail(ctx, "vt_slice_collide: slice too large");
    const vt_phys_object *both[2] = {self, other};
    for (int k = 0; k < 2; k++) {
        if (!both[k]->dev_grid) return fail(ctx, "vt_slice_collide: object without a voxel grid");
        for (int a2 = 0; a2 < 3; a2++) {
            if (both[k]->size[a2] <= 0 || both[k]->grid_dims[a2] <= 0) return fail(ctx, "vt_slice_collide: sprite size / grid dims must be positive");
            if (both[k]->angle[a2] < 0 || both[k]->angle[a2] > 3) return fail(ctx, "vt_slice_collide: angle must be 0..3");
        }
    }
    a.scale = friction_scale; a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    void *p1 = nullptr, *p2 = nullptr;
    if (int rc = scene_scratch(ctx, 1, (size_t)n_cells + 64, &p1)) return rc;
    if (int rc = scene_scratch(ctx, 2, (size_t)n_cells * sizeof(double2) + 64, &p2)) return rc;
    if (!ctx->h_word) VT_CHECK(ctx, cudaMallocHost(&ctx->h_word, 64));
    if (!ctx->d_phys_out) VT_CHECK(ctx, cudaMalloc(&ctx->d_phys_out, 4 * sizeof(double)));
    double *h = reinterpret_cast<double *>(ctx->h_word);
    h[0] = inout_sums[0]; h[1] = inout_sums[1]; h[2] = 0.0;
    VT_CHECK(ctx, cudaMemcpyAsync(ctx->d_phys_out, h, 3 * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    const int sms = ctx->sm_count > 0 ? ctx->sm_count : 148;
    long long want = (n_cells + 255) / 256;
    const int grid = (int)(want < (long long)sms * 8 ? want : (long long)sms * 8);
    vtphys::slice_cells_kernel<<<grid, 256, 0, ctx->stream>>>(a, reinterpret_cast<const double *>(ctx->d_phys), n_cells, static_cast<uint8_t *>(p1), static_cast<double2 *>(p2));
    vtphys::slice_sum_kernel<<<1, 1024, 0, ctx->stream>>>(static_cast<const uint8_t *>(p1), static_cast<double2 *>(p2), n_cells, ctx->d_phys_out);
    ctx->launches += 2;
    VT_CHECK(ctx, cudaGetLastError());
    VT_CHECK(ctx, cudaMemcpyAsync(h, ctx->d_phys_out, 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    inout_sums[0] = h[0]; inout_sums[1] = h[1];
    *out_blocked = h[2] != 0.0 ? 1 : 0;
    return 0;
}

// ---- allocation-free variant of the slice test ---------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  vt_slice_collide takes its per-cell scratch (one
// flag byte + two fp64 products per cell) from grow-only cudaMalloc'ed buffers; here the scratch is this file-scope array,
// VT_STATIC_SLICE_MAX_CELLS cells = 1 114 112 bytes held by every GPU that loads the library until the process exits.
// In host code `vt_static_slice_scratch` is a host-side stand-in, not a device address: the address is looked up once per
// device with cudaGetSymbolAddress (that device current) and that pointer goes to the kernels.
static __device__ __align__(256) unsigned char vt_static_slice_scratch[(size_t)VT_STATIC_SLICE_MAX_CELLS * (sizeof(double2) + 1)];
static unsigned char *g_static_scratch_addr[64];            // per CUDA device ordinal, resolved on first use

extern "C" int vt_slice_collide_static(vt_ctx *ctx, const vt_phys_object *self, const vt_phys_object *other, const int32_t lo[3],
                                       const int32_t hi[3], const int32_t vel_dir[3], double friction_scale, uint64_t seed,
                                       const vt_phys_material *dev_materials, double *dev_inout) {
    if (!ctx || !self || !other || !lo || !hi || !vel_dir || !dev_materials || !dev_inout) return 1;
    if (ctx->device < 0 || ctx->device >= 64) return fail(ctx, "vt_slice_collide_static: device ordinal out of range");
    vtphys::SliceArgs a;
    memset(&a, 0, sizeof a);
    a.self = *self; a.other = *other;
    long long n_cells = 1;
    for (int k = 0; k < 3; k++) {
        if (hi[k] < lo[k]) n_cells = 0;
        a.lo[k] = lo[k]; a.n[k] = hi[k] < lo[k] ? 0 : hi[k] - lo[k] + 1; a.dir[k] = vel_dir[k];
        n_cells *= a.n[k];
    }
    if (n_cells > VT_STATIC_SLICE_MAX_CELLS)
        return fail(ctx, "vt_slice_collide_static: slice larger than VT_STATIC_SLICE_MAX_CELLS (65536) cells; use vt_slice_collide");
    const vt_phys_object *both[2] = {self, other};
    for (int k = 0; k < 2; k++) {
        if (!both[k]->dev_grid) return fail(ctx, "vt_slice_collide_static: object without a voxel grid");
        for (int a2 = 0; a2 < 3; a2++) {
            if (both[k]->size[a2] <= 0 || both[k]->grid_dims[a2] <= 0) return fail(ctx, "vt_slice_collide_static: sprite size / grid dims must be positive");
            if (both[k]->angle[a2] < 0 || both[k]->angle[a2] > 3) return fail(ctx, "vt_slice_collide_static: angle must be 0..3");
        }
    }
    a.scale = friction_scale; a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    if (!g_static_scratch_addr[ctx->device]) {               // once per device, with that device current
        void *ptr = nullptr;
        VT_CHECK(ctx, cudaGetSymbolAddress(&ptr, vt_static_slice_scratch));
        if (!ptr) return fail(ctx, "vt_slice_collide_static: cudaGetSymbolAddress returned a null address");
        g_static_scratch_addr[ctx->device] = static_cast<unsigned char *>(ptr);
    }
    unsigned char *base = g_static_scratch_addr[ctx->device];
    double2 *contrib = reinterpret_cast<double2 *>(base);                                     // [MAX] 16-byte aligned (base is 256-aligned)
    uint8_t *flag = base + (size_t)VT_STATIC_SLICE_MAX_CELLS * sizeof(double2);               // [MAX]
    // Nothing in the scratch has to start at zero: slice_cells_kernel writes flag[i] and contrib[i] for every i < n_cells
    // before slice_sum_kernel reads them, so the previous call's contents are never seen.
    const int sms = ctx->sm_count > 0 ? ctx->sm_count : 148;
    if (n_cells > 0) {
        long long want = (n_cells + 255) / 256;
        const int grid = (int)(want < (long long)sms * 8 ? want : (long long)sms * 8);
        vtphys::slice_cells_kernel<<<grid, 256, 0, ctx->stream>>>(a, reinterpret_cast<const double *>(dev_materials), n_cells, flag, contrib);
        ctx->launches++;
    }
    vtphys::slice_sum_kernel<<<1, 1024, 0, ctx->stream>>>(flag, contrib, n_cells, dev_inout);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    return 0;
}

// ---- post-FX (vt_postfx.cuh) ------------------------------------------------------------------------------
#include "vt_postfx.cuh"

static int fx_grid(long long n) { long long g = (n + 255) / 256; return (int)(g < 148LL * 16 ? (g > 0 ? g : 1) : 148LL * 16); }

static int fx_buffer(vt_ctx *ctx, int k, long long bytes, uchar4 **out) {
    if (bytes > ctx->fx_cap[k]) {
        VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->fx_buf[k]); ctx->fx_buf[k] = nullptr; ctx->fx_cap[k] = 0;
        VT_CHECK(ctx, cudaMalloc(&ctx->fx_buf[k], (size_t)bytes));
        ctx->fx_cap[k] = bytes;
    }
    *out = reinterpret_cast<uchar4 *>(ctx->fx_buf[k]);
    return 0;
}

extern "C" int vt_fx_compose(vt_ctx *ctx, uint8_t *dev_canvas, const uint8_t *dev_tile, int64_t n) {
    if (!ctx || !dev_canvas || !dev_tile || n < 0) return 1;
    if (n == 0) return 0;
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    vtfx::compose_kernel<<<fx_grid(n), 256, 0, ctx->stream>>>(reinterpret_cast<uchar4 *>(dev_canvas), reinterpret_cast<const uchar4 *>(dev_tile), n);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    return 0;
}

// one pygame scaling call: nearest (transform.scale), or "smooth" = area average when shrinking on both axes, bilinear otherwise
static int fx_scale(vt_ctx *ctx, const uchar4 *src, int sw, int sh, uchar4 *dst, int dw, int dh, bool smooth) {
    const long long n = (long long)dw * dh;
    if (smooth && dw < sw && dh < sh) vtfx::shrink_kernel<<<fx_grid(n), 256, 0, ctx->stream>>>(src, sw, sh, dst, dw, dh);
    else vtfx::resample_kernel<<<fx_grid(n), 256, 0, ctx->stream>>>(src, sw, sh, dst, dw, dh, smooth ? 1 : 0, nullptr);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    return 0;
}

extern "C" int vt_fx_draw(vt_ctx *ctx, const vt_fx *fx, const uint8_t *dev_canvas, uint8_t *dev_out, int32_t host_avg[4]) {
    if (!ctx || !fx || !dev_canvas || !dev_out) return 1;
    const int W = fx->width, H = fx->height, OW = fx->out_width, OH = fx->out_height;
    if (W <= 0 || H <= 0 || OW <= 0 || OH <= 0 || W > 65535 || H > 65535 || OW > 65535 || OH > 65535)
        return fail(ctx, "vt_fx_draw: canvas / window size out of range");
    if (fx->smooth < 0 || fx->subsamples < 0 || fx->subsamples > 15) return fail(ctx, "vt_fx_draw: smooth / subsamples out of range");
    VT_CHECK(ctx, cudaSetDevice(ctx->device));
    const long long n = (long long)W * H;
    if (!ctx->d_fx_sums) {
        VT_CHECK(ctx, cudaMalloc(&ctx->d_fx_sums, 4 * sizeof(unsigned long long)));
        VT_CHECK(ctx, cudaMalloc(&ctx->d_fx_avg, 4 * sizeof(int)));
    }
    const uchar4 *canvas = reinterpret_cast<const uchar4 *>(dev_canvas);
    VT_CHECK(ctx, cudaMemsetAsync(ctx->d_fx_sums, 0, 4 * sizeof(unsigned long long), ctx->stream));
    vtfx::average_kernel<<<fx_grid(n), 256, 0, ctx->stream>>>(canvas, n, ctx->d_fx_sums);
    ctx->launches++;
    vtfx::FxParams p;
    p.spill_fac = fx->spill != 0.0 ? 255 - (int)nearbyint(fx->spill * 255.0) : -1;
    p.iris_on = fx->iris_on ? 1 : 0;
    p.iris_add = fx->iris > 0.0 ? 1 : 0;
    p.iris_fac = (int)nearbyint(fabs(fx->iris * 255.0));
    const bool bloom = fx->bloom != 0.0 && fx->bloom_blur != 0.0;
    p.bloom_fac = bloom ? (int)nearbyint((1.0 - fx->bloom) * 255.0) : -1;
    uchar4 *a = nullptr, *dark = nullptr, *small = nullptr, *big = nullptr;
    if (int rc = fx_buffer(ctx, 0, n * 4, &a)) return rc;
    if (bloom) if (int rc = fx_buffer(ctx, 1, n * 4, &dark)) return rc;
    vtfx::spill_iris_kernel<<<fx_grid(n), 256, 0, ctx->stream>>>(canvas, n, ctx->d_fx_sums, p, a, dark, ctx->d_fx_avg);
    ctx->launches++;
    VT_CHECK(ctx, cudaGetLastError());
    if (bloom) {                                                                  // init.py:232-239
        double div = fx->bloom_blur > 1.0 ? fx->bloom_blur : 1.0;
        int bw = (int)nearbyint((double)W / div), bh = (int)nearbyint((double)H / div);
        if (bw < 1) bw = 1;
        if (bh < 1) bh = 1;
        if (int rc = fx_buffer(ctx, 2, (long long)bw * bh * 4, &small)) return rc;
        if (int rc = fx_scale(ctx, dark, W, H, small, bw, bh, true)) return rc;
        vtfx::resample_kernel<<<fx_grid(n), 256, 0, ctx->stream>>>(small, bw, bh, a, W, H, 1, a);      // smoothscale up, blit ADD
        ctx->launches++;
        VT_CHECK(ctx, cudaGetLastError());
    }
    const uchar4 *cur = a;
    int cw = W, ch = H;
    uchar4 *out = reinterpret_cast<uchar4 *>(dev_out);
    if (fx->subsamples != 0.0) {                                                  // init.py:242-244
        const double f = 1.0 + fx->subsamples;
        int nw = (int)((double)W * f), nh = (int)((double)H * f);
        if (int rc = fx_buffer(ctx, 1, (long long)nw * nh * 4, &dark)) return rc;
        if (int rc = fx_scale(ctx, cur, cw, ch, dark, nw, nh, true)) return rc;
        cur = dark; cw = nw; ch = nh;
    }
    if (fx->smooth == 0.0) {                                                      // init.py:247-254
        if (int rc = fx_scale(ctx, cur, cw, ch, out, OW, OH, false)) return rc;
    } else if (fx->smooth == 1.0) {
        if (int rc = fx_scale(ctx, cur, cw, ch, out, OW, OH, true)) return rc;
    } else {
        int f = (int)(1.0 / fx->smooth);
        if (f < 1) f = 1;
        const long long nw = (long long)cw * f, nh = (long long)ch * f;
        if (nw > 65535 || nh > 65535) return fail(ctx, "vt_fx_draw: sharp pre-scale too large");
        if (int rc = fx_buffer(ctx, 3, nw * nh * 4, &big)) return rc;
        if (int rc = fx_scale(ctx, cur, cw, ch, big, (int)nw, (int)nh, false)) return rc;
        if (int rc = fx_scale(ctx, big, (int)nw, (int)nh, out, OW, OH, true)) return rc;
    }
    int avg[4] = {0, 0, 0, 0};
    VT_CHECK(ctx, cudaMemcpyAsync(avg, ctx->d_fx_avg, sizeof avg, cudaMemcpyDeviceToHost, ctx->stream));
    VT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    if (host_avg) for (int k = 0; k < 4; k++) host_avg[k] = avg[k];
    return 0;
}
