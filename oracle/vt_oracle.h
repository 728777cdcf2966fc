/* vt_oracle.h — CPU restatement of the reference trace loop.  TEST INFRASTRUCTURE ONLY.
 *
 * This is the oracle of the voxtrace-b200 build: a plain-C, fp64 restatement of
 * python_raytracer's Camera.tile / Camera.trace / Frame.get_voxel / lib.material /
 * lib.material_background (reference init.py:37-150, data.py:136-145, lib.py:448-476).
 * It is NOT part of the product: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.
 *
 * Parity is PINNED: tests/golden/ holds outputs of the reference's own Python code run in the build
 * container (oracle/gen_golden.py); tests/test_oracle_golden.py checks this file against them bit for bit.
 */
#ifndef VT_ORACLE_H
#define VT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* data.settings subset read by the path (reference data.py:18-50, 64-68) */
typedef struct {
    int32_t width, height, samples, static_noise;
    int32_t chunk_size, chunk_radius;
    int32_t background;          /* 0 = data.background is None, 1 = lib.material_background */
    int32_t pad_;
    double shutter, falloff, dof, dist_min, dist_max, max_light, max_bounces;
    double lod_bounces, lod_samples, lod_random, lod_edge, proportions;
} vto_settings;

typedef struct {
    double pos[3];
    double rot[4];               /* quaternion x, y, z, w (reference lib.py:341-348) */
    double lens;                 /* Camera.lens = fov*pi/8 (init.py:17) */
} vto_camera;

/* World = union of the camera's chunk frames, keyed by world cell (init.py:417-424), stored dense. */
typedef struct {
    int32_t origin[3];           /* world cell of grid[0][0][0]; multiple of chunk_size */
    int32_t dims[3];             /* cells per axis; multiple of chunk_size */
    const uint8_t *cells;        /* [dims0][dims1][dims2], z fastest; 0 = empty else material id */
    const uint8_t *chunk_res;    /* [dims0/cs][dims1/cs][dims2/cs]; 0 = chunk absent else Frame.resolution */
    const double *materials;     /* [n_materials+1][7]: albedo r,g,b, roughness, absorption, ior, energy */
    int32_t n_materials;
    int32_t pad_;
} vto_scene;

enum { VTO_RNG_MT19937 = 0, VTO_RNG_PHILOX = 1 };

/* Per-pixel outputs, all aligned with the pixel list:
 *   mean[n][4]   per-pixel RGBA mean handed to Surface.set_at (init.py:145-146)
 *   rgba8[n][4]  truncating quantisation of mean
 *   nsamp[n]     samples traced
 *   ray0[n][15]  sample-0 ray: r,g,b,energy,bounces,step,hits,fx,fy,fz,fflag,first_step,vx,vy,vz
 *   visited[nchunks] 1 where a ray re-evaluated into that chunk slot (ray.traversed, init.py:72-73)
 *   counters[8]  rays, in-chunk samples, void samples, hits, neighbour probes, chunk re-evaluations, 0, 0
 * Any output pointer may be NULL.  Returns 0 on success. */
int vto_render(const vto_settings *s, const vto_camera *cam, const vto_scene *scene,
               const int32_t *pixels, int64_t n_pixels, int rng_mode, uint64_t seed, int threads,
               double *mean, uint8_t *rgba8, int32_t *nsamp, double *ray0,
               uint8_t *visited, uint64_t *counters);

/* Camera.trace (init.py:37-121) for n caller-given rays: dirs [n][3] = dir_x, dir_y, detail.  records [n][20]: r, g, b,
 * energy, bounces, step, hits, first-hit x, y, z, flag, first-hit step, vel x, y, z, life, pos x, y, z, 0.  Philox mode:
 * ray i draws from the stream keyed i; MT19937 mode: one generator seeded with `seed` runs on across the rays. */
int vto_trace_rays(const vto_settings *s, const vto_camera *cam, const vto_scene *scene, const double *dirs, int64_t n,
                   int rng_mode, uint64_t seed, double *records, uint8_t *visited, uint64_t *counters);

/* Known-answer hooks for the primitives (tests compare them with the imported lib.py / CPython). */
double vto_py_floordiv(double x, double w);
double vto_py_mod(double x, double w);
double vto_mt_random(uint64_t seed, int skip);      /* random.seed(seed); skip draws; random.random() */
void vto_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]);
void vto_ray_dir(const vto_camera *cam, double lens_x, double lens_y, double out[3]);

#ifdef __cplusplus
}
#endif
#endif
