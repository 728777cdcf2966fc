/* vt_oracle.c — CPU restatement of the reference trace loop.  TEST INFRASTRUCTURE ONLY (see vt_oracle.h).
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference).
 * All ray state is fp64 like Python floats; every a*b+c is two roundings (build with
 * -ffp-contract=off); Python round() == rint() in the default rounding mode (ties-to-even).
 * Build: see oracle/Makefile.
 */
#include "vt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <stdatomic.h>

/* ---------------------------------------------------------------- Python float // and % ---------- */
/* CPython floatobject.c float_divmod semantics, used by vec3.snapped / __floordiv__ (lib.py:204-208,
 * 316-320) and by the void step's float modulo (init.py:114). */
static void py_divmod(double vx, double wx, double *fdiv, double *fmodv) {
    double mod = fmod(vx, wx);
    double div = (vx - mod) / wx;
    if (mod) {
        if ((wx < 0) != (mod < 0)) { mod += wx; div -= 1.0; }
    } else {
        mod = copysign(0.0, wx);
    }
    double floordiv;
    if (div) {
        floordiv = floor(div);
        if (div - floordiv > 0.5) floordiv += 1.0;
    } else {
        floordiv = copysign(0.0, vx / wx);
    }
    *fdiv = floordiv; *fmodv = mod;
}
double vto_py_floordiv(double x, double w) { double d, m; py_divmod(x, w, &d, &m); return d; }
double vto_py_mod(double x, double w) { double d, m; py_divmod(x, w, &d, &m); return m; }

/* Python int floor division (data.py:137 `pos // self.resolution` on int vec3) */
static inline int64_t ifloordiv(int64_t a, int64_t b) {
    int64_t q = a / b, r = a % b;
    return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q;
}

/* ---------------------------------------------------------------- RNG ---------------------------- */
/* MT19937 exactly as CPython's _randommodule.c: random.seed(int) -> init_by_array(32-bit limbs of |a|),
 * random.random() -> genrand_res53.  Used to pin the stochastic paths against the reference when
 * settings.static is on (init.py:136-137). */
typedef struct { uint32_t mt[624]; int idx; } mt_t;
static void mt_init_genrand(mt_t *r, uint32_t s) {
    r->mt[0] = s;
    for (int i = 1; i < 624; i++) r->mt[i] = 1812433253U * (r->mt[i - 1] ^ (r->mt[i - 1] >> 30)) + (uint32_t)i;
    r->idx = 624;
}
static void mt_init_by_array(mt_t *r, const uint32_t *key, int len) {
    mt_init_genrand(r, 19650218U);
    int i = 1, j = 0, k = (624 > len ? 624 : len);
    for (; k; k--) {
        r->mt[i] = (r->mt[i] ^ ((r->mt[i - 1] ^ (r->mt[i - 1] >> 30)) * 1664525U)) + key[j] + (uint32_t)j;
        i++; j++;
        if (i >= 624) { r->mt[0] = r->mt[623]; i = 1; }
        if (j >= len) j = 0;
    }
    for (k = 623; k; k--) {
        r->mt[i] = (r->mt[i] ^ ((r->mt[i - 1] ^ (r->mt[i - 1] >> 30)) * 1566083941U)) - (uint32_t)i;
        i++;
        if (i >= 624) { r->mt[0] = r->mt[623]; i = 1; }
    }
    r->mt[0] = 0x80000000U;
}
static void mt_seed_u64(mt_t *r, uint64_t a) {
    uint32_t key[2] = { (uint32_t)a, (uint32_t)(a >> 32) };
    mt_init_by_array(r, key, key[1] ? 2 : 1);
}
static uint32_t mt_u32(mt_t *r) {
    static const uint32_t mag01[2] = { 0U, 0x9908b0dfU };
    if (r->idx >= 624) {
        int kk; uint32_t y; uint32_t *mt = r->mt;
        for (kk = 0; kk < 624 - 397; kk++) { y = (mt[kk] & 0x80000000U) | (mt[kk + 1] & 0x7fffffffU); mt[kk] = mt[kk + 397] ^ (y >> 1) ^ mag01[y & 1U]; }
        for (; kk < 623; kk++) { y = (mt[kk] & 0x80000000U) | (mt[kk + 1] & 0x7fffffffU); mt[kk] = mt[kk + (397 - 624)] ^ (y >> 1) ^ mag01[y & 1U]; }
        y = (mt[623] & 0x80000000U) | (mt[0] & 0x7fffffffU); mt[623] = mt[396] ^ (y >> 1) ^ mag01[y & 1U];
        r->idx = 0;
    }
    uint32_t y = r->mt[r->idx++];
    y ^= (y >> 11); y ^= (y << 7) & 0x9d2c5680U; y ^= (y << 15) & 0xefc60000U; y ^= (y >> 18);
    return y;
}
static inline double u53(uint32_t hi, uint32_t lo) {
    uint32_t a = hi >> 5, b = lo >> 6;
    return (a * 67108864.0 + b) * (1.0 / 9007199254740992.0);
}

/* Philox4x32-10 (Salmon et al. 2011), the GPU path's counter-based generator (north_star). */
void vto_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53U * c0, p1 = (uint64_t)0xCD9E8D57U * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9U; k1 += 0xBB67AE85U;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

typedef struct {
    int mode;
    mt_t mt;
    uint32_t key0, key1, sample, seed_lo, seed_hi, draw;   /* philox */
    uint32_t w[4];
} rng_t;

/* one uniform in [0,1) with 53 random bits, like random.random() */
static double rng_next(rng_t *g) {
    if (g->mode == VTO_RNG_MT19937) {
        uint32_t a = mt_u32(&g->mt), b = mt_u32(&g->mt);
        return u53(a, b);
    }
    uint32_t d = g->draw++;
    if ((d & 1U) == 0) {
        vto_philox(d >> 1, g->sample, g->seed_lo, g->seed_hi, g->key0, g->key1, g->w);
        return u53(g->w[0], g->w[1]);
    }
    return u53(g->w[2], g->w[3]);
}
/* lib.py:431-434 rand(amp): returns 0 WITHOUT drawing when amp is falsy */
static inline double rng_rand(rng_t *g, double amp) {
    if (amp == 0.0) return 0.0;
    return (-1 + rng_next(g) * 2) * amp;
}
double vto_mt_random(uint64_t seed, int skip) {
    rng_t g; g.mode = VTO_RNG_MT19937; mt_seed_u64(&g.mt, seed);
    for (int i = 0; i < skip; i++) rng_next(&g);
    return rng_next(&g);
}

/* ---------------------------------------------------------------- ray direction ------------------ */
/* init.py:43-45: vec3(0, -lens_x, +lens_y).quaternion() (lib.py:322-338, degrees), cam.rot.multiply(.)
 * (lib.py:353-358, the reference's own product), .vec_forward() (lib.py:372-376).
 * `q ** 2` in vec_forward is CPython float_pow -> libm pow(q, 2.0): kept literally. */
void vto_ray_dir(const vto_camera *cam, double lens_x, double lens_y, double out[3]) {
    const double deg2rad = 3.14159265358979323846 / 180.0;   /* math.radians: x * (pi/180) */
    double ex = 0.0, ey = -lens_x, ez = +lens_y;
    double rad_x = ex * deg2rad, rad_y = ey * deg2rad, rad_z = ez * deg2rad;
    double sin_x = sin(rad_x / 2), cos_x = cos(rad_x / 2);
    double sin_y = sin(rad_y / 2), cos_y = cos(rad_y / 2);
    double sin_z = sin(rad_z / 2), cos_z = cos(rad_z / 2);
    double ox = sin_x * cos_y * cos_z - cos_x * sin_y * sin_z;
    double oy = cos_x * sin_y * cos_z - sin_x * cos_y * sin_z;
    double oz = cos_x * cos_y * sin_z + sin_x * sin_y * cos_z;
    double ow = cos_x * cos_y * cos_z + sin_x * sin_y * sin_z;
    double sx = cam->rot[0], sy = cam->rot[1], sz = cam->rot[2], sw = cam->rot[3];
    double x = sw * ox + sz * oy - sy * oz + sx * ow;
    double y = sz * ox + sw * oy + sx * oz + sy * ow;
    double z = sy * ox - sx * oy + sw * oz + sz * ow;
    double w = sx * ox - sy * oy - sz * oz + sw * ow;
    out[0] = 2 * (z * x + w * y);
    out[1] = 2 * (y * x - w * z);
    out[2] = 1 - 2 * (pow(z, 2.0) + pow(y, 2.0));
}

/* ---------------------------------------------------------------- world lookups ------------------ */
typedef struct {
    const vto_settings *s;
    const vto_scene *sc;
    int32_t nc[3];               /* chunk slots per axis */
    uint8_t *visited;            /* shared; benign races (bytes set to 1) */
} world_t;

/* `post_chunk in self.chunks` (init.py:71, init.py:28-33): chunk_min is a float triple that is a multiple of
 * chunk_size; returns the chunk-slot index or -1. */
static int64_t chunk_slot(const world_t *w, const double cmin[3]) {
    int64_t idx = 0;
    for (int a = 0; a < 3; a++) {
        double rel = (cmin[a] - (double)w->sc->origin[a]) / (double)w->s->chunk_size;
        if (!(rel >= 0.0) || !(rel < (double)w->nc[a])) return -1;
        idx = idx * w->nc[a] + (int64_t)rel;
    }
    return idx;
}

/* Frame.get_voxel (data.py:136-145) on the LOD-r frame of the chunk whose origin is c[]:
 * key = floor(p) // r; that key exists iff the full-resolution voxel r*key belongs to this chunk's own voxel
 * set (init.py:417-424 fills only cells inside the chunk; data.py:165-169 keeps cells with all coords % r == 0). */
static int get_voxel(const world_t *w, const int64_t c[3], int r, const int64_t p[3]) {
    int64_t idx = 0;
    for (int a = 0; a < 3; a++) {
        int64_t q = (r > 1) ? ifloordiv(p[a], r) * r : p[a];
        if (q < c[a] || q >= c[a] + w->s->chunk_size) return 0;
        idx = idx * w->sc->dims[a] + (q - w->sc->origin[a]);
    }
    return w->sc->cells[idx];
}

typedef struct {
    double cr, cg, cb;           /* rgb ints held in doubles */
    double energy, pos[3], vel[3], step, life, bounces;
    int hits, fflag; int64_t fhit[3]; double first_step;
} ray_t;

static inline double py_round(double x) { return rint(x); }   /* round-half-even, default FE_TONEAREST */
static inline double mixd(double a, double b, double bias1) { double bias2 = 1 - bias1; return a * bias2 + b * bias1; }

/* lib.material (lib.py:448-460).  mat = row of 7 doubles.  Returns the bounce weight. */
static double material(ray_t *ray, const double *mat, const vto_settings *s, rng_t *g) {
    double absorption = fmin(1.0, mat[4] / pow(1 + ray->bounces, 1 + s->falloff));
    double bias2 = 1 - absorption;
    ray->cr = py_round(ray->cr * bias2 + mat[0] * absorption);
    ray->cg = py_round(ray->cg * bias2 + mat[1] * absorption);
    ray->cb = py_round(ray->cb * bias2 + mat[2] * absorption);
    ray->energy = mixd(ray->energy, mat[6], absorption);
    ray->life *= 1 - (mat[3] * absorption);
    double rx = rng_rand(g, mat[3]), ry = rng_rand(g, mat[3]), rz = rng_rand(g, mat[3]);
    ray->vel[0] += rx; ray->vel[1] += ry; ray->vel[2] += rz;
    return mat[4];
}

/* lib.material_background (lib.py:463-476) */
static void material_background(ray_t *ray, const vto_settings *s) {
    double absorption = fmin(1.0, 1 / pow(1 + ray->bounces, 1 + s->falloff));
    double up = fmax(0.0, +ray->vel[1]);
    double sr = 127, sg = 127 + up * 64, sb = 127 + up * 128;
    double energy = 1 + up;
    double bias2 = 1 - absorption;
    ray->cr = py_round(ray->cr * bias2 + sr * absorption);
    ray->cg = py_round(ray->cg * bias2 + sg * absorption);
    ray->cb = py_round(ray->cb * bias2 + sb * absorption);
    ray->energy = mixd(ray->energy, energy, absorption);
    ray->cr = fmin(255.0, py_round(ray->cr * ray->energy));
    ray->cg = fmin(255.0, py_round(ray->cg * ray->energy));
    ray->cb = fmin(255.0, py_round(ray->cb * ray->energy));
}

static inline int inside_incl(const double p[3], const double lo[3], const double hi[3]) {
    /* `pos >= chunk_min and pos <= chunk_max`, all-component compares (lib.py:228-244) */
    return p[0] >= lo[0] && p[1] >= lo[1] && p[2] >= lo[2] && p[0] <= hi[0] && p[1] <= hi[1] && p[2] <= hi[2];
}

/* Camera.trace (init.py:37-121) */
static void trace(const world_t *w, const vto_camera *cam, double dir_x, double dir_y, double detail,
                  rng_t *g, ray_t *ray, uint64_t *cnt) {
    const vto_settings *s = w->s;
    const double cs = (double)s->chunk_size;
    double lens_x = (dir_x / s->proportions) * cam->lens + rng_rand(g, s->dof);
    double lens_y = (dir_y * s->proportions) * cam->lens + rng_rand(g, s->dof);
    double dir[3];
    vto_ray_dir(cam, lens_x, lens_y, dir);
    double cmin[3] = { 0, 0, 0 }, cmax[3] = { 0, 0, 0 };
    int64_t corg[3] = { 0, 0, 0 };
    int res = 0;                                     /* 0 == `chunk is None` */

    memset(ray, 0, sizeof *ray);
    for (int a = 0; a < 3; a++) { ray->pos[a] = cam->pos[a] + dir[a] * s->dist_min; ray->vel[a] = dir[a]; }
    ray->life = (s->dist_max - s->dist_min) * detail;
    ray->first_step = -1;

    while (ray->step < ray->life) {
        if (!inside_incl(ray->pos, cmin, cmax)) {                           /* init.py:67-73 */
            for (int a = 0; a < 3; a++) {
                cmin[a] = vto_py_floordiv(ray->pos[a], cs) * cs;             /* vec3.snapped, lib.py:316-320 */
                cmax[a] = cmin[a] + cs;
            }
            int64_t slot = chunk_slot(w, cmin);
            res = slot >= 0 ? w->sc->chunk_res[slot] : 0;
            if (slot >= 0) {
                if (w->visited) w->visited[slot] = 1;
                for (int a = 0; a < 3; a++) corg[a] = (int64_t)cmin[a];
            }
            cnt[5]++;
        }
        if (res) {
            int64_t p[3] = { (int64_t)floor(ray->pos[0]), (int64_t)floor(ray->pos[1]), (int64_t)floor(ray->pos[2]) };
            int id = get_voxel(w, corg, res, p);
            cnt[1]++;
            if (id) {
                const double *mat = w->sc->materials + 7 * (size_t)id;
                if (!ray->fflag) { ray->fflag = 1; ray->fhit[0] = p[0]; ray->fhit[1] = p[1]; ray->fhit[2] = p[2]; ray->first_step = ray->step; }
                ray->hits++; cnt[3]++;
                double bounce = material(ray, mat, s, g);                    /* init.py:81 */
                ray->bounces += bounce;
                ray->life /= res + bounce * s->lod_bounces;                  /* init.py:83 */
                double ref = fmax(fabs(ray->vel[0]), fmax(fabs(ray->vel[1]), fabs(ray->vel[2])));   /* lib.py:310-314 */
                if (ref != 0.0 && ref != 1.0) { ray->vel[0] /= ref; ray->vel[1] /= ref; ray->vel[2] /= ref; }
                if (ray->step >= ray->life || ray->energy >= s->max_light || ray->bounces >= s->max_bounces + 1)
                    break;                                                   /* init.py:85-86 */
                double ior = mat[5];
                if (ior != 0.0) {                                            /* init.py:92-111 */
                    double direction = (ior - 0.5) * 2;
                    int flip[3];
                    for (int a = 0; a < 3; a++) {
                        double pp[3] = { ray->pos[0], ray->pos[1], ray->pos[2] };
                        pp[a] = (ray->vel[a] < direction) ? pp[a] + 1 : pp[a] - 1;
                        int64_t pi[3] = { (int64_t)floor(pp[0]), (int64_t)floor(pp[1]), (int64_t)floor(pp[2]) };
                        int nid = 0;
                        if (inside_incl(pp, cmin, cmax)) {
                            nid = get_voxel(w, corg, res, pi); cnt[4]++;
                        } else {                                             /* Camera.chunk_get, init.py:28-33 */
                            double nmin[3]; int64_t norg[3];
                            for (int b = 0; b < 3; b++) { nmin[b] = vto_py_floordiv(pp[b], cs) * cs; norg[b] = (int64_t)nmin[b]; }
                            int64_t nslot = chunk_slot(w, nmin);
                            int nres = nslot >= 0 ? w->sc->chunk_res[nslot] : 0;
                            if (nres) { nid = get_voxel(w, norg, nres, pi); cnt[4]++; }
                        }
                        flip[a] = (!nid) || (w->sc->materials[7 * (size_t)nid + 5] != ior);
                    }
                    for (int a = 0; a < 3; a++)
                        if (flip[a]) ray->vel[a] -= ray->vel[a] * ior * 2;
                }
            }
        } else {
            cnt[2]++;
        }
        /* init.py:114-116 */
        double step;
        if (res) step = (double)res;
        else {
            double m = fmin(ray->pos[0], fmin(ray->pos[1], ray->pos[2]));
            step = 1 + fabs(s->chunk_radius - vto_py_mod(m + s->chunk_radius, cs));
        }
        ray->step += step;
        for (int a = 0; a < 3; a++) ray->pos[a] += ray->vel[a] * step;
    }
    if (s->background) material_background(ray, s);                          /* init.py:119-120 */
    cnt[0]++;
}

/* Camera.tile (init.py:126-150) over an explicit pixel list.  The reference fans pixel sets out over an
 * mp.Pool (init.py:162, 205); here worker threads pull blocks of 64 pixels from an atomic counter. */
typedef struct {
    const vto_settings *s; const vto_camera *cam; const world_t *w;
    const int32_t *pixels; int64_t n_pixels; int rng_mode; uint64_t seed;
    double *mean; uint8_t *rgba8; int32_t *nsamp; double *ray0;
    atomic_llong *next; uint64_t cnt[8];
} job_t;

static void tile_pixel(job_t *j, rng_t *g, int64_t i) {
    const vto_settings *s = j->s;
    int32_t x = j->pixels[2 * i], y = j->pixels[2 * i + 1];
    double dir_x = -1 + ((double)x / s->width) * 2;
    double dir_y = -1 + ((double)y / s->height) * 2;
    double detail = 1 - fabs(dir_x * dir_y) * s->lod_edge;
    int64_t n = (int64_t)py_round(s->samples * detail);
    if (n < 1) n = 1;
    double sum[4] = { 0, 0, 0, 0 };
    for (int64_t sample = 0; sample < n; sample++) {
        if (j->rng_mode == VTO_RNG_MT19937) {
            if (s->static_noise) mt_seed_u64(&g->mt, (uint64_t)(1 + x) * (uint64_t)(1 + y) * (uint64_t)(1 + sample));
        } else {
            uint64_t pix = (uint64_t)y * (uint64_t)s->width + (uint64_t)x;
            g->key0 = (uint32_t)pix; g->key1 = (uint32_t)(pix >> 32);
            g->sample = (uint32_t)sample; g->seed_lo = (uint32_t)j->seed; g->seed_hi = (uint32_t)(j->seed >> 32);
            g->draw = 0;
        }
        double u = rng_next(g);                                      /* always drawn, init.py:139 */
        double ray_detail = detail / (1 + sample * s->lod_samples) * (1 - s->lod_random * u);
        ray_t ray;
        trace(j->w, j->cam, dir_x, dir_y, ray_detail, g, &ray, j->cnt);
        double alpha = py_round(fmin(1.0, ray.energy + s->shutter) * 255);
        sum[0] += ray.cr; sum[1] += ray.cg; sum[2] += ray.cb; sum[3] += alpha;
        if (sample == 0 && j->ray0) {
            double *o = j->ray0 + 15 * i;
            o[0] = ray.cr; o[1] = ray.cg; o[2] = ray.cb; o[3] = ray.energy; o[4] = ray.bounces; o[5] = ray.step;
            o[6] = ray.hits; o[7] = (double)ray.fhit[0]; o[8] = (double)ray.fhit[1]; o[9] = (double)ray.fhit[2];
            o[10] = ray.fflag; o[11] = ray.first_step; o[12] = ray.vel[0]; o[13] = ray.vel[1]; o[14] = ray.vel[2];
        }
    }
    for (int c = 0; c < 4; c++) {
        double m = sum[c] / (double)n;                                /* lib.average, lib.py:419-428 */
        if (j->mean) j->mean[4 * i + c] = m;
        if (j->rgba8) { double t = trunc(m); j->rgba8[4 * i + c] = (uint8_t)(t < 0 ? 0 : t > 255 ? 255 : t); }
    }
    if (j->nsamp) j->nsamp[i] = (int32_t)n;
}

static void *tile_worker(void *arg) {
    job_t *j = (job_t *)arg;
    rng_t *g = (rng_t *)malloc(sizeof(rng_t));
    memset(g, 0, sizeof *g);
    g->mode = j->rng_mode;
    if (j->rng_mode == VTO_RNG_MT19937) mt_seed_u64(&g->mt, 0);
    for (;;) {
        int64_t b = atomic_fetch_add(j->next, 64);
        if (b >= j->n_pixels) break;
        int64_t e = b + 64 < j->n_pixels ? b + 64 : j->n_pixels;
        for (int64_t i = b; i < e; i++) tile_pixel(j, g, i);
    }
    free(g);
    return NULL;
}

int vto_render(const vto_settings *s, const vto_camera *cam, const vto_scene *scene,
               const int32_t *pixels, int64_t n_pixels, int rng_mode, uint64_t seed, int threads,
               double *mean, uint8_t *rgba8, int32_t *nsamp, double *ray0,
               uint8_t *visited, uint64_t *counters) {
    if (!s || !cam || !scene || !pixels || s->chunk_size <= 0) return 1;
    world_t w; w.s = s; w.sc = scene; w.visited = visited;
    int64_t nslots = 1;
    for (int a = 0; a < 3; a++) { w.nc[a] = scene->dims[a] / s->chunk_size; nslots *= w.nc[a]; }
    if (visited) memset(visited, 0, (size_t)nslots);
    if (threads < 1) threads = 1;
    if (threads > 1024) threads = 1024;
    atomic_llong next = 0;
    job_t *jobs = (job_t *)calloc((size_t)threads, sizeof(job_t));
    pthread_t *tids = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    for (int t = 0; t < threads; t++) {
        job_t *j = &jobs[t];
        j->s = s; j->cam = cam; j->w = &w; j->pixels = pixels; j->n_pixels = n_pixels; j->rng_mode = rng_mode;
        j->seed = seed; j->mean = mean; j->rgba8 = rgba8; j->nsamp = nsamp; j->ray0 = ray0; j->next = &next;
    }
    if (threads == 1) tile_worker(&jobs[0]);
    else {
        for (int t = 0; t < threads; t++) pthread_create(&tids[t], NULL, tile_worker, &jobs[t]);
        for (int t = 0; t < threads; t++) pthread_join(tids[t], NULL);
    }
    uint64_t total[8] = { 0 };
    for (int t = 0; t < threads; t++) for (int k = 0; k < 8; k++) total[k] += jobs[t].cnt[k];
    if (counters) memcpy(counters, total, sizeof total);
    free(jobs); free(tids);
    return 0;
}

/* Camera.trace (init.py:37-121) for caller-given (dir_x, dir_y, detail) triples, one after the other.  In Philox mode
 * ray i draws from the stream keyed i (sample 0, draw 0 first: trace() itself draws dof x, dof y, then the hits'
 * jitter -- the lod_random draw belongs to tile, init.py:139); in MT19937 mode the generator is seeded once with
 * `seed` like random.seed(seed) and runs on across rays like the interpreter's global generator would.
 * records [n][20]: r, g, b, energy, bounces, step, hits, first-hit x, y, z, flag, first-hit step, vel x, y, z, life,
 * pos x, y, z, 0. */
int vto_trace_rays(const vto_settings *s, const vto_camera *cam, const vto_scene *scene, const double *dirs, int64_t n,
                   int rng_mode, uint64_t seed, double *records, uint8_t *visited, uint64_t *counters) {
    if (!s || !cam || !scene || !dirs || !records || s->chunk_size <= 0) return 1;
    world_t w; w.s = s; w.sc = scene; w.visited = visited;
    int64_t nslots = 1;
    for (int a = 0; a < 3; a++) { w.nc[a] = scene->dims[a] / s->chunk_size; nslots *= w.nc[a]; }
    if (visited) memset(visited, 0, (size_t)nslots);
    rng_t *g = (rng_t *)calloc(1, sizeof(rng_t));
    g->mode = rng_mode;
    if (rng_mode == VTO_RNG_MT19937) mt_seed_u64(&g->mt, seed);
    uint64_t cnt[8] = { 0 };
    for (int64_t i = 0; i < n; i++) {
        if (rng_mode != VTO_RNG_MT19937) {
            g->key0 = (uint32_t)i; g->key1 = (uint32_t)((uint64_t)i >> 32);
            g->sample = 0; g->seed_lo = (uint32_t)seed; g->seed_hi = (uint32_t)(seed >> 32); g->draw = 0;
        }
        ray_t ray;
        trace(&w, cam, dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2], g, &ray, cnt);
        double *o = records + 20 * i;
        o[0] = ray.cr; o[1] = ray.cg; o[2] = ray.cb; o[3] = ray.energy; o[4] = ray.bounces; o[5] = ray.step;
        o[6] = ray.hits; o[7] = (double)ray.fhit[0]; o[8] = (double)ray.fhit[1]; o[9] = (double)ray.fhit[2];
        o[10] = ray.fflag; o[11] = ray.first_step; o[12] = ray.vel[0]; o[13] = ray.vel[1]; o[14] = ray.vel[2];
        o[15] = ray.life; o[16] = ray.pos[0]; o[17] = ray.pos[1]; o[18] = ray.pos[2]; o[19] = 0.0;
    }
    if (counters) memcpy(counters, cnt, sizeof cnt);
    free(g);
    return 0;
}
