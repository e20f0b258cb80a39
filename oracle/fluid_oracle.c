/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C, runtime-sized) of the
 * reference's per-timestep solver, used solely as the checker in tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg.  The product path
 * (fluid_sim_b200/) never links, loads or calls this file.
 *
 * Parity status: PINNED.  This file is gated bit-for-bit against the real
 * reference compiled from /root/reference (oracle/_ref, see oracle/build_ref.py)
 * by tests/test_oracle.py, and against SURVEY.md A.4's known answers for the demo
 * run.  The reference ships no tests or golden vectors of its own.
 *
 * Why it exists beside oracle/_ref: the reference's grid size and sweep counts are
 * compile-time literals (include/fluid_sim.h:11-12, src/fluid_sim.cpp:50,319);
 * this restatement takes them at run time, so parity cases are not limited to the
 * variants that were pre-built here for the GPU box.
 *
 * Floating point: every expression keeps the reference's association order and is
 * compiled with -O2 -ffp-contract=off (the reference's x86-64 -O2 build has no FMA).
 * Storage is f[i*H + j], i = x (slow), j = y (contiguous), like float f[W][H]
 * (include/fluid_sim.h:42-53).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

enum { F_D = 0, F_U, F_V, F_PD, F_PU, F_PV, F_T1, F_T2, F_COUNT };

typedef struct {
    int W, H;            /* width (i extent), height (j extent) */
    int P, D;            /* projection sweeps (:319), diffusion sweeps (:50) */
    float visc, diff;    /* VISCOSITY, DIFFUSION (include/fluid_sim.h:13-14) */
    float omega;         /* SOR factor (:47) */
    int scheme;          /* 0 RK4 (:219, live), 1 MacCormack (:223), 2 semi-Lagrangian (:227) */
    int noise_mode;      /* 0 glibc default-seed stream, 1 zero noise */
    /* glibc TYPE_3 additive-feedback generator state: last 34 raw words */
    uint32_t rs[34];
    int rpos;
    uint64_t rcount;     /* outputs consumed so far */
    float *f[F_COUNT];
} ora_t;

/* ---- glibc rand(), default seed (the reference never calls srand) ---------------
 * Restates glibc's random_r TYPE_3 generator as used by rand(): degree 31,
 * separation 3; srandom_r fills r[1..30] with the Lehmer step 16807*r mod (2^31-1)
 * evaluated by Schrage's method, then discards 310 outputs.  Output = word >> 1. */
static void ora_rand_seed(ora_t *s, uint32_t seed)
{
    int32_t r[34];
    int i;
    if (seed == 0) seed = 1;
    r[0] = (int32_t)seed;
    for (i = 1; i < 31; i++) {
        int32_t hi = r[i - 1] / 127773, lo = r[i - 1] % 127773;
        int32_t w = 16807 * lo - 2836 * hi;
        if (w < 0) w += 2147483647;
        r[i] = w;
    }
    for (i = 31; i < 34; i++) r[i] = r[i - 31];
    for (i = 0; i < 34; i++) s->rs[i] = (uint32_t)r[i];
    s->rpos = 0;      /* rs is a ring over the last 34 words; rs[rpos] is the oldest */
    s->rcount = 0;
    /* ring holds words 0..33; word n = word[n-31] + word[n-3]; discard words 34..343 */
    for (i = 34; i < 344; i++) {
        uint32_t w = s->rs[(s->rpos + 3) % 34] + s->rs[(s->rpos + 31) % 34];
        s->rs[s->rpos] = w;
        s->rpos = (s->rpos + 1) % 34;
    }
}

static int ora_rand_next(ora_t *s)
{
    /* oldest entry is word n-34: word[n-31] sits 3 after it, word[n-3] sits 31 after it */
    uint32_t w = s->rs[(s->rpos + 3) % 34] + s->rs[(s->rpos + 31) % 34];
    s->rs[s->rpos] = w;
    s->rpos = (s->rpos + 1) % 34;
    s->rcount++;
    return (int)(w >> 1);
}

/* ---- boundary conditions (src/fluid_sim.cpp:347-367) ------------------------------ */
static void set_boundary(const ora_t *s, int b, float *x)
{
    const int W = s->W, H = s->H;
    int i, j;
    for (i = 1; i < W - 1; i++) {
        x[i * H + 0] = (b == 2) ? -x[i * H + 1] : x[i * H + 1];
        x[i * H + H - 1] = (b == 2) ? -x[i * H + H - 2] : x[i * H + H - 2];
    }
    for (j = 1; j < H - 1; j++) {
        x[0 * H + j] = (b == 1) ? -x[1 * H + j] : x[1 * H + j];
        x[(W - 1) * H + j] = (b == 1) ? -x[(W - 2) * H + j] : x[(W - 2) * H + j];
    }
    x[0] = 0.5f * (x[1 * H + 0] + x[0 * H + 1]);
    x[H - 1] = 0.5f * (x[1 * H + H - 1] + x[0 * H + H - 2]);
    x[(W - 1) * H] = 0.5f * (x[(W - 2) * H] + x[(W - 1) * H + 1]);
    x[(W - 1) * H + H - 1] = 0.5f * (x[(W - 2) * H + H - 1] + x[(W - 1) * H + H - 2]);
}

/* ---- SOR diffusion (src/fluid_sim.cpp:43-62): lexicographic, in place ---------------- */
static void diffuse(const ora_t *s, int b, float *x, const float *src, float diff, float dt)
{
    const int W = s->W, H = s->H;
    const float a = dt * diff * W * H;          /* ((dt*diff)*W)*H, ints converted to float */
    const float c_recip = 1.0f / (1 + 4 * a);
    const float omega = s->omega;
    int k, i, j;
    for (k = 0; k < s->D; k++) {
        for (i = 1; i < W - 1; i++) {
            for (j = 1; j < H - 1; j++) {
                float nb = x[(i + 1) * H + j] + x[(i - 1) * H + j] + x[i * H + j + 1] + x[i * H + j - 1];
                float nv = (src[i * H + j] + a * nb) * c_recip;
                x[i * H + j] = x[i * H + j] + omega * (nv - x[i * H + j]);
            }
        }
        set_boundary(s, b, x);
    }
}

/* ---- bilinear sample (src/fluid_sim.cpp:231-252) ------------------------------------------ */
static float bilinear(const ora_t *s, const float *f, float x, float y)
{
    const int W = s->W, H = s->H;
    float xm = W - 1.5f, ym = H - 1.5f;
    int i0, i1, j0, j1;
    float s0, s1, t0, t1;
    x = (x < xm) ? x : xm;   /* std::min(width - 1.5f, x) */
    x = (0.5f < x) ? x : 0.5f; /* std::max(0.5f, .) */
    y = (y < ym) ? y : ym;
    y = (0.5f < y) ? y : 0.5f;
    i0 = (int)x; i1 = i0 + 1;
    j0 = (int)y; j1 = j0 + 1;
    s1 = x - i0; s0 = 1 - s1;
    t1 = y - j0; t0 = 1 - t1;
    return s0 * (t0 * f[i0 * H + j0] + t1 * f[i0 * H + j1]) +
           s1 * (t0 * f[i1 * H + j0] + t1 * f[i1 * H + j1]);
}

/* ---- RK4 back-trace (src/fluid_sim.cpp:261-298), the live scheme (:219) ---------------------- */
static void advect_rk4(const ora_t *s, int b, float *dst, const float *src, const float *u, const float *v, float dt)
{
    const int W = s->W, H = s->H;
    const float dt0 = dt * W;
    const float ndt0 = -dt0;
    int i, j;
    for (i = 1; i < W - 1; i++) {
        for (j = 1; j < H - 1; j++) {
            float x = (float)i, y = (float)j;
            float k1x = bilinear(s, u, x, y) * ndt0, k1y = bilinear(s, v, x, y) * ndt0;
            float p2x = x + k1x * 0.5f, p2y = y + k1y * 0.5f;
            float k2x = bilinear(s, u, p2x, p2y) * ndt0, k2y = bilinear(s, v, p2x, p2y) * ndt0;
            float p3x = x + k2x * 0.5f, p3y = y + k2y * 0.5f;
            float k3x = bilinear(s, u, p3x, p3y) * ndt0, k3y = bilinear(s, v, p3x, p3y) * ndt0;
            float p4x = x + k3x, p4y = y + k3y;
            float k4x = bilinear(s, u, p4x, p4y) * ndt0, k4y = bilinear(s, v, p4x, p4y) * ndt0;
            float dx = (k1x + 2.0f * k2x + 2.0f * k3x + k4x) / 6.0f;
            float dy = (k1y + 2.0f * k2y + 2.0f * k3y + k4y) / 6.0f;
            dst[i * H + j] = bilinear(s, src, x + dx, y + dy);
        }
    }
    set_boundary(s, b, dst);
}

/* One semi-Lagrangian sweep with displacement sign sgn: the pattern shared by
 * src/fluid_sim.cpp:65-100 (sgn=-1) and both MacCormack half steps (:110-168). */
static void sl_pass(const ora_t *s, float *dst, const float *src, const float *u, const float *v, float dt0, float sgn)
{
    const int W = s->W, H = s->H;
    int i, j;
    for (i = 1; i < W - 1; i++) {
        for (j = 1; j < H - 1; j++) {
            float x = (sgn < 0) ? i - dt0 * u[i * H + j] : i + dt0 * u[i * H + j];
            float y = (sgn < 0) ? j - dt0 * v[i * H + j] : j + dt0 * v[i * H + j];
            dst[i * H + j] = bilinear(s, src, x, y);
        }
    }
}

static void advect_sl(const ora_t *s, int b, float *dst, const float *src, const float *u, const float *v, float dt)
{
    sl_pass(s, dst, src, u, v, dt * s->W, -1.0f);
    set_boundary(s, b, dst);
}

/* MacCormack with neighbourhood clamp (src/fluid_sim.cpp:103-209) */
static void advect_maccormack(ora_t *s, int b, float *dst, const float *src, const float *u, const float *v, float dt)
{
    const int W = s->W, H = s->H;
    const float dt0 = dt * W;
    float *t1 = s->f[F_T1], *t2 = s->f[F_T2];
    int i, j, di, dj;
    sl_pass(s, t1, src, u, v, dt0, -1.0f);
    set_boundary(s, b, t1);
    sl_pass(s, t2, t1, u, v, dt0, +1.0f);
    set_boundary(s, b, t2);
    for (i = 1; i < W - 1; i++) {
        for (j = 1; j < H - 1; j++) {
            float err = src[i * H + j] - t2[i * H + j];
            float r = t1[i * H + j] + 0.5f * err;
            float lo = src[i * H + j], hi = src[i * H + j];
            for (di = -1; di <= 1; di++)
                for (dj = -1; dj <= 1; dj++) {
                    int ni = i + di, nj = j + dj;
                    if (ni >= 0 && ni < W && nj >= 0 && nj < H) {
                        float c = src[ni * H + nj];
                        lo = (c < lo) ? c : lo;   /* std::min(minVal, c) */
                        hi = (hi < c) ? c : hi;   /* std::max(maxVal, c) */
                    }
                }
            r = (r < hi) ? r : hi;                /* std::min(maxVal, r) */
            r = (lo < r) ? r : lo;                /* std::max(minVal, .) */
            dst[i * H + j] = r;
        }
    }
    set_boundary(s, b, dst);
}

static void advect(ora_t *s, int b, float *dst, const float *src, const float *u, const float *v, float dt)
{
    if (s->scheme == 0) advect_rk4(s, b, dst, src, u, v, dt);
    else if (s->scheme == 1) advect_maccormack(s, b, dst, src, u, v, dt);
    else advect_sl(s, b, dst, src, u, v, dt);
}

/* ---- projection (src/fluid_sim.cpp:301-344): lexicographic Gauss-Seidel, in place -------------- */
static void project(const ora_t *s, float *u, float *v, float *p, float *div)
{
    const int W = s->W, H = s->H;
    const float h = 1.0f / W;
    int i, j, k;
    for (i = 1; i < W - 1; i++)
        for (j = 1; j < H - 1; j++) {
            div[i * H + j] = -0.5f * h * (u[(i + 1) * H + j] - u[(i - 1) * H + j] + v[i * H + j + 1] - v[i * H + j - 1]);
            p[i * H + j] = 0;
        }
    set_boundary(s, 0, div);
    set_boundary(s, 0, p);
    for (k = 0; k < s->P; k++) {
        for (i = 1; i < W - 1; i++)
            for (j = 1; j < H - 1; j++)
                p[i * H + j] = (div[i * H + j] + p[(i + 1) * H + j] + p[(i - 1) * H + j] + p[i * H + j + 1] + p[i * H + j - 1]) / 4;
        set_boundary(s, 0, p);
    }
    for (i = 1; i < W - 1; i++)
        for (j = 1; j < H - 1; j++) {
            u[i * H + j] -= 0.5f * (p[(i + 1) * H + j] - p[(i - 1) * H + j]) / h;
            v[i * H + j] -= 0.5f * (p[i * H + j + 1] - p[i * H + j - 1]) / h;
        }
    set_boundary(s, 1, u);
    set_boundary(s, 2, v);
}

/* ---- velocity noise (src/fluid_sim.cpp:374-383): full grid, X then Y per cell --------------------- */
static void add_noise(ora_t *s)
{
    const size_t n = (size_t)s->W * s->H;
    size_t c;
    float *u = s->f[F_U], *v = s->f[F_V];
    for (c = 0; c < n; c++) {
        int rx = (s->noise_mode == 0) ? ora_rand_next(s) : 0x3FFFFFFF;
        int ry = (s->noise_mode == 0) ? ora_rand_next(s) : 0x3FFFFFFF;
        float nx = ((float)rx / 2147483647 - 0.5f) * 1e-4f;  /* (float)RAND_MAX == 2^31 */
        float ny = ((float)ry / 2147483647 - 0.5f) * 1e-4f;
        u[c] += nx;
        v[c] += ny;
    }
}

static void velocity_step(ora_t *s, float dt)
{
    const size_t bytes = sizeof(float) * (size_t)s->W * s->H;
    float *u = s->f[F_U], *v = s->f[F_V], *pu = s->f[F_PU], *pv = s->f[F_PV];
    add_noise(s);
    memcpy(pu, u, bytes); memcpy(pv, v, bytes);
    diffuse(s, 1, u, pu, s->visc, dt);
    diffuse(s, 2, v, pv, s->visc, dt);
    project(s, u, v, pu, pv);
    memcpy(pu, u, bytes); memcpy(pv, v, bytes);
    advect(s, 1, u, pu, pu, pv, dt);
    advect(s, 2, v, pv, pu, pv, dt);
    project(s, u, v, pu, pv);
}

static void density_step(ora_t *s, float dt)
{
    const size_t bytes = sizeof(float) * (size_t)s->W * s->H;
    float *d = s->f[F_D], *pd = s->f[F_PD];
    memcpy(pd, d, bytes);
    diffuse(s, 0, d, pd, s->diff, dt);
    memcpy(pd, d, bytes);
    advect(s, 0, d, pd, s->f[F_U], s->f[F_V], dt);
}

/* ================================ C interface ========================================= */

void *ora_create(int W, int H, int P, int D)
{
    ora_t *s = (ora_t *)calloc(1, sizeof(ora_t));
    int k;
    if (!s) return NULL;
    s->W = W; s->H = H; s->P = P; s->D = D;
    s->visc = 0.0001f; s->diff = 0.0f; s->omega = 1.5f;
    s->scheme = 0; s->noise_mode = 0;
    ora_rand_seed(s, 1);
    for (k = 0; k < F_COUNT; k++) {
        s->f[k] = (float *)calloc((size_t)W * H, sizeof(float));
        if (!s->f[k]) return NULL;
    }
    return s;
}

void ora_destroy(void *h)
{
    ora_t *s = (ora_t *)h;
    int k;
    if (!s) return;
    for (k = 0; k < F_COUNT; k++) free(s->f[k]);
    free(s);
}

void ora_reset(void *h)  /* `sim = FluidSim()` (src/main.cpp:289): zero fields, rand() state untouched */
{
    ora_t *s = (ora_t *)h;
    int k;
    for (k = 0; k < F_COUNT; k++) memset(s->f[k], 0, sizeof(float) * (size_t)s->W * s->H);
}

void ora_set_params(void *h, float visc, float diff, float omega, int scheme)
{
    ora_t *s = (ora_t *)h;
    s->visc = visc; s->diff = diff; s->omega = omega; s->scheme = scheme;
}

void ora_set_noise(void *h, int mode, uint32_t seed, uint64_t skip)
{
    ora_t *s = (ora_t *)h;
    uint64_t k;
    s->noise_mode = mode;
    ora_rand_seed(s, seed);
    for (k = 0; k < skip; k++) (void)ora_rand_next(s);
}

int ora_rand(void *h) { return ora_rand_next((ora_t *)h); }
uint64_t ora_rand_count(void *h) { return ((ora_t *)h)->rcount; }

void ora_step(void *h, float dt)  /* src/fluid_sim.cpp:23-28 */
{
    velocity_step((ora_t *)h, dt);
    density_step((ora_t *)h, dt);
}
void ora_velocity_step(void *h, float dt) { velocity_step((ora_t *)h, dt); }
void ora_density_step(void *h, float dt) { density_step((ora_t *)h, dt); }

void ora_add_density(void *h, int x, int y, float a)  /* :424-430 */
{
    ora_t *s = (ora_t *)h;
    if (x >= 0 && x < s->W && y >= 0 && y < s->H) s->f[F_D][x * s->H + y] += a;
}

void ora_add_velocity(void *h, int x, int y, float ax, float ay)  /* :432-439 */
{
    ora_t *s = (ora_t *)h;
    if (x >= 0 && x < s->W && y >= 0 && y < s->H) {
        s->f[F_U][x * s->H + y] += ax;
        s->f[F_V][x * s->H + y] += ay;
    }
}

void ora_read_field(void *h, int id, float *out)
{
    ora_t *s = (ora_t *)h;
    memcpy(out, s->f[id], sizeof(float) * (size_t)s->W * s->H);
}

void ora_write_field(void *h, int id, const float *in)
{
    ora_t *s = (ora_t *)h;
    memcpy(s->f[id], in, sizeof(float) * (size_t)s->W * s->H);
}

void ora_set_boundary(void *h, int b, int id) { ora_t *s = (ora_t *)h; set_boundary(s, b, s->f[id]); }
void ora_diffuse(void *h, int b, int dst, int src, float diff, float dt)
{
    ora_t *s = (ora_t *)h;
    diffuse(s, b, s->f[dst], s->f[src], diff, dt);
}
void ora_project(void *h, int u, int v, int p, int div)
{
    ora_t *s = (ora_t *)h;
    project(s, s->f[u], s->f[v], s->f[p], s->f[div]);
}
void ora_advect(void *h, int b, int dst, int src, int u, int v, float dt)
{
    ora_t *s = (ora_t *)h;
    advect(s, b, s->f[dst], s->f[src], s->f[u], s->f[v], dt);
}
void ora_add_noise(void *h) { add_noise((ora_t *)h); }
float ora_bilinear(void *h, int id, float x, float y)
{
    ora_t *s = (ora_t *)h;
    return bilinear(s, s->f[id], x, y);
}

/* ---- visualisation getters (src/fluid_sim.cpp:459-519) ---------------------------------------- */
float ora_velocity_magnitude(void *h, int x, int y)
{
    ora_t *s = (ora_t *)h;
    if (x >= 0 && x < s->W && y >= 0 && y < s->H) {
        float a = s->f[F_U][x * s->H + y], b = s->f[F_V][x * s->H + y];
        return sqrtf(a * a + b * b);
    }
    return 0.0f;
}

void ora_normalized_velocity(void *h, int x, int y, float *out)
{
    ora_t *s = (ora_t *)h;
    out[0] = out[1] = 0.0f;
    if (x >= 0 && x < s->W && y >= 0 && y < s->H) {
        float a = s->f[F_U][x * s->H + y], b = s->f[F_V][x * s->H + y];
        float m = sqrtf(a * a + b * b);
        if (m > 0.001f) { out[0] = a / m; out[1] = b / m; }
    }
}

void ora_velocity_color(void *h, int x, int y, float *rgb)
{
    float speed = ora_velocity_magnitude(h, x, y);
    float q = speed / 10.0f;
    float n = (q < 1.0f) ? q : 1.0f;
    if (n < 0.25f)      { float t = n / 0.25f;           rgb[0] = 0.0f; rgb[1] = t;        rgb[2] = 1.0f; }
    else if (n < 0.5f)  { float t = (n - 0.25f) / 0.25f; rgb[0] = 0.0f; rgb[1] = 1.0f;     rgb[2] = 1.0f - t; }
    else if (n < 0.75f) { float t = (n - 0.5f) / 0.25f;  rgb[0] = t;    rgb[1] = 1.0f;     rgb[2] = 0.0f; }
    else                { float t = (n - 0.75f) / 0.25f; rgb[0] = 1.0f; rgb[1] = 1.0f - t; rgb[2] = 0.0f; }
}
