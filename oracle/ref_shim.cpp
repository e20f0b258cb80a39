// TEST INFRASTRUCTURE ONLY.  Never linked into, loaded by, or called from the
// product path (fluid_sim_b200/); only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load the library this
// file is built into.
//
// extern "C" shim around the *unmodified* reference class
// (/root/reference/include/fluid_sim.h:17-77, /root/reference/src/fluid_sim.cpp),
// so that Python (ctypes) can drive the real reference through its public API and,
// for per-kernel parity tests, through its private methods.  Private access uses
// the `#define private public` trick in this translation unit only; the reference
// source file itself is compiled untouched (see oracle/build_ref.py).
//
// REF_PROJ_ITERS / REF_DIFF_ITERS are informational: they record which literal
// sweep counts the reference source compiled beside this shim carries
// (src/fluid_sim.cpp:319 and :50).
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>

#define private public
#include "fluid_sim.h"
#undef private

#ifndef REF_PROJ_ITERS
#define REF_PROJ_ITERS 20
#endif
#ifndef REF_DIFF_ITERS
#define REF_DIFF_ITERS 5
#endif

static int g_noise_mode = 0;  // 0: libc rand(); 1: constant giving exactly zero noise

// The library is linked with -Bsymbolic-functions, so the reference's call to
// rand() (src/fluid_sim.cpp:378-379) binds here.  Mode 0 forwards to libc.
// Mode 1 returns 0x3FFFFFFF: (float)0x3FFFFFFF rounds to 2^30, and
// 2^30 / (float)RAND_MAX == 0.5f exactly, so the injected noise is exactly 0.
extern "C" int rand(void) noexcept
{
    typedef int (*rand_fn)(void);
    static rand_fn real = nullptr;
    if (g_noise_mode == 1) return 0x3FFFFFFF;
    if (!real) real = (rand_fn)dlsym(RTLD_NEXT, "rand");
    return real();
}

typedef float (*field_t)[GRID_SIZE_Y];

static field_t field_of(FluidSim *s, int id)
{
    switch (id) {
    case 0: return s->density;
    case 1: return s->velocityX;
    case 2: return s->velocityY;
    case 3: return s->prevDensity;
    case 4: return s->prevVelocityX;
    case 5: return s->prevVelocityY;
    case 6: return s->tempField1;
    case 7: return s->tempField2;
    }
    return nullptr;
}

extern "C" {

int ref_grid_w(void) { return GRID_SIZE_X; }
int ref_grid_h(void) { return GRID_SIZE_Y; }
int ref_proj_iters(void) { return REF_PROJ_ITERS; }
int ref_diff_iters(void) { return REF_DIFF_ITERS; }

void *ref_create(void) { return new FluidSim(); }
void ref_destroy(void *h) { delete static_cast<FluidSim *>(h); }
void ref_reset(void *h)
{
    // main.cpp:289 does `sim = FluidSim();` with a stack temporary; the temporary is
    // heap-allocated here because sizeof(FluidSim) is 2 GiB at 8192 x 8192.
    FluidSim *fresh = new FluidSim();
    *static_cast<FluidSim *>(h) = *fresh;
    delete fresh;
}

// copy construction (`FluidSim b = a;`): a second object with the same fields; rand() stays process-global
void *ref_copy(void *h) { return new FluidSim(*static_cast<FluidSim *>(h)); }

void ref_srand(unsigned seed) { srand(seed); }
void ref_noise_mode(int m) { g_noise_mode = m; }
int ref_rand(void) { return rand(); }

void ref_step(void *h, float dt) { static_cast<FluidSim *>(h)->step(dt); }
void ref_add_density(void *h, int x, int y, float a) { static_cast<FluidSim *>(h)->addDensity(x, y, a); }
void ref_add_velocity(void *h, int x, int y, float ax, float ay) { static_cast<FluidSim *>(h)->addVelocity(x, y, ax, ay); }
float ref_get_density(void *h, int x, int y) { return static_cast<FluidSim *>(h)->getDensity(x, y); }
void ref_get_velocity(void *h, int x, int y, float *out)
{
    glm::vec2 v = static_cast<FluidSim *>(h)->getVelocity(x, y);
    out[0] = v.x; out[1] = v.y;
}
void ref_get_normalized_velocity(void *h, int x, int y, float *out)
{
    glm::vec2 v = static_cast<FluidSim *>(h)->getNormalizedVelocity(x, y);
    out[0] = v.x; out[1] = v.y;
}
float ref_get_velocity_magnitude(void *h, int x, int y) { return static_cast<FluidSim *>(h)->getVelocityMagnitude(x, y); }
void ref_get_velocity_color(void *h, int x, int y, float *out)
{
    glm::vec3 c = static_cast<FluidSim *>(h)->getVelocityColor(x, y);
    out[0] = c.x; out[1] = c.y; out[2] = c.z;
}
int ref_get_width(void *h) { return static_cast<FluidSim *>(h)->getWidth(); }
int ref_get_height(void *h) { return static_cast<FluidSim *>(h)->getHeight(); }

// ---- raw field access (private members) -------------------------------------
void ref_read_field(void *h, int id, float *out)
{
    std::memcpy(out, field_of(static_cast<FluidSim *>(h), id), sizeof(float) * (size_t)GRID_SIZE_X * GRID_SIZE_Y);
}
void ref_write_field(void *h, int id, const float *in)
{
    std::memcpy(field_of(static_cast<FluidSim *>(h), id), in, sizeof(float) * (size_t)GRID_SIZE_X * GRID_SIZE_Y);
}

// ---- private numerical methods, for per-kernel parity -------------------------
void ref_set_boundary(void *h, int b, int id)
{
    FluidSim *s = static_cast<FluidSim *>(h);
    s->setBoundary(b, field_of(s, id));
}
void ref_diffuse(void *h, int b, int dst, int src, float diff, float dt)
{
    FluidSim *s = static_cast<FluidSim *>(h);
    s->diffuse(b, field_of(s, dst), field_of(s, src), diff, dt);
}
void ref_project(void *h, int u, int v, int p, int div)
{
    FluidSim *s = static_cast<FluidSim *>(h);
    s->project(field_of(s, u), field_of(s, v), field_of(s, p), field_of(s, div));
}
void ref_advect(void *h, int b, int dst, int src, int u, int v, float dt)
{
    FluidSim *s = static_cast<FluidSim *>(h);
    s->advect(b, field_of(s, dst), field_of(s, src), field_of(s, u), field_of(s, v), dt);
}
// scheme: 0 = RK4 (the live path, :219), 1 = MacCormack (:223), 2 = semi-Lagrangian (:227)
void ref_advect_scheme(void *h, int scheme, int b, int dst, int src, int u, int v, float dt)
{
    FluidSim *s = static_cast<FluidSim *>(h);
    if (scheme == 0) s->rk4Advect(b, field_of(s, dst), field_of(s, src), field_of(s, u), field_of(s, v), dt);
    else if (scheme == 1) s->macCormackAdvect(b, field_of(s, dst), field_of(s, src), field_of(s, u), field_of(s, v), dt);
    else s->semiLagrangianAdvect(b, field_of(s, dst), field_of(s, src), field_of(s, u), field_of(s, v), dt);
}
float ref_bilinear(void *h, int id, float x, float y)
{
    FluidSim *s = static_cast<FluidSim *>(h);
    return s->bilinearInterpolate(field_of(s, id), x, y);
}
void ref_velocity_step(void *h, float dt) { static_cast<FluidSim *>(h)->velocityStep(dt); }
void ref_density_step(void *h, float dt) { static_cast<FluidSim *>(h)->densityStep(dt); }

}  // extern "C"
