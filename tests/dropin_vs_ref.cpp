// Side-by-side run of the drop-in class (include/fluid_sim.h -> libfluidsim_dropin.so -> C ABI -> CUDA) and the
// compiled reference (oracle/_ref/libref_200x200_p20_d5.so, reached through its extern "C" shim with dlopen, so the
// two FluidSim classes never meet in one symbol table).  Checks the object semantics the reference gets from its
// process-global rand() (src/fluid_sim.cpp:378-379) and value-type class (src/main.cpp:7, :289):
//   1. copy-construct after 3 steps, then step original and copy in turn;
//   2. two objects with different sources stepped alternately;
//   3. `sim = FluidSim();` followed by new sources and steps (the noise stream is not rewound).
// All fields must agree bit for bit after every step.  usage: dropin_vs_ref <path to libref_200x200_p20_d5.so>
#include "fluid_sim.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <vector>

namespace {
void *lib;
template <typename T> T sym(const char *n)
{
    void *p = dlsym(lib, n);
    if (!p) { std::fprintf(stderr, "missing %s\n", n); std::exit(2); }
    return reinterpret_cast<T>(p);
}
void *(*ref_create)();
void *(*ref_copy)(void *);
void (*ref_destroy)(void *);
void (*ref_reset)(void *);
void (*ref_step)(void *, float);
void (*ref_add_density)(void *, int, int, float);
void (*ref_add_velocity)(void *, int, int, float, float);
void (*ref_read_field)(void *, int, float *);

int failures = 0;

void compare(const FluidSim &a, void *r, const char *what)
{
    const size_t n = (size_t)GRID_SIZE_X * GRID_SIZE_Y;
    std::vector<float> d(n), u(n), v(n), rd(n), ru(n), rv(n);
    a.readDensity(d.data());
    a.readVelocity(u.data(), v.data());
    ref_read_field(r, 0, rd.data());
    ref_read_field(r, 1, ru.data());
    ref_read_field(r, 2, rv.data());
    const bool ok = !std::memcmp(d.data(), rd.data(), 4 * n) && !std::memcmp(u.data(), ru.data(), 4 * n) &&
                    !std::memcmp(v.data(), rv.data(), 4 * n);
    std::printf("%s: %s\n", what, ok ? "identical" : "DIFFERENT");
    if (!ok) failures++;
}

void inject(FluidSim &a, void *r, int cx, int cy, float dens, float vx, float vy)
{
    for (int i = -5; i <= 5; i++)
        for (int j = -5; j <= 5; j++)
            if (i * i + j * j < 25) {
                a.addDensity(cx + i, cy + j, dens);
                a.addVelocity(cx + i, cy + j, vx, vy);
                ref_add_density(r, cx + i, cy + j, dens);
                ref_add_velocity(r, cx + i, cy + j, vx, vy);
            }
}
}  // namespace

int main(int argc, char **argv)
{
    if (argc < 2) return 2;
    lib = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!lib) { std::fprintf(stderr, "%s\n", dlerror()); return 2; }
    ref_create = sym<void *(*)()>("ref_create");
    ref_copy = sym<void *(*)(void *)>("ref_copy");
    ref_destroy = sym<void (*)(void *)>("ref_destroy");
    ref_reset = sym<void (*)(void *)>("ref_reset");
    ref_step = sym<void (*)(void *, float)>("ref_step");
    ref_add_density = sym<void (*)(void *, int, int, float)>("ref_add_density");
    ref_add_velocity = sym<void (*)(void *, int, int, float, float)>("ref_add_velocity");
    ref_read_field = sym<void (*)(void *, int, float *)>("ref_read_field");

    // 1. copy after three steps
    FluidSim a;
    void *ra = ref_create();
    inject(a, ra, 100, 100, 10.0f, 0.0f, 2.0f);
    for (int k = 0; k < 3; ++k) { a.step(0.01f); ref_step(ra, 0.01f); }
    compare(a, ra, "original after 3 steps");
    FluidSim b = a;
    void *rb = ref_copy(ra);
    compare(b, rb, "copy right after construction");
    for (int k = 0; k < 2; ++k) {
        a.step(0.01f); ref_step(ra, 0.01f);
        b.step(0.01f); ref_step(rb, 0.01f);
    }
    compare(a, ra, "original, stepped in turn with its copy");
    compare(b, rb, "copy, stepped in turn with the original");

    // 2. two unrelated objects stepped alternately
    FluidSim x, y;
    void *rx = ref_create(), *ry = ref_create();
    inject(x, rx, 60, 80, 5.0f, 1.0f, -1.5f);
    inject(y, ry, 140, 120, 7.0f, -2.0f, 0.5f);
    for (int k = 0; k < 2; ++k) {
        x.step(0.01f); ref_step(rx, 0.01f);
        y.step(0.01f); ref_step(ry, 0.01f);
    }
    compare(x, rx, "first of two objects stepped alternately");
    compare(y, ry, "second of two objects stepped alternately");

    // 3. reset idiom, then go on
    a = FluidSim();
    ref_reset(ra);
    compare(a, ra, "after sim = FluidSim()");
    inject(a, ra, 90, 110, 10.0f, 0.0f, 2.0f);
    for (int k = 0; k < 2; ++k) { a.step(0.01f); ref_step(ra, 0.01f); }
    compare(a, ra, "two steps after the reset (noise stream continues)");

    // copy assignment between existing objects
    y = x;
    ref_destroy(ry);
    ry = ref_copy(rx);
    y.step(0.01f); ref_step(ry, 0.01f);
    compare(y, ry, "copy-assigned object after one step");

    ref_destroy(ra); ref_destroy(rb); ref_destroy(rx); ref_destroy(ry);
    std::printf(failures ? "FAILED %d\n" : "ALL IDENTICAL\n", failures);
    return failures ? 1 : 0;
}
