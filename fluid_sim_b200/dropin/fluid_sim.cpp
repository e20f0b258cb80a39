// FluidSim (include/fluid_sim.h) implemented on the C ABI of libfluidsim_b200.so.
// Error convention: the reference class has no error channel; a CUDA failure here
// prints the library's message and aborts.
#include "fluid_sim.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "fluid_sim_b200.h"

namespace {
void check(int rc, const char *what)
{
    if (rc != FSB_OK) {
        std::fprintf(stderr, "FluidSim: %s failed (%d): %s\n", what, rc, fsb_last_error());
        std::abort();
    }
}
}  // namespace

FluidSim::FluidSim()
    : width(GRID_SIZE_X), height(GRID_SIZE_Y), projSweeps(20), diffSweeps(5), visc(VISCOSITY), diff(DIFFUSION), sim(nullptr)
{
    create();
}

FluidSim::FluidSim(int w, int h, int projectionSweeps, int diffusionSweeps, float viscosity, float diffusion)
    : width(w), height(h), projSweeps(projectionSweeps), diffSweeps(diffusionSweeps), visc(viscosity), diff(diffusion), sim(nullptr)
{
    create();
}

void FluidSim::create()
{
    fsb_params p;
    fsb_default_params(&p);
    p.width = width;
    p.height = height;
    p.proj_iters = projSweeps;
    p.diff_iters = diffSweeps;
    p.viscosity = visc;
    p.diffusion = diff;
    // rand() is process-global in the reference (src/fluid_sim.cpp:378-379, no srand anywhere): every FluidSim of
    // the process draws from one stream, so copies, resets and objects stepped in turn continue it
    p.noise_process_stream = 1;
    check(fsb_create(&p, &sim), "fsb_create");
}

void FluidSim::destroy()
{
    if (sim) fsb_destroy(sim);
    sim = nullptr;
}

void FluidSim::copyStateFrom(const FluidSim &o)
{
    // the three live fields; the position in the rand() stream is shared by all objects (see create())
    std::vector<float> buf((size_t)width * height);
    for (int f = FSB_DENSITY; f <= FSB_VELOCITY_Y; ++f) {
        check(fsb_read_field(o.sim, f, buf.data()), "fsb_read_field");
        check(fsb_write_field(sim, f, buf.data()), "fsb_write_field");
    }
}

FluidSim::FluidSim(const FluidSim &o)
    : width(o.width), height(o.height), projSweeps(o.projSweeps), diffSweeps(o.diffSweeps), visc(o.visc), diff(o.diff), sim(nullptr)
{
    create();
    copyStateFrom(o);
}

FluidSim::FluidSim(FluidSim &&o) noexcept
    : width(o.width), height(o.height), projSweeps(o.projSweeps), diffSweeps(o.diffSweeps), visc(o.visc), diff(o.diff), sim(o.sim)
{
    o.sim = nullptr;
}

FluidSim &FluidSim::operator=(const FluidSim &o)
{
    if (this == &o) return *this;
    if (width != o.width || height != o.height || projSweeps != o.projSweeps || diffSweeps != o.diffSweeps ||
        visc != o.visc || diff != o.diff) {
        destroy();
        width = o.width; height = o.height; projSweeps = o.projSweeps; diffSweeps = o.diffSweeps;
        visc = o.visc; diff = o.diff;
        create();
    }
    copyStateFrom(o);
    return *this;
}

FluidSim &FluidSim::operator=(FluidSim &&o) noexcept
{
    if (this == &o) return *this;
    // `sim = FluidSim();` (src/main.cpp:289): take the other object's (zero) fields; the rand() stream is the
    // process's, so it simply continues
    destroy();
    width = o.width; height = o.height; projSweeps = o.projSweeps; diffSweeps = o.diffSweeps;
    visc = o.visc; diff = o.diff;
    sim = o.sim;
    o.sim = nullptr;
    return *this;
}

FluidSim::~FluidSim() { destroy(); }

void FluidSim::step(float dt) { check(fsb_step(sim, dt), "fsb_step"); }
void FluidSim::addDensity(int x, int y, float amount) { check(fsb_add_density(sim, x, y, amount), "fsb_add_density"); }
void FluidSim::addVelocity(int x, int y, float ax, float ay) { check(fsb_add_velocity(sim, x, y, ax, ay), "fsb_add_velocity"); }

float FluidSim::getDensity(int x, int y) const
{
    float v = 0.0f;
    check(fsb_get_density(sim, x, y, &v), "fsb_get_density");
    return v;
}

glm::vec2 FluidSim::getVelocity(int x, int y) const
{
    float v[2];
    check(fsb_get_velocity(sim, x, y, v), "fsb_get_velocity");
    return glm::vec2(v[0], v[1]);
}

glm::vec2 FluidSim::getNormalizedVelocity(int x, int y) const
{
    float v[2];
    check(fsb_get_normalized_velocity(sim, x, y, v), "fsb_get_normalized_velocity");
    return glm::vec2(v[0], v[1]);
}

float FluidSim::getVelocityMagnitude(int x, int y) const
{
    float v = 0.0f;
    check(fsb_get_velocity_magnitude(sim, x, y, &v), "fsb_get_velocity_magnitude");
    return v;
}

glm::vec3 FluidSim::getVelocityColor(int x, int y) const
{
    float c[3];
    check(fsb_get_velocity_color(sim, x, y, c), "fsb_get_velocity_color");
    return glm::vec3(c[0], c[1], c[2]);
}

void FluidSim::readDensity(float *host) const { check(fsb_read_field(sim, FSB_DENSITY, host), "fsb_read_field"); }
void FluidSim::readVelocity(float *hx, float *hy) const
{
    check(fsb_read_field(sim, FSB_VELOCITY_X, hx), "fsb_read_field");
    check(fsb_read_field(sim, FSB_VELOCITY_Y, hy), "fsb_read_field");
}
