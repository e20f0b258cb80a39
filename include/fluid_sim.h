// Drop-in replacement for the reference's include/fluid_sim.h (reference: include/fluid_sim.h:1-78).
//
// Same class name, same eleven public members with the same meaning, same grid-size constants, so the
// reference's src/main.cpp compiles against this header unchanged.  The object no longer holds eight
// in-object float[200][200] arrays: it owns a handle to a simulation that lives on one B200
// (libfluidsim_b200.so, C ABI in include/fluid_sim_b200.h), and there is no CPU fallback.
//
// Differences a caller can observe:
//   * copying a FluidSim copies the whole device state (the reference's `sim = FluidSim();` reset idiom,
//     src/main.cpp:289, keeps working: assigning a fresh object clears the fields); the velocity noise comes from one
//     rand() stream per process, as in the reference, so copies and objects stepped in turn continue it
//     (tests/dropin_vs_ref.cpp runs those cases side by side with the compiled reference);
//   * step() returns before the GPU has finished; every getter synchronises;
//   * an additional constructor takes run-time sizes and sweep counts (the reference's are literals).
#ifndef FLUID_SIM_H
#define FLUID_SIM_H

#if defined(__has_include)
#if __has_include(<glm/glm.hpp>)
#include <glm/glm.hpp>
#define FSB_HAVE_GLM 1
#endif
#endif
#ifndef FSB_HAVE_GLM
#include "fsb_mini_glm.h"   // vec2 / vec3 value types only, for machines without libglm-dev
#endif

struct fsb_sim;

// Grid-based Eulerian fluid simulation parameters (reference literals, include/fluid_sim.h:11-15)
const int GRID_SIZE_X = 200;
const int GRID_SIZE_Y = 200;
const float VISCOSITY = 0.0001f;
const float DIFFUSION = 0.0f;
const float PRESSURE = 0.5f;   // unused by the reference as well

class FluidSim
{
public:
    FluidSim();                                                  // 200 x 200, 20 / 5 sweeps, like the reference
    FluidSim(int width, int height, int projectionSweeps = 20, int diffusionSweeps = 5,
             float viscosity = VISCOSITY, float diffusion = DIFFUSION);
    FluidSim(const FluidSim &other);
    FluidSim(FluidSim &&other) noexcept;
    FluidSim &operator=(const FluidSim &other);
    FluidSim &operator=(FluidSim &&other) noexcept;
    ~FluidSim();

    void step(float dt);

    // Methods for interacting with the fluid
    void addDensity(int x, int y, float amount);
    void addVelocity(int x, int y, float amountX, float amountY);

    // Methods to access grid data for rendering
    float getDensity(int x, int y) const;
    glm::vec2 getVelocity(int x, int y) const;

    // Velocity visualization methods
    glm::vec2 getNormalizedVelocity(int x, int y) const;
    float getVelocityMagnitude(int x, int y) const;
    glm::vec3 getVelocityColor(int x, int y) const;

    int getWidth() const { return width; }
    int getHeight() const { return height; }

    // Bulk forms for callers that render from whole fields (width*height floats, y contiguous).
    void readDensity(float *host) const;
    void readVelocity(float *hostX, float *hostY) const;
    fsb_sim *handle() const { return sim; }

private:
    int width, height;
    int projSweeps, diffSweeps;
    float visc, diff;
    fsb_sim *sim;
    void create();
    void destroy();
    void copyStateFrom(const FluidSim &other);
};

#endif // FLUID_SIM_H
