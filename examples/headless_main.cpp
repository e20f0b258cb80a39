// Headless counterpart of the reference's viewer loop (reference: src/main.cpp:7, :350-363, :369-396):
// the same global object, the same start-up injection, step(0.01f) per frame and a full per-cell
// density read-back, with the GLFW / OpenGL calls left out (no graphics stack on the GPU box).
// Compiles unchanged against the reference's header + src/fluid_sim.cpp or against this repository's
// include/fluid_sim.h + libfluidsim_dropin.so.
#include "fluid_sim.h"

#include <cstdio>
#include <cstdlib>

FluidSim sim;   // namespace-scope object, constructed before main (src/main.cpp:7)

int main(int argc, char **argv)
{
    const int frames = argc > 1 ? std::atoi(argv[1]) : 100;
    int centerX = sim.getWidth() / 2;
    int centerY = sim.getHeight() / 2;
    for (int i = -5; i <= 5; i++)
        for (int j = -5; j <= 5; j++)
            if (i * i + j * j < 25) {
                sim.addDensity(centerX + i, centerY + j, 10.0f);
                sim.addVelocity(centerX + i, centerY + j, 0.0f, 2.0f);
            }
    double sumD = 0.0, sumU = 0.0, sumV = 0.0;
    for (int f = 0; f < frames; ++f) {
        sim.step(0.01f);
        sumD = sumU = sumV = 0.0;
        for (int i = 0; i < sim.getWidth(); i++)
            for (int j = 0; j < sim.getHeight(); j++) {
                sumD += sim.getDensity(i, j);
                glm::vec2 v = sim.getVelocity(i, j);
                sumU += v.x;
                sumV += v.y;
            }
    }
    std::printf("frames %d  sum(d) %.8f  sum(u) %.8f  sum(v) %.8f  d[100][110] %.9g\n", frames, sumD, sumU, sumV,
                (double)sim.getDensity(100, 110));
    sim = FluidSim();   // reset idiom of src/main.cpp:289
    std::printf("after reset d[100][110] %.9g\n", (double)sim.getDensity(100, 110));
    return 0;
}
