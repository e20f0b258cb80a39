// The reference viewer's frame, without a window: run the demo for a number of frames and write what the viewer would
// show (density quads, optionally the velocity vectors: src/main.cpp:388-514) as a binary PPM.  Uses the drop-in class
// for the simulation and the C ABI underneath it (FluidSim::handle()) for the image.
//   usage: headless_view [frames=100] [out.ppm=frame.ppm] [vectors=0] [width=800] [height=600]
#include "fluid_sim.h"
#include "fluid_sim_b200.h"

#include <cstdio>
#include <cstdlib>
#include <vector>

int main(int argc, char **argv)
{
    const int frames = argc > 1 ? std::atoi(argv[1]) : 100;
    const char *path = argc > 2 ? argv[2] : "frame.ppm";
    const bool vectors = argc > 3 && std::atoi(argv[3]) != 0;
    const int pw = argc > 4 ? std::atoi(argv[4]) : 800, ph = argc > 5 ? std::atoi(argv[5]) : 600;   // src/main.cpp:9-10
    FluidSim sim;
    const int cx = sim.getWidth() / 2, cy = sim.getHeight() / 2;
    for (int i = -5; i <= 5; i++)                                  // start-up injection, src/main.cpp:351-363
        for (int j = -5; j <= 5; j++)
            if (i * i + j * j < 25) {
                sim.addDensity(cx + i, cy + j, 10.0f);
                sim.addVelocity(cx + i, cy + j, 0.0f, 2.0f);
            }
    for (int f = 0; f < frames; ++f) sim.step(0.01f);
    std::vector<unsigned char> rgba((size_t)pw * ph * 4);
    const int rc = vectors ? fsb_view_velocity(sim.handle(), pw, ph, rgba.data())
                           : fsb_view_density(sim.handle(), pw, ph, rgba.data());
    if (rc != 0) {
        std::fprintf(stderr, "view failed: %s\n", fsb_last_error());
        return 1;
    }
    std::FILE *out = std::fopen(path, "wb");
    if (!out) {
        std::perror(path);
        return 1;
    }
    std::fprintf(out, "P6\n%d %d\n255\n", pw, ph);
    for (int y = ph - 1; y >= 0; --y)                              // image row 0 is the bottom window row
        for (int x = 0; x < pw; ++x) std::fwrite(&rgba[((size_t)y * pw + x) * 4], 1, 3, out);
    std::fclose(out);
    std::printf("%s: %d x %d after %d frames%s\n", path, pw, ph, frames, vectors ? ", velocity vectors on" : "");
    return 0;
}
