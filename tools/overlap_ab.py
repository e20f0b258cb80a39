"""ms per step with overlapping steps (the default) and without, on several grid shapes (development aid)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fluid_sim_b200 as fsb
from tests.util import vortex_lattice

for (W, H, P, n) in ((200, 200, 20, 400), (512, 512, 20, 100), (2048, 2048, 20, 30), (1024, 8192, 40, 20), (4096, 4096, 40, 20),
                     (8192, 8192, 40, 20), (32768, 4096, 100, 5), (16384, 2048, 20, 6)):
    d, u, v = vortex_lattice(W, H)
    res = []
    for overlap in (False, True):
        sim = fsb.FluidSim(W, H, P, 5, overlap=overlap)
        sim.write(0, d); sim.write(1, u); sim.write(2, v)
        for _ in range(3):
            sim.step(0.01)
        sim.sync()
        n0 = sim.launch_count()
        t = time.perf_counter()
        for _ in range(n):
            sim.step(0.01)
        sim.sync()
        res.append(((time.perf_counter() - t) / n * 1e3, (sim.launch_count() - n0) / n))
        sim.close()
    print(f"{W} x {H} P={P}: {res[0][0]:.3f} ms/step ({res[0][1]:g} launches) -> overlapping {res[1][0]:.3f} ms/step ({res[1][1]:g} launches)", flush=True)
