"""ms per step on several grid shapes (development aid); the environment selects the variant (FSB_CHAIN_OVERLAP,
FSB_CHAIN_HELPERS, ...), so run it once per variant."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fluid_sim_b200 as fsb
from tests.util import vortex_lattice

tag = " ".join(f"{k}={os.environ[k]}" for k in ("FSB_CHAIN_HELPERS", "FSB_CHAIN_HELPER_SHARE", "FSB_CHAIN_OVERLAP") if k in os.environ)
for (W, H, P, n) in ((200, 200, 20, 400), (512, 512, 20, 100), (2048, 2048, 20, 30), (1024, 8192, 40, 20), (4096, 4096, 40, 20),
                     (8192, 8192, 40, 20), (32768, 4096, 100, 5), (16384, 2048, 20, 6)):
    d, u, v = vortex_lattice(W, H)
    sim = fsb.FluidSim(W, H, P, 5)
    sim.write(0, d); sim.write(1, u); sim.write(2, v)
    for _ in range(3):
        sim.step(0.01)
    sim.sync()
    n0 = sim.launch_count()
    t = time.perf_counter()
    for _ in range(n):
        sim.step(0.01)
    sim.sync()
    print(f"[{tag or 'default'}] {W} x {H} P={P}: {(time.perf_counter() - t) / n * 1e3:.3f} ms/step, {(sim.launch_count() - n0) / n:g} launches per step", flush=True)
    sim.close()
