"""Do consecutive steps overlap on the GPU (development aid)?  ms per step with and without the overlap, with and
without the per-launch profiling events, and the timeline of the last of several back-to-back launches.
usage: overlap_check.py [W H P]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import fluid_sim_b200 as fsb
from tests.util import vortex_lattice

W, H, P = (int(a) for a in (sys.argv[1:4] + ["8192", "8192", "40"][len(sys.argv) - 1:]))
d, u, v = vortex_lattice(W, H)


def run(overlap, profile, n=8):
    sim = fsb.FluidSim(W, H, P, 5, overlap=overlap)
    sim.write(fsb.DENSITY, d); sim.write(fsb.VELOCITY_X, u); sim.write(fsb.VELOCITY_Y, v)
    for _ in range(3):
        sim.step(0.01)
    sim.sync()
    if profile:
        sim.profile_enable(True)
    t = time.perf_counter()
    for _ in range(n):
        sim.step(0.01)
    t_sub = time.perf_counter() - t
    sim.sync()
    ms = (time.perf_counter() - t) / n * 1e3
    print(f"overlap {overlap} profiling {profile}: {ms:.3f} ms/step (host submitted {n} steps in {t_sub * 1e3:.2f} ms)", flush=True)
    return sim


run(False, False).close()
run(True, True).close()
sim = run(True, False)
tr = sim.chain_trace(0).astype(np.float64)
t0 = tr[tr > 0].min()
print("last launch: phases", tr.shape[0])
for p in range(tr.shape[0]):
    ran = np.nonzero(tr[p, :, 2] > 0)[0]
    if len(ran) == 0:
        continue
    print(f" phase {p}: {len(ran)} items; work mean {np.mean(tr[p, ran, 2] - tr[p, ran, 1]) / 1e6:.3f} ms, wait mean "
          f"{np.mean(tr[p, ran, 1] - tr[p, ran, 0]) / 1e6:.3f} ms; first / last item done {(tr[p, ran[0], 2] - t0) / 1e6:.3f} / {(tr[p, ran, 2].max() - t0) / 1e6:.3f} ms")
    if p == 0:
        for i in sorted(set([ran[0], ran[10], ran[19], ran[20], ran[len(ran) // 2], ran[-1]])):
            print(f"   item {i:4d}: CTA in at {(tr[p, i, 0] - t0) / 1e6:8.3f}  work from {(tr[p, i, 1] - t0) / 1e6:8.3f} to {(tr[p, i, 2] - t0) / 1e6:8.3f} ms")
sim.close()
