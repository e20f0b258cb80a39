"""Timeline of the chained launches of one step (development aid): per phase, when the work items (blocks of columns,
left to right) waited, started and finished.  usage: chain_timeline.py [W H P]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import fluid_sim_b200 as fsb
from tests.util import vortex_lattice

W, H, P = (int(a) for a in (sys.argv[1:4] + ["8192", "8192", "40"][len(sys.argv) - 1:]))
sim = fsb.FluidSim(W, H, P, 5)
d, u, v = vortex_lattice(W, H)
sim.write(fsb.DENSITY, d); sim.write(fsb.VELOCITY_X, u); sim.write(fsb.VELOCITY_Y, v)
for _ in range(3):
    sim.step(0.01)
sim.sync()
sim.step(0.01)
for name, tr in (("diffuse + project", sim.chain_trace(0).astype(np.float64)), ("project alone", sim.chain_trace(1).astype(np.float64))):
    if not (tr > 0).any():
        continue
    t0 = tr[tr > 0].min()
    print(name, "phases", tr.shape[0])
    for p in range(tr.shape[0]):
        ran = np.nonzero(tr[p, :, 2] > 0)[0]
        if len(ran) == 0:
            continue
        sel = sorted(set([ran[0], ran[len(ran) // 4], ran[len(ran) // 2], ran[3 * len(ran) // 4], ran[-1]]))
        print(f" phase {p}: {len(ran)} items; work time mean {np.mean(tr[p, ran, 2] - tr[p, ran, 1]) / 1e6:.3f} ms, "
              f"wait mean {np.mean(tr[p, ran, 1] - tr[p, ran, 0]) / 1e6:.3f} ms, last item done at {(tr[p, ran, 2].max() - t0) / 1e6:.3f} ms")
        for i in sel:
            print(f"   item {i:4d}: wait from {(tr[p, i, 0] - t0) / 1e6:8.3f}  start {(tr[p, i, 1] - t0) / 1e6:8.3f}  end {(tr[p, i, 2] - t0) / 1e6:8.3f} ms")
sim.close()
