"""ms per step of back-to-back steps at 8192^2, P=40 and the per-phase means of the last launch (development aid; the
environment selects the variant: FSB_CHAIN_HELPERS, FSB_CHAIN_HELPER_SHARE, FSB_CHAIN_OVERLAP)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import fluid_sim_b200 as fsb
from tests.util import vortex_lattice

W, H, P = 8192, 8192, 40
n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
d, u, v = vortex_lattice(W, H)
sim = fsb.FluidSim(W, H, P, 5)
sim.write(fsb.DENSITY, d); sim.write(fsb.VELOCITY_X, u); sim.write(fsb.VELOCITY_Y, v)
for _ in range(4):
    sim.step(0.01)
sim.sync()
t = time.perf_counter()
for _ in range(n):
    sim.step(0.01)
sim.sync()
ms = (time.perf_counter() - t) / n * 1e3
tr = sim.chain_trace(0, max_items=288).astype(np.float64)
means = []
t0 = tr[tr > 0].min()
for p in range(tr.shape[0]):
    ran = np.nonzero(tr[p, :, 2] > 0)[0]
    if len(ran):
        means.append(f"{np.mean(tr[p, ran, 2] - tr[p, ran, 1]) / 1e6:.2f}(w{np.mean(tr[p, ran, 1] - tr[p, ran, 0]) / 1e6:.2f})")
if os.environ.get("FSB_QUICK_DETAIL"):
    for p in range(tr.shape[0]):
        ran = np.nonzero(tr[p, :, 2] > 0)[0]
        if not len(ran):
            continue
        sel = [ran[0], ran[1], ran[len(ran) // 2], ran[len(ran) // 2 + 1], ran[-2], ran[-1]]
        print(f" phase {p} ({len(ran)} items):", "; ".join(f"{i}: {(tr[p, i, 0] - t0) / 1e6:.2f} {(tr[p, i, 1] - t0) / 1e6:.2f} {(tr[p, i, 2] - t0) / 1e6:.2f}" for i in sel))
tag = " ".join(f"{k}={os.environ[k]}" for k in ("FSB_CHAIN_HELPERS", "FSB_CHAIN_HELPER_SHARE", "FSB_CHAIN_OVERLAP") if k in os.environ)
print(f"[{tag or 'default'}] {ms:.3f} ms/step over {n} steps; work per item by phase: {' '.join(means)}", flush=True)
sim.close()
