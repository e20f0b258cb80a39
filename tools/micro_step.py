"""Development aid: per-kernel-class times of a full step for given tuning.
usage: micro_step.py W H P spw strips [steps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fluid_sim_b200 as fsb
from tests.util import vortex_lattice
W, H, P, spw, strips = (int(a) for a in sys.argv[1:6])
steps = int(sys.argv[6]) if len(sys.argv) > 6 else 3
sim = fsb.FluidSim(W, H, P, 5, sweeps_per_warp=spw, strips_per_cta=strips)
d, u, v = vortex_lattice(W, H)
sim.write(0, d); sim.write(1, u); sim.write(2, v)
sim.step(0.01); sim.sync()
sim.profile_enable(True); sim.profile_read(reset=True)
for _ in range(steps):
    sim.step(0.01)
pr = sim.profile_read()
print(f"W={W} H={H} P={P} spw={spw} strips={strips}: " + " ".join(f"{k}={v[0]/steps:.2f}" for k, v in pr.items()) +
      f" total={sum(v[0] for v in pr.values())/steps:.2f} ms/step", flush=True)
