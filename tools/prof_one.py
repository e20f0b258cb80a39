"""Development aid: run a few projections on one grid shape (for ncu captures).
usage: prof_one.py W H P sweeps_per_warp strips_per_cta"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fluid_sim_b200 as fsb
W, H, P, spw, strips = (int(a) for a in sys.argv[1:6])
sim = fsb.FluidSim(W, H, P, 5, sweeps_per_warp=spw, strips_per_cta=strips, noise="zero")
rng = np.random.default_rng(0)
sim.write(fsb.VELOCITY_X, rng.standard_normal((W, H)).astype(np.float32))
sim.write(fsb.VELOCITY_Y, rng.standard_normal((W, H)).astype(np.float32))
for _ in range(3):
    sim.project()
sim.sync()
print("done")
sim.close()
