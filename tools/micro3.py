"""Development aid: time one projection for several (sweeps per warp, strips per CTA) configurations.
usage: micro3.py [W H P spw strips]...   (default: a sweep over 8192^2 P=40 and 2048^2 P=20)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fluid_sim_b200 as fsb

def run(W, H, P, spw, strips, reps=3):
    sim = fsb.FluidSim(W, H, P, 5, strips_per_cta=strips, sweeps_per_warp=spw, noise="zero")
    rng = np.random.default_rng(0)
    sim.write(fsb.VELOCITY_X, rng.standard_normal((W, H)).astype(np.float32))
    sim.write(fsb.VELOCITY_Y, rng.standard_normal((W, H)).astype(np.float32))
    sim.project(); sim.sync()
    sim.profile_enable(True); sim.profile_read(reset=True)
    for _ in range(reps):
        sim.project()
    pr = sim.profile_read()
    ms, n = pr["gs"]
    tot = ms / reps
    print(f"W={W:6d} H={H:6d} P={P:3d} S/warp={spw:2d} strips/cta={strips}: {tot:8.3f} ms/projection ({n//reps} launches) "
          f"{W*H*P/tot/1e6:8.1f} G sweep-cells/s", flush=True)
    sim.close()

if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    cfgs = [tuple(a[i:i + 5]) for i in range(0, len(a), 5)] if a else \
        [(W, H, P, spw, st) for (W, H, P) in [(8192, 8192, 40), (2048, 2048, 20)] for spw in (20, 10, 5) for st in (1, 2, 4)]
    for cfg in cfgs:
        try:
            run(*cfg)
        except Exception as e:
            print("FAIL", cfg, str(e)[:100], flush=True)
