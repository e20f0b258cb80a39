"""Development aid: time one fused Gauss-Seidel pass (P=20) on different grid shapes, to separate the
per-step cost of a strip from the cost of the CTA chain.  Prints ns per global time step."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fluid_sim_b200 as fsb

def run(W, H, P=20, nw=0, cap=0, reps=3):
    sim = fsb.FluidSim(W, H, P, 5, warps_per_cta=nw, max_sweeps_per_pass=cap, noise="zero")
    rng = np.random.default_rng(0)
    sim.write(fsb.VELOCITY_X, rng.standard_normal((W, H)).astype(np.float32))
    sim.write(fsb.VELOCITY_Y, rng.standard_normal((W, H)).astype(np.float32))
    sim.project(); sim.sync()
    sim.profile_enable(True); sim.profile_read(reset=True)
    for _ in range(reps):
        sim.project()
    pr = sim.profile_read()
    ms, n = pr["gs"]
    S = min(P, cap or 20)
    steps = W + H + S + 64
    per = ms / n
    print(f"W={W:6d} H={H:6d} P={P} nw={nw} cap={cap}: {per:8.3f} ms/pass  {per*1e6/steps:8.1f} ns/step  "
          f"({per*1e6/steps*1.965:7.0f} cyc/step)  {W*H*S/per/1e6:8.1f} G sweep-cells/s", flush=True)
    sim.close()

if __name__ == "__main__":
    for args in [(32768, 96), (32768, 224), (32768, 480), (16384, 2048), (8192, 8192), (2048, 2048)]:
        for nw in (1, 2, 4):
            run(*args, nw=nw)
    run(32768, 96, cap=10, nw=4)
    run(32768, 96, cap=5, nw=4)
    run(8192, 8192, cap=10, nw=4)
