"""Development aid: time the projection (Gauss-Seidel) and the diffusion (SOR) of one step for different
(strips per CTA, sweeps per warp, stage cap) settings of the sweep pipeline.  Prints ms per projection /
per velocity diffusion from the library's own CUDA-event profile."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fluid_sim_b200 as fsb


def run(W, H, P, D, strips, levels, cap, reps=3):
    sim = fsb.FluidSim(W, H, P, D, noise="zero")
    sim.set_tuning(strips, levels, cap)
    rng = np.random.default_rng(0)
    sim.write(fsb.VELOCITY_X, rng.standard_normal((W, H)).astype(np.float32))
    sim.write(fsb.VELOCITY_Y, rng.standard_normal((W, H)).astype(np.float32))
    try:
        sim.step(0.01); sim.sync()
        sim.profile_enable(True); sim.profile_read(reset=True)
        for _ in range(reps):
            sim.step(0.01)
        pr = sim.profile_read()
    except Exception as e:   # a setting that does not fit the SM
        print(f"W={W} H={H} P={P} D={D} strips={strips} levels={levels} cap={cap}: {e}", flush=True)
        sim.close()
        return
    gs_ms, gs_n = pr["gs"]
    sor_ms, sor_n = pr["sor"]
    print(f"W={W} H={H} P={P} D={D} strips={strips} levels={levels} cap={cap}: projection {gs_ms / (2 * reps):7.3f} ms "
          f"({gs_n // (2 * reps)} launch)  diffusion {sor_ms / reps:7.3f} ms ({sor_n // reps} launch)", flush=True)
    sim.close()


if __name__ == "__main__":
    W = int(os.environ.get("TW", "8192")); H = int(os.environ.get("TH", "8192"))
    P = int(os.environ.get("TP", "40")); D = int(os.environ.get("TD", "5"))
    cases = os.environ.get("TCASES", "2:10:0,2:5:0,4:5:0,8:5:0,16:5:0,1:2:0,2:2:0")
    for c in cases.split(","):
        strips, levels, cap = (int(v) for v in c.split(":"))
        run(W, H, P, D, strips, levels, cap)
