"""Multi-GPU parity check, run under torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/mgpu_check.py [W H P D steps] [visc=..] [diff=..] [halo=..] [demo=1]

Every rank steps its column block (fluid_sim_b200.multi.BlockFluidSim); rank 0 also steps the same problem on one
GPU and compares the gathered fields bit for bit.  Prints MGPU-PARITY PASS / FAIL.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fluid_sim_b200 as fsb  # noqa: E402
from fluid_sim_b200.multi import BlockFluidSim  # noqa: E402
from tests.util import vortex_lattice  # noqa: E402


def main():
    pos = [v for v in sys.argv[1:] if "=" not in v]
    opt = dict(v.split("=", 1) for v in sys.argv[1:] if "=" in v)
    a = [int(v) for v in pos]
    W, H, P, D, steps = (a + [256, 512, 20, 5, 2][len(a):])[:5]
    visc, diff = float(opt.get("visc", 1e-4)), float(opt.get("diff", 0.0))
    halo = int(opt.get("halo", 128))
    demo = int(opt.get("demo", 0))           # 1: the reference's start-up injection (|v| = 2, src/main.cpp:351-360)
    chain = int(opt.get("chain", 1))         # 0: pass-by-pass launches with NCCL halo exchanges
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    d, u, v = vortex_lattice(W, H)
    rng = np.random.default_rng(7)
    u = (u + rng.standard_normal((W, H)).astype(np.float32) * 0.05).astype(np.float32)
    if demo:
        d, u, v = np.zeros_like(d), np.zeros_like(u), np.zeros_like(v)
        for i in range(-5, 6):
            for j in range(-5, 6):
                if i * i + j * j < 25:
                    d[W // 2 + i, H // 2 + j] += np.float32(10.0)
                    v[W // 2 + i, H // 2 + j] += np.float32(2.0)
    sim = BlockFluidSim(W, H, P, D, viscosity=visc, diffusion=diff, device=local, halo=halo, chain=bool(chain))
    for f, arr in ((fsb.DENSITY, d), (fsb.VELOCITY_X, u), (fsb.VELOCITY_Y, v)):
        sim.set_from_global(f, arr)
    reaches = []
    for _ in range(steps):
        sim.step(0.01)
        reaches.append("chained" if sim.chained else sim.reach)
    sim.sync()
    got = [sim.gather(f) for f in (fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y)]
    ok = True
    if rank == 0:
        ref = fsb.FluidSim(W, H, P, D, viscosity=visc, diffusion=diff, device=local)
        ref.write(fsb.DENSITY, d); ref.write(fsb.VELOCITY_X, u); ref.write(fsb.VELOCITY_Y, v)
        for _ in range(steps):
            ref.step(0.01)
        want = ref.fields()
        for name, g, w_ in zip(("density", "velX", "velY"), got, want):
            same = np.array_equal(g.view(np.uint32), w_.view(np.uint32))
            if not same:
                bad = np.argwhere(g.view(np.uint32) != w_.view(np.uint32))
                print(f"{name}: {len(bad)} cells differ, first {tuple(bad[0])}, cols {bad[:,1].min()}..{bad[:,1].max()} "
                      f"rows {bad[:,0].min()}..{bad[:,0].max()}", flush=True)
            ok = ok and same
        print(f"MGPU-PARITY {'PASS' if ok else 'FAIL'} world={dist.get_world_size()} grid={W}x{H} P={P} D={D} steps={steps} "
              f"visc={visc} diff={diff} halo={halo} demo={demo} exchanged columns per step={reaches}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
