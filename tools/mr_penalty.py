"""Cost of the multi-rank instantiation of the chained launch, measured on ONE GPU: a column block that spans the whole
grid (world size 1, no neighbours) runs chain_kernel<..., MR=true>, the plain simulation chain_kernel<..., MR=false>.
usage: python -m torch.distributed.run --nproc-per-node 1 --master-addr 127.0.0.1 tools/mr_penalty.py [W H P]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import fluid_sim_b200 as fsb
from fluid_sim_b200.multi import BlockFluidSim
from tests.util import vortex_lattice

W, H, P = (int(a) for a in (sys.argv[1:4] + ["8192", "8192", "40"][len(sys.argv) - 1:]))
torch.cuda.set_device(0)
dist.init_process_group("nccl", device_id=torch.device("cuda", 0))
d, u, v = vortex_lattice(W, H)
for name, sim in (("block (MR kernel)", BlockFluidSim(W, H, P, 5, device=0, chain=True)), ("plain", fsb.FluidSim(W, H, P, 5, chain=3))):
    if isinstance(sim, BlockFluidSim):
        for f, arr in ((fsb.DENSITY, d), (fsb.VELOCITY_X, u), (fsb.VELOCITY_Y, v)):
            sim.set_from_global(f, arr)
        print("chained:", sim.chained)
    else:
        sim.write(fsb.DENSITY, d); sim.write(fsb.VELOCITY_X, u); sim.write(fsb.VELOCITY_Y, v)
    for _ in range(3):
        sim.step(0.01)
    sim.sync()
    import ctypes as C
    sim._check(sim._lib.fsb_profile_enable(sim._h, 1))
    t = time.perf_counter()
    for _ in range(5):
        sim.step(0.01)
    sim.sync()
    names = fsb.FluidSim.PROF_NAMES
    ms = (C.c_double * len(names))(); cnt = (C.c_uint64 * len(names))()
    sim._check(sim._lib.fsb_profile_read(sim._h, ms, cnt, 1))
    print(name, W, H, P, "ms/step %.3f" % ((time.perf_counter() - t) / 5 * 1e3),
          {n: round(ms[i] / 5, 3) for i, n in enumerate(names) if ms[i] > 0}, flush=True)
    sim.close()
dist.destroy_process_group()
