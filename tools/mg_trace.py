"""Development aid: per-phase GPU time of the multi-GPU step on every rank (run under torchrun)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import fluid_sim_b200 as fsb
from fluid_sim_b200.multi import BlockFluidSim
from tests.util import vortex_lattice_block
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
W, P = 32768, 100
H = int(os.environ.get("MG_COLS", "4096")) * world
sim = BlockFluidSim(W, H, P, 5, device=local)
d, u, v = vortex_lattice_block(W, H, sim.jbase, sim.jbase + sim.pitch)
for fid, arr in ((fsb.DENSITY, d), (fsb.VELOCITY_X, u), (fsb.VELOCITY_Y, v)):
    sim._check(sim._lib.fsb_write_field(sim._h, fid, np.ascontiguousarray(arr).ctypes.data))
sim.step(0.01); sim.sync(); dist.barrier()
sim.trace = {}
steps = 2
for _ in range(steps):
    sim.step(0.01)
if os.environ.get("MG_TIMELINE"):
    # one more step, started from a barrier, as a timeline (device clocks differ; the barrier aligns them to ~10 us)
    sim.sync(); sim._ev = []
    with torch.cuda.stream(sim.stream):
        dist.barrier()
    sim.step(0.01)
    tl = sim.trace_timeline()
    for r in range(world):
        dist.barrier()
        if r == rank:
            print(f"timeline rank {rank}: " + " ".join(f"{n}@{t:.1f}" for n, t in tl), flush=True)
    sim._ev = []
    for _ in range(steps):
        sim.step(0.01)
rep = sim.trace_report()
tot = sum(rep.values())
for r in range(world):
    dist.barrier()
    if r == rank:
        print(f"rank {rank}: total {tot/steps:7.1f} ms/step  " + "  ".join(f"{k} {v/steps:6.1f}" for k, v in sorted(rep.items())), flush=True)
dist.destroy_process_group()
