import sys, os, time
sys.path.insert(0, os.getcwd())
import fluid_sim_b200 as fsb, numpy as np
from tests.util import vortex_lattice
for (W,H,P,n) in ((200,200,20,200),(2048,2048,20,20),(1024,8192,40,10),(32768,4096,100,3),(4096,4096,40,10),(16384,2048,20,5)):
    sim=fsb.FluidSim(W,H,P,5)
    d,u,v=vortex_lattice(W,H)
    sim.write(0,d);sim.write(1,u);sim.write(2,v)
    for _ in range(3): sim.step(0.01)
    sim.sync(); t=time.perf_counter()
    for _ in range(n): sim.step(0.01)
    sim.sync(); print(os.environ.get("FSB_CHAIN","2"), W,H,P, "ms/step %.3f" % ((time.perf_counter()-t)/n*1e3), flush=True)
    sim.close()
