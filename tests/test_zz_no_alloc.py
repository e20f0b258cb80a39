"""GPU test of fsb_pressure_sweeps_noalloc, the entry point for hosts that cannot let the library allocate
(engines that gave all GPU memory to their own pools at start-up, loops that must never meet the device-wide
synchronisation of cudaMalloc / cudaFree).  It runs the projection's pressure sweeps on buffers the caller owns
and keeps its inter-CTA scratch in a fixed array inside the library image.

Checked here, through the C ABI: the result equals, bit for bit, the pressure the ordinary entry point
(fsb_op_project) leaves in FSB_PREV_VEL_X for the same divergence, and the pressure the CPU oracle's project()
computes - on two calls in a row, because every call reuses the same scratch array and must cope with what
the previous call left in it; calls of different shapes can alternate on one device; a warmed-up call leaves the
amount of free device memory unchanged; and requests beyond the fixed array are refused with a clear message.
The file sorts last so that a run that stops at the first failure has already executed every other test.
"""
import ctypes as C

import numpy as np
import pytest

import fluid_sim_b200 as fsb
from oracle.orasim import OracleSim
from oracle import orasim
from tests.util import assert_bits, random_field

pytestmark = pytest.mark.gpu


def _reference_pressure(w, h, sweeps, seed):
    """divergence (device pointer kept alive by the returned sim) and pressure from the ordinary path"""
    sim = fsb.FluidSim(w, h, sweeps, 5, noise="zero")
    rng = np.random.default_rng(seed)
    sim.write(fsb.VELOCITY_X, random_field(rng, w, h))
    sim.write(fsb.VELOCITY_Y, random_field(rng, w, h))
    u, v = sim.read(fsb.VELOCITY_X), sim.read(fsb.VELOCITY_Y)
    sim.project()
    sim.sync()
    want = sim.read(fsb.PREV_VEL_X)
    # the same projection on the CPU oracle (src/fluid_sim.cpp:301-344): its pressure is the second, independent check
    ora = OracleSim(w, h, sweeps, 5, noise="zero")
    ora.write(orasim.VEL_X, u)
    ora.write(orasim.VEL_Y, v)
    ora.project()
    assert_bits(want, ora.read(orasim.PREV_VEL_X), "fsb_op_project pressure against the oracle")
    ora.close()
    return sim, sim.device_field(fsb.PREV_VEL_Y), want


@pytest.mark.parametrize("w,h,sweeps", [(96, 200, 20), (300, 1100, 40), (64, 80, 7)])
def test_noalloc_matches_project_twice(w, h, sweeps):
    lib = fsb.load_library()
    holder = fsb.FluidSim(w, h, noise="zero")            # only lends a caller-owned device buffer for the output
    out_ptr = holder.device_field(fsb.DENSITY)
    for call in range(2):                                 # two calls in a row share the built-in scratch
        src, div_ptr, want = _reference_pressure(w, h, sweeps, seed=10 * call + w)
        holder.write(fsb.DENSITY, np.full((w, h), np.float32(123.0)))   # stale content must be overwritten
        rc = lib.fsb_pressure_sweeps_noalloc(C.c_void_p(div_ptr), C.c_void_p(out_ptr), w, h, sweeps, None)
        assert rc == 0, lib.fsb_last_error().decode()
        got = holder.read(fsb.DENSITY)                    # synchronises the holder's stream ...
        import torch
        torch.cuda.synchronize()                          # ... and the default stream the sweeps were queued on
        got = holder.read(fsb.DENSITY)
        assert_bits(got, want, f"no-alloc pressure, call {call}")
        src.close()
    holder.close()


def _run_noalloc(lib, holder, w, h, sweeps, seed):
    import torch
    src, div_ptr, want = _reference_pressure(w, h, sweeps, seed)
    out_ptr = holder.device_field(fsb.DENSITY)
    holder.write(fsb.DENSITY, np.full((holder.width, holder.height), np.float32(-7.0)))
    rc = lib.fsb_pressure_sweeps_noalloc(C.c_void_p(div_ptr), C.c_void_p(out_ptr), w, h, sweeps, None)
    assert rc == 0, lib.fsb_last_error().decode()
    torch.cuda.synchronize()
    got = holder.read(fsb.DENSITY).reshape(-1)[: w * h].reshape(w, h)
    src.close()
    return got, want


def test_noalloc_alternating_shapes_share_the_array():
    """Shapes A, B, A, B on one device: every shape lays its rings out differently inside the one array, so a change
    of shape must not let a call pick up ring words an earlier shape left there."""
    lib = fsb.load_library()
    holder = fsb.FluidSim(128, 640, noise="zero")        # large enough for both shapes' output
    shapes = [(128, 640, 40), (90, 333, 13), (128, 640, 40), (90, 333, 13)]
    for k, (w, h, sweeps) in enumerate(shapes):
        got, want = _run_noalloc(lib, holder, w, h, sweeps, seed=100 + k)
        assert_bits(got, want, f"no-alloc pressure, shape {k} {w}x{h} {sweeps} sweeps")
    holder.close()


def test_noalloc_warm_call_reserves_no_memory():
    """After one warm-up call per sweep count, a call must leave cudaMemGetInfo exactly where it was: neither this
    library nor (any longer) CUDA's lazy module loading takes device memory."""
    import torch
    lib = fsb.load_library()
    w, h, sweeps = 200, 512, 20
    holder = fsb.FluidSim(w, h, noise="zero")
    _run_noalloc(lib, holder, w, h, sweeps, seed=5)       # warm-up: array + kernel variants reach the device
    src, div_ptr, want = _reference_pressure(w, h, sweeps, seed=6)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        rc = lib.fsb_pressure_sweeps_noalloc(C.c_void_p(div_ptr), C.c_void_p(holder.device_field(fsb.DENSITY)), w, h,
                                             sweeps, None)
        assert rc == 0, lib.fsb_last_error().decode()
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before
    assert_bits(holder.read(fsb.DENSITY), want, "no-alloc pressure after warm-up")
    src.close()
    holder.close()


def test_noalloc_rejects_what_does_not_fit():
    lib = fsb.load_library()
    w, h, sweeps = 16, 4500, 80                           # 80 sweeps over 4500 columns: rings larger than the fixed array
    a = fsb.FluidSim(w, h, noise="zero")
    rc = lib.fsb_pressure_sweeps_noalloc(C.c_void_p(a.device_field(fsb.PREV_VEL_Y)),
                                         C.c_void_p(a.device_field(fsb.DENSITY)), w, h, sweeps, None)
    assert rc == -4                                       # FSB_ELIMIT
    assert b"built-in scratch" in lib.fsb_last_error()
    rc = lib.fsb_pressure_sweeps_noalloc(None, C.c_void_p(a.device_field(fsb.DENSITY)), w, h, sweeps, None)
    assert rc == -1                                       # FSB_EINVAL
    a.close()
