"""GPU test of fsb_pressure_sweeps_noalloc, the entry point for hosts that cannot let the library allocate
(engines that gave all GPU memory to their own pools at start-up, loops that must never meet the device-wide
synchronisation of cudaMalloc / cudaFree).  It runs the projection's pressure sweeps on buffers the caller owns
and keeps its inter-CTA scratch in a fixed array inside the library image.

Checked here, through the C ABI: the result equals, bit for bit, the pressure the ordinary entry point
(fsb_op_project) leaves in FSB_PREV_VEL_X for the same divergence - on two calls in a row, because every call
reuses the same scratch array and must cope with what the previous call left in it - and requests beyond the
fixed array are refused with a clear message.  The file sorts last so that a run that stops at the first
failure has already executed every other test.
"""
import ctypes as C

import numpy as np
import pytest

import fluid_sim_b200 as fsb
from tests.util import assert_bits, random_field

pytestmark = pytest.mark.gpu


def _reference_pressure(w, h, sweeps, seed):
    """divergence (device pointer kept alive by the returned sim) and pressure from the ordinary path"""
    sim = fsb.FluidSim(w, h, sweeps, 5, noise="zero")
    rng = np.random.default_rng(seed)
    sim.write(fsb.VELOCITY_X, random_field(rng, w, h))
    sim.write(fsb.VELOCITY_Y, random_field(rng, w, h))
    sim.project()
    sim.sync()
    return sim, sim.device_field(fsb.PREV_VEL_Y), sim.read(fsb.PREV_VEL_X)


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
