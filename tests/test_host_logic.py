"""CPU tests of the host-side logic: the stream-position model of glibc rand(), the numpy model of the
strip-pipelined sweep (the algorithm the CUDA kernel implements), and the C ABI's symbol table."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import fluid_sim_b200 as fsb
from oracle import orasim
from oracle.orasim import OracleSim
from tests.util import ROOT, assert_bits, random_field

sys.path.insert(0, os.path.join(ROOT, "tools"))
import wavefront_emu as emu  # noqa: E402


def test_library_exports_every_declared_symbol():
    """The C-ABI library loads (no GPU needed) and exports everything include/fluid_sim_b200.h declares."""
    lib = fsb.load_library()
    hdr = open(os.path.join(ROOT, "include", "fluid_sim_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(fsb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    bound = {name for name, _, _ in fsb.ABI}
    assert declared == bound, (declared - bound, bound - declared)
    for name in declared:
        assert getattr(lib, name) is not None
    assert b"sm_100a" in lib.fsb_version()


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(fsb.FsbError, match="no CUDA device"):
        fsb.FluidSim()


def test_rand_jump_model():
    """glibc_rand.h (host side of the device noise kernel): sequential model and jump-ahead vs libc."""
    src = os.path.join(ROOT, "tests", "rand_jump_check.cpp")
    exe = "/tmp/fsb_rand_jump_check"
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "fluid_sim_b200", "csrc"), src, "-o", exe],
                   check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr


def test_noise_phase_stream_positions():
    """Noise as a phase of the chained launch (noise.cuh, chain.cu): every thread's schedule of polynomial jumps and
    its circular generation reproduce the step's rand() outputs at the cells of its row pieces."""
    src = os.path.join(ROOT, "tests", "noise_phase_check.cpp")
    exe = "/tmp/fsb_noise_phase_check"
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "fluid_sim_b200", "csrc"), src, "-o", exe],
                   check=True)
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr


@pytest.mark.parametrize("w,h,P", [(20, 24, 3), (40, 70, 20), (33, 31, 7)])
def test_wavefront_model_projection(w, h, P):
    rng = np.random.default_rng(w + h)
    o = OracleSim(w, h, P, 5)
    u, v = random_field(rng, w, h), random_field(rng, w, h)
    o.write(orasim.VEL_X, u); o.write(orasim.VEL_Y, v)
    o.project()
    for chunks in (None, [1] * P, [P - 2, 2]):
        ue, ve, pe = emu.project(u, v, P, chunks=chunks)
        assert_bits(ue, o.read(orasim.VEL_X), f"u chunks={chunks}")
        assert_bits(ve, o.read(orasim.VEL_Y), f"v chunks={chunks}")
        assert_bits(pe, o.read(orasim.PREV_VEL_X), f"p chunks={chunks}")


@pytest.mark.parametrize("b", [0, 1, 2])
def test_wavefront_model_diffusion(b):
    w, h, D = 37, 45, 5
    rng = np.random.default_rng(b)
    o = OracleSim(w, h, 20, D)
    x = random_field(rng, w, h)
    o.write(orasim.VEL_X, x); o.write(orasim.PREV_VEL_X, x)
    o.diffuse(b, orasim.VEL_X, orasim.PREV_VEL_X, 0.005, 0.01)
    for chunks in (None, [2, 2, 1]):
        xe = emu.diffuse(x, b, 0.005, 0.01, D, chunks=chunks)
        assert_bits(xe, o.read(orasim.VEL_X), f"b={b} chunks={chunks}")
