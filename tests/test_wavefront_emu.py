"""CPU tests of the numpy model of the sweep pipeline (tools/wavefront_emu.py): the skewed, time-pipelined update
order the CUDA kernel uses -- one column per lane as shipped, two columns per lane as planned -- gives exactly the
fields of the reference's lexicographic loops (C restatement, oracle/fluid_oracle.c)."""
import numpy as np
import pytest

from oracle import orasim
from oracle.orasim import OracleSim
from tools import wavefront_emu as emu

SWEEPS = {"one column per lane": emu.sweep_pass, "two columns per lane": emu.sweep_pass_2col}


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("name", list(SWEEPS))
@pytest.mark.parametrize("w,h,p,chunks", [(12, 17, 6, None), (9, 40, 7, [5, 2]), (21, 34, 20, [10, 10]), (8, 7, 3, None)])
def test_projection_order(name, w, h, p, chunks):
    rng = np.random.default_rng(w * h + p)
    u = rng.standard_normal((w, h)).astype(np.float32)
    v = rng.standard_normal((w, h)).astype(np.float32)
    ora = OracleSim(w, h, p, 5)
    ora.write(orasim.VEL_X, u); ora.write(orasim.VEL_Y, v)
    ora.project()
    gu, gv, _ = emu.project(u, v, p, chunks=chunks, sweep=SWEEPS[name])
    assert np.array_equal(bits(gu), bits(ora.read(orasim.VEL_X)))
    assert np.array_equal(bits(gv), bits(ora.read(orasim.VEL_Y)))


@pytest.mark.parametrize("name", list(SWEEPS))
@pytest.mark.parametrize("w,h,d,b", [(12, 17, 5, 1), (9, 40, 5, 2), (16, 33, 3, 0)])
def test_diffusion_order(name, w, h, d, b):
    rng = np.random.default_rng(w + h + d)
    x = rng.standard_normal((w, h)).astype(np.float32)
    ora = OracleSim(w, h, 20, d)
    ora.write(orasim.VEL_X, x)
    ora.write(orasim.PREV_VEL_X, x)
    ora.diffuse(b, orasim.VEL_X, orasim.PREV_VEL_X, 0.02, 0.01)
    got = emu.diffuse(x, b, 0.02, 0.01, d, sweep=SWEEPS[name])
    assert np.array_equal(bits(got), bits(ora.read(orasim.VEL_X)))
