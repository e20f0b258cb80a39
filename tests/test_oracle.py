"""CPU tests: the oracle (oracle/fluid_oracle.c) against the reference's own outputs.

The reference ships no tests; its outputs are pinned as golden files generated
from the unmodified source (tests/golden/make_golden.py) and, when oracle/_ref is present, by running
the real reference side by side.
"""
import ctypes
import os

import numpy as np
import pytest

from oracle import orasim, refsim
from oracle.orasim import OracleSim
from tests.util import apply_golden_inputs, assert_bits, load_golden, random_field

GOLDENS = ["config1_200x200_p20_d5", "small_48x40_p3_d2", "odd_257x131_p20_d5", "tall_131x257_p7_d3"]


@pytest.mark.parametrize("name", GOLDENS)
def test_oracle_matches_golden(name):
    g = load_golden(name)
    w, h, p, d = int(g["w"]), int(g["h"]), int(g["p"]), int(g["d"])
    sim = OracleSim(w, h, p, d)
    apply_golden_inputs(sim, g)
    checkpoints = sorted(int(k[2:]) for k in g.files if k.startswith("d_"))
    step = 0
    for cp in checkpoints:
        while step < cp:
            sim.step(0.01)
            step += 1
        dd, uu, vv = sim.fields()
        assert_bits(dd, g[f"d_{cp}"], f"{name} density step {cp}")
        assert_bits(uu, g[f"u_{cp}"], f"{name} velX step {cp}")
        assert_bits(vv, g[f"v_{cp}"], f"{name} velY step {cp}")


def test_known_answers_config1():
    """values obtained from the verbatim reference, 100 steps of the demo run."""
    g = load_golden("config1_200x200_p20_d5")
    d, u, v = g["d_100"], g["u_100"], g["v_100"]
    assert abs(d.astype(np.float64).sum() - 1074.24332) < 1e-4
    assert abs(u.astype(np.float64).sum() - 1.11766667) < 1e-6
    assert abs(v.astype(np.float64).sum() - 111.47682) < 1e-4
    assert d[100, 100] == np.float32(4.46316655e-08)
    assert d[100, 110] == np.float32(0.000305137684)
    assert u[90, 100] == np.float32(0.0174698886)
    assert v[100, 100] == np.float32(0.0318563022)
    assert d[0, 0] == 0


def test_rand_restatement_matches_libc():
    """The oracle's glibc generator against this machine's libc rand() (default seed)."""
    libc = ctypes.CDLL(None)
    libc.srand(1)
    sim = OracleSim(8, 8)
    for _ in range(5000):
        assert sim.rand() == libc.rand()


def test_rand_first_outputs():
    sim = OracleSim(8, 8)
    assert [sim.rand() for _ in range(6)] == [1804289383, 846930886, 1681692777, 1714636915, 1957747793, 424238335]


@pytest.mark.parametrize("variant", [(48, 40, 3, 2), (64, 64, 20, 5), (257, 131, 20, 5), (96, 1100, 20, 5)])
def test_oracle_vs_real_reference(variant):
    w, h, p, d = variant
    if not refsim.available(w, h, p, d):
        pytest.skip("oracle/_ref not built for this variant")
    rng = np.random.default_rng(w * 7 + h)
    ref = refsim.RefSim(w, h, p, d, seed=1)
    ora = OracleSim(w, h, p, d)
    for fid in (0, 1, 2):
        f = random_field(rng, w, h, 0.5)
        ref.write(fid, f)
        ora.write(fid, f)
    for _ in range(3):
        ref.step(0.01)
        ora.step(0.01)
    for a, b, n in zip(ora.fields(), ref.fields(), "duv"):
        assert_bits(a, b, f"{variant} field {n}")


@pytest.mark.parametrize("scheme", [1, 2])
def test_oracle_alt_schemes_vs_real_reference(scheme):
    """MacCormack / semi-Lagrangian bodies (unreachable in the reference, src/fluid_sim.cpp:223,227)."""
    w, h, p, d = 64, 64, 20, 5
    if not refsim.available(w, h, p, d):
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(scheme)
    ref = refsim.RefSim(w, h, p, d)
    ora = OracleSim(w, h, p, d, scheme=scheme)
    for fid in (0, 1, 2):
        f = random_field(rng, w, h, 2.0)
        ref.write(fid, f)
        ora.write(fid, f)
    ref.advect(0, refsim.PREV_DENSITY, refsim.DENSITY, refsim.VEL_X, refsim.VEL_Y, 0.01, scheme=scheme)
    ora.advect(0, orasim.PREV_DENSITY, orasim.DENSITY, orasim.VEL_X, orasim.VEL_Y, 0.01)
    assert_bits(ora.read(orasim.PREV_DENSITY), ref.read(refsim.PREV_DENSITY), f"scheme {scheme}")


def test_zero_noise_mode_matches_reference():
    w, h, p, d = 48, 40, 3, 2
    if not refsim.available(w, h, p, d):
        pytest.skip("oracle/_ref not built")
    ref = refsim.RefSim(w, h, p, d, noise="zero")
    ora = OracleSim(w, h, p, d, noise="zero")
    ref.addVelocity(10, 10, 1.0, -2.0); ora.addVelocity(10, 10, 1.0, -2.0)
    ref.addDensity(12, 9, 3.0); ora.addDensity(12, 9, 3.0)
    for _ in range(5):
        ref.step(0.01); ora.step(0.01)
    for a, b, n in zip(ora.fields(), ref.fields(), "duv"):
        assert_bits(a, b, f"zero-noise field {n}")
    ref.lib.ref_noise_mode(0)


def test_getters_match_reference():
    if not refsim.available(200, 200, 20, 5):
        pytest.skip("oracle/_ref not built")
    ref = refsim.RefSim()
    ora = OracleSim()
    for s in (ref, ora):
        s.addVelocity(5, 6, 3.0, 4.0); s.addVelocity(7, 8, 0.0005, 0.0); s.addVelocity(9, 9, 30.0, 40.0)
        s.addVelocity(1, 1, 2.0, 1.0); s.addVelocity(2, 2, 5.5, 2.0)
    for (x, y) in [(5, 6), (7, 8), (9, 9), (1, 1), (2, 2), (0, 0), (-1, 3), (200, 0)]:
        assert ora.getVelocityMagnitude(x, y) == ref.getVelocityMagnitude(x, y)
        assert ora.getNormalizedVelocity(x, y) == ref.getNormalizedVelocity(x, y)
        assert ora.getVelocityColor(x, y) == ref.getVelocityColor(x, y)
