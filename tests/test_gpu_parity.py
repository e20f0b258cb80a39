"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle on the same
seeded inputs.  The bar is bit-exact (float32 fields compared as uint32); north_star's tolerance of
1e-5 relative max-abs per field is therefore met with zero error.
"""
import numpy as np
import pytest

import fluid_sim_b200 as fsb
from oracle import orasim, refsim
from oracle.orasim import OracleSim
from tests.util import (apply_golden_inputs, assert_bits, demo_injection, load_golden, random_field, rel_max_abs,
                        vortex_lattice)

pytestmark = pytest.mark.gpu

TOL = 1e-5  # north_star: relative max-abs per field (expected: exactly 0)

SIZES = [(16, 16), (48, 40), (33, 70), (131, 257), (200, 200), (96, 1100)]


def make_pair(w, h, p=20, d=5, **kw):
    gpu = fsb.FluidSim(w, h, p, d, **kw)
    okw = {}
    for k in ("viscosity", "diffusion", "omega"):
        if k in kw:
            okw[{"viscosity": "visc", "diffusion": "diff", "omega": "omega"}[k]] = kw[k]
    if "advect_scheme" in kw:
        okw["scheme"] = kw["advect_scheme"]
    if "noise" in kw:
        okw["noise"] = kw["noise"]
    ora = OracleSim(w, h, p, d, **okw)
    return gpu, ora


def seed_state(gpu, ora, rng, w, h, scale=0.5):
    for fid in (0, 1, 2):
        f = random_field(rng, w, h, scale)
        if fid == 0:
            f = np.abs(f)
        gpu.write(fid, f)
        ora.write(fid, f)


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("b", [0, 1, 2])
def test_set_boundary(w, h, b):
    gpu, ora = make_pair(w, h)
    rng = np.random.default_rng(b + w)
    f = random_field(rng, w, h)
    gpu.write(fsb.VELOCITY_X, f); ora.write(orasim.VEL_X, f)
    gpu.setBoundary(b, fsb.VELOCITY_X); ora.setBoundary(b, orasim.VEL_X)
    assert_bits(gpu.read(fsb.VELOCITY_X), ora.read(orasim.VEL_X), f"setBoundary b={b}")


@pytest.mark.parametrize("w,h", SIZES + [(300, 300)])
def test_noise_matches_glibc_stream(w, h):
    gpu, ora = make_pair(w, h)
    rng = np.random.default_rng(w)
    seed_state(gpu, ora, rng, w, h)
    for _ in range(3):   # three steps' worth: checks the position bookkeeping between steps
        gpu.addNoise(); ora.addNoise()
    assert gpu.noise_pos() == 3 * 2 * w * h
    assert_bits(gpu.read(fsb.VELOCITY_X), ora.read(orasim.VEL_X), "noise velX")
    assert_bits(gpu.read(fsb.VELOCITY_Y), ora.read(orasim.VEL_Y), "noise velY")


def test_noise_at_arbitrary_stream_position():
    w, h = 64, 48
    pos = 1234567
    gpu = fsb.FluidSim(w, h, noise_pos=pos)
    ora = OracleSim(w, h, skip=pos)
    gpu.addNoise(); ora.addNoise()
    assert_bits(gpu.read(fsb.VELOCITY_X), ora.read(orasim.VEL_X), "velX")
    assert_bits(gpu.read(fsb.VELOCITY_Y), ora.read(orasim.VEL_Y), "velY")


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("b", [0, 1, 2])
def test_diffuse(w, h, b):
    gpu, ora = make_pair(w, h)
    rng = np.random.default_rng(10 * b + h)
    f = random_field(rng, w, h)
    gpu.write(fsb.VELOCITY_X, f); ora.write(orasim.VEL_X, f); ora.write(orasim.PREV_VEL_X, f)
    diff = 0.0001 * 30
    gpu.diffuse(b, fsb.VELOCITY_X, diff, 0.01)
    ora.diffuse(b, orasim.VEL_X, orasim.PREV_VEL_X, diff, 0.01)
    assert_bits(gpu.read(fsb.VELOCITY_X), ora.read(orasim.VEL_X), f"diffuse b={b}")


@pytest.mark.parametrize("D,spw,stg", [(1, 0, 0), (3, 0, 0), (7, 0, 0), (5, 2, 0), (12, 5, 1), (5, 1, 0), (10, 5, 0), (0, 0, 0)])
def test_diffuse_sweep_counts(D, spw, stg):
    """Sweep counts split over several stages and / or several launches (the source field must stay the
    pre-diffusion copy across launches)."""
    w, h = 70, 90
    gpu = fsb.FluidSim(w, h, 20, D, sweeps_per_warp=spw, max_stages=stg)
    ora = OracleSim(w, h, 20, D)
    rng = np.random.default_rng(D)
    f = random_field(rng, w, h)
    gpu.write(fsb.VELOCITY_Y, f); ora.write(orasim.VEL_Y, f); ora.write(orasim.PREV_VEL_Y, f)
    gpu.diffuse(2, fsb.VELOCITY_Y, 0.002, 0.01)
    ora.diffuse(2, orasim.VEL_Y, orasim.PREV_VEL_Y, 0.002, 0.01)
    assert_bits(gpu.read(fsb.VELOCITY_Y), ora.read(orasim.VEL_Y), f"diffuse D={D}")


def test_diffuse_zero_coefficient():
    """DIFFUSION == 0 (include/fluid_sim.h:14): identity on the interior plus the boundary update."""
    w, h = 50, 64
    gpu, ora = make_pair(w, h)
    rng = np.random.default_rng(5)
    f = random_field(rng, w, h)
    f[3, 4] = -0.0
    gpu.write(fsb.DENSITY, f); ora.write(orasim.DENSITY, f); ora.write(orasim.PREV_DENSITY, f)
    gpu.diffuse(0, fsb.DENSITY, 0.0, 0.01)
    ora.diffuse(0, orasim.DENSITY, orasim.PREV_DENSITY, 0.0, 0.01)
    assert_bits(gpu.read(fsb.DENSITY), ora.read(orasim.DENSITY), "diffuse a=0")


@pytest.mark.parametrize("w,h", SIZES)
def test_project(w, h):
    gpu, ora = make_pair(w, h)
    rng = np.random.default_rng(w * h)
    seed_state(gpu, ora, rng, w, h)
    gpu.project(); ora.project()
    assert_bits(gpu.read(fsb.PREV_VEL_X), ora.read(orasim.PREV_VEL_X), "pressure")
    assert_bits(gpu.read(fsb.VELOCITY_X), ora.read(orasim.VEL_X), "velX after project")
    assert_bits(gpu.read(fsb.VELOCITY_Y), ora.read(orasim.VEL_Y), "velY after project")


@pytest.mark.parametrize("P,spw,stg,strips", [(1, 0, 0, 0), (3, 0, 0, 1), (7, 0, 0, 2), (40, 0, 0, 0), (20, 5, 0, 2),
                                               (33, 10, 0, 4), (20, 20, 0, 4), (40, 10, 2, 2), (16, 1, 0, 1),
                                               (100, 10, 0, 1), (0, 0, 0, 0)])
def test_project_sweep_counts(P, spw, stg, strips):
    w, h = 90, 150
    gpu = fsb.FluidSim(w, h, P, 5, sweeps_per_warp=spw, max_stages=stg, strips_per_cta=strips)
    ora = OracleSim(w, h, P, 5)
    rng = np.random.default_rng(P)
    seed_state(gpu, ora, rng, w, h)
    gpu.project(); ora.project()
    assert_bits(gpu.read(fsb.VELOCITY_X), ora.read(orasim.VEL_X), f"velX P={P}")
    assert_bits(gpu.read(fsb.VELOCITY_Y), ora.read(orasim.VEL_Y), f"velY P={P}")


@pytest.mark.parametrize("w,h", SIZES)
@pytest.mark.parametrize("scheme", [fsb.ADVECT_RK4, fsb.ADVECT_MACCORMACK, fsb.ADVECT_SEMI_LAGRANGIAN])
def test_advect(w, h, scheme):
    gpu, ora = make_pair(w, h, advect_scheme=scheme)
    rng = np.random.default_rng(scheme + w)
    seed_state(gpu, ora, rng, w, h, scale=3.0)
    for b in (0, 1, 2):
        gpu.advect(b, fsb.PREV_DENSITY, fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y, 0.01)
        ora.advect(b, orasim.PREV_DENSITY, orasim.DENSITY, orasim.VEL_X, orasim.VEL_Y, 0.01)
        assert_bits(gpu.read(fsb.PREV_DENSITY), ora.read(orasim.PREV_DENSITY), f"advect scheme={scheme} b={b}")


@pytest.mark.parametrize("chain", [0, 1, 3])
@pytest.mark.parametrize("name", ["config1_200x200_p20_d5", "small_48x40_p3_d2", "odd_257x131_p20_d5",
                                  "tall_131x257_p7_d3"])
def test_step_against_reference_golden(name, chain):
    """Whole-step parity against fields produced by the unmodified reference (tests/golden), with every pass as its own
    launch (chain 0), diffuse + project chained (1) and the whole step as one chained launch wherever one exists (3)."""
    g = load_golden(name)
    w, h, p, d = int(g["w"]), int(g["h"]), int(g["p"]), int(g["d"])
    gpu = fsb.FluidSim(w, h, p, d, chain=chain)
    apply_golden_inputs(gpu, g)
    checkpoints = sorted(int(k[2:]) for k in g.files if k.startswith("d_"))
    step = 0
    for cp in checkpoints:
        while step < cp:
            gpu.step(0.01)
            step += 1
        dd, uu, vv = gpu.fields()
        for got, key in ((dd, "d"), (uu, "u"), (vv, "v")):
            want = g[f"{key}_{cp}"]
            assert rel_max_abs(got, want) <= TOL, f"{name} {key} step {cp}"
            assert_bits(got, want, f"{name} {key} step {cp}")


@pytest.mark.parametrize("scheme", [fsb.ADVECT_MACCORMACK, fsb.ADVECT_SEMI_LAGRANGIAN])
def test_step_alternative_schemes(scheme):
    w, h = 80, 72
    gpu, ora = make_pair(w, h, advect_scheme=scheme)
    for s in (gpu, ora):
        demo_injection(s, w, h)
    for _ in range(5):
        gpu.step(0.01); ora.step(0.01)
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"scheme {scheme} field {n}")


def test_step_with_diffusion_and_zero_noise():
    w, h = 60, 84
    gpu, ora = make_pair(w, h, 7, 3, diffusion=0.0003, viscosity=0.002, noise="zero")
    d, u, v = vortex_lattice(w, h)
    for s, ids in ((gpu, (fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y)), (ora, (0, 1, 2))):
        s.write(ids[0], d); s.write(ids[1], u); s.write(ids[2], v)
    for _ in range(4):
        gpu.step(0.01); ora.step(0.01)
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"field {n}")


def test_async_read_is_a_snapshot():
    """fsb_read_field_async returns the field as of the call, even though the following steps (and source injections)
    are queued while the copy is still in flight.  After the first read the steps' launches leave a second copy of the
    density for the transfer and apply the pending sources themselves (so consecutive steps still overlap); a step with
    more sources than a launch carries, and a source added between a step and its read, take the other paths."""
    w, h = 512, 384
    gpu, ora = make_pair(w, h, 20, 5)
    demo_injection(gpu, w, h); demo_injection(ora, w, h)
    bufs = [gpu.pinned_array(), gpu.pinned_array()]
    snaps = []
    n = 8
    for k in range(n):
        for s in (gpu, ora):
            s.addDensity(w // 3 + k, h // 2, 2.5)
            s.addVelocity(w // 3, h // 2 + k, 0.3, -0.2)
            s.addVelocity(w // 3, h // 2 + k, -0.1, 0.05)       # twice the same cell: applied in call order
            s.addDensity(0, 5 + k, 1.0)                          # a wall cell
            if k == 4:
                for q in range(300):
                    s.addDensity(10 + q % 50, 20 + q // 50, 0.01 * (q + 1))
            s.step(0.01)
            if k == 5:
                s.addDensity(w // 2, h // 2, 4.0)                # after the step, before its read
        gpu.read_wait()                                  # the previous snapshot has arrived ...
        if k:
            assert_bits(bufs[(k - 1) % 2].reshape(w, h), snaps[-1], f"density snapshot of step {k - 1}")
        gpu.read_async(fsb.DENSITY, bufs[k % 2])         # ... queue this one and go on
        snaps.append(ora.read(0).copy())
    gpu.read_wait()
    assert_bits(bufs[(n - 1) % 2].reshape(w, h), snaps[-1], "density snapshot of the last step")
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"field {n}")
    # velocity snapshots take the conservative path (the next step waits for the copy)
    gpu.read_async(fsb.VELOCITY_X, bufs[0]); gpu.step(0.01); gpu.read_wait()
    assert_bits(bufs[0].reshape(w, h), ora.read(1), "velocity snapshot")


@pytest.mark.parametrize("seed", range(40))
def test_random_configurations(seed):
    """Random small grids, sweep counts, coefficients and tuning: two whole steps against the C restatement.  Covers
    the planner's mixed plans (sweep counts that are not multiples of 10), odd sizes and every strips-per-CTA choice."""
    rng = np.random.default_rng(1000 + seed)
    w, h = int(rng.integers(3, 140)), int(rng.integers(3, 400))
    p, d = int(rng.integers(0, 46)), int(rng.integers(0, 13))
    kw = dict(viscosity=float(rng.choice([0.0, 1e-4, 3e-3])), diffusion=float(rng.choice([0.0, 0.0, 2e-4])),
              noise=str(rng.choice(["glibc", "zero"])), strips_per_cta=int(rng.choice([0, 1, 2, 4])), chain=3)
    gpu, ora = make_pair(w, h, p, d, **kw)
    seed_state(gpu, ora, rng, w, h, scale=2.0)
    for k in range(2):
        x, y = int(rng.integers(0, w)), int(rng.integers(0, h))
        gpu.addDensity(x, y, 3.0); ora.addDensity(x, y, 3.0)
        gpu.addVelocity(x, y, -1.5, 0.75); ora.addVelocity(x, y, -1.5, 0.75)
        gpu.step(0.01); ora.step(0.01)
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"{w}x{h} P={p} D={d} {kw}: field {n}")


@pytest.mark.parametrize("w,h,p,d,steps", [(96, 200, 20, 5, 3), (64, 300, 100, 5, 2), (300, 1100, 40, 10, 2), (40, 257, 30, 5, 3),
                                           (2048, 2048, 20, 5, 2), (37, 4100, 50, 5, 2)])
@pytest.mark.parametrize("inject", ["lattice", "demo"])
def test_whole_step_as_one_chained_launch(w, h, p, d, steps, inject):
    """The whole step after the noise as ONE launch (chain level 3: also on shapes where the default keeps separate
    launches): several passes per projection (P = 100, 50, 30), two SOR stages (D = 10), odd widths, a grid wide enough
    for several CTAs per phase, and the reference's injection, whose |v| = 2 makes the back-traces leave their column
    block (the advection phases then wait for more gradient blocks and take the global-memory path)."""
    gpu, ora = make_pair(w, h, p, d, chain=3)
    if inject == "demo":
        for s in (gpu, ora):
            demo_injection(s, w, h)
    else:
        dd, uu, vv = vortex_lattice(w, h)
        for s, ids in ((gpu, (fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y)), (ora, (0, 1, 2))):
            s.write(ids[0], dd); s.write(ids[1], uu); s.write(ids[2], vv)
    n0 = gpu.launch_count()
    for _ in range(steps):
        gpu.step(0.01); ora.step(0.01)
    assert gpu.launch_count() - n0 <= 3 * steps            # noise (+ its zero-mode twin) + the chained launch
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"{w}x{h} P={p} D={d} {inject}: field {n}")


@pytest.mark.parametrize("w,h,p,d", [(300, 1100, 40, 10), (96, 200, 20, 5), (40, 257, 30, 5), (1024, 4096, 40, 5),
                                     (400, 7168, 20, 5)])      # the last one: 64-column blocks, as at 8192^2
@pytest.mark.parametrize("noise", ["glibc", "zero"])
def test_overlapping_steps(w, h, p, d, noise):
    """fsb_step as one launch that contains the noise, consecutive steps overlapping on the GPU (the default where a
    chained configuration exists): seven steps back to back (the velocity sets rotate with period 2, the densities with
    period 3), then calls that put other work between the steps (sources, reads, a velocityStep of its own) and more
    steps; compared with the oracle after every stage, and with the same library run without the overlap."""
    gpu, ora = make_pair(w, h, p, d, chain=3, noise=noise)
    ser = fsb.FluidSim(w, h, p, d, chain=3, noise=noise, overlap=False)
    dd, uu, vv = vortex_lattice(w, h)
    for s, ids in ((gpu, (fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y)), (ser, (fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y)),
                   (ora, (0, 1, 2))):
        s.write(ids[0], dd); s.write(ids[1], uu); s.write(ids[2], vv)
    n0 = gpu.launch_count()
    for _ in range(7):
        gpu.step(0.01); ser.step(0.01); ora.step(0.01)
    assert gpu.launch_count() - n0 == 7                     # one launch per step, noise included
    for a, b, c, n in zip(gpu.fields(), ora.fields(), ser.fields(), "duv"):
        assert_bits(a, b, f"{w}x{h} after 7 steps: field {n}")
        assert_bits(c, b, f"{w}x{h} after 7 steps without overlap: field {n}")
    for k in range(3):
        for s in (gpu, ser, ora):
            s.addDensity(w // 2, h // 3 + k, 2.0)
            s.addVelocity(w // 3, h // 2, 0.5, -0.25)
            s.step(0.01)
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"{w}x{h} with sources between the steps: field {n}")
    for s in (gpu, ora):
        s.velocityStep(0.01); s.densityStep(0.01)
        s.step(0.01); s.step(0.01)
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"{w}x{h} after velocityStep / densityStep / two more steps: field {n}")
    ser.close()


def test_velocity_and_density_step_entry_points():
    """fsb_op_velocity_step / fsb_op_density_step (velocityStep, densityStep of src/fluid_sim.cpp:370-422) called one after
    the other equal step(); checked against the oracle's two halves after each call."""
    w, h = 90, 140
    gpu, ora = make_pair(w, h, 20, 5)
    for s in (gpu, ora):
        demo_injection(s, w, h)
    for k in range(3):
        gpu.velocityStep(0.01); ora.velocityStep(0.01)
        for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
            assert_bits(a, b, f"after velocityStep {k}: field {n}")
        gpu.densityStep(0.01); ora.densityStep(0.01)
        for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
            assert_bits(a, b, f"after densityStep {k}: field {n}")


def test_single_cell_reads_after_a_step():
    """A caller that probes a few cells gets them without a whole-field download (the first reads after a change fetch
    single cells); a caller that keeps reading is served from the host mirror.  Same values either way, and sources
    added in between are visible."""
    w, h = 70, 66
    gpu, ora = make_pair(w, h, 10, 5)
    for s in (gpu, ora):
        demo_injection(s, w, h)
    gpu.step(0.01); ora.step(0.01)
    d, u, v = ora.fields()
    cells = [(i, j) for i in range(0, w, 3) for j in range(0, h, 5)]          # more than the probe budget
    for (i, j) in cells:
        assert gpu.getDensity(i, j).view(np.uint32) == d[i, j].view(np.uint32)
        gu, gv = gpu.getVelocity(i, j)
        assert gu.view(np.uint32) == u[i, j].view(np.uint32) and gv.view(np.uint32) == v[i, j].view(np.uint32)
    gpu.addDensity(7, 9, 1.25); ora.addDensity(7, 9, 1.25)
    assert gpu.getDensity(7, 9).view(np.uint32) == ora.read(0)[7, 9].view(np.uint32)
    gpu.step(0.01); ora.step(0.01)
    assert gpu.getDensity(35, 33).view(np.uint32) == ora.read(0)[35, 33].view(np.uint32)     # one probe after the step
    assert gpu.getDensity(-1, 3) == 0 and gpu.getDensity(3, h) == 0


def test_many_launches_ring_tags_wrap():
    """The rings between CTAs carry a tag made of the launch number (14 bits) and the hand-over unit: more than 2^14
    launches on one simulation must stay exact.  One strip per CTA, so every strip boundary crosses a ring."""
    w, h = 40, 100
    gpu, ora = make_pair(w, h, 10, 5, noise="zero", strips_per_cta=1)
    d, u, v = vortex_lattice(w, h)
    for s, ids in ((gpu, (fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y)), (ora, (0, 1, 2))):
        s.write(ids[0], d); s.write(ids[1], u); s.write(ids[2], v)
    steps = 5600                                   # 3 sweep launches per step: 16800 launches
    for _ in range(steps):
        gpu.step(0.01); ora.step(0.01)
    assert gpu.launch_count() > 3 * steps
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"field {n} after {steps} steps")


def test_step_2048_config2():
    """Config 2 (2048^2, P=20): two steps of the vortex lattice against the oracle."""
    w = h = 2048
    gpu, ora = make_pair(w, h)
    d, u, v = vortex_lattice(w, h)
    gpu.add_field(fsb.DENSITY, d); gpu.add_field(fsb.VELOCITY_X, u); gpu.add_field(fsb.VELOCITY_Y, v)
    ora.write(0, d); ora.write(1, u); ora.write(2, v)
    for _ in range(2):
        gpu.step(0.01); ora.step(0.01)
    for a, b, n in zip(gpu.fields(), ora.fields(), "duv"):
        assert_bits(a, b, f"2048^2 field {n}")


def test_step_8192_config3_full_size():
    """Config 3 at its full size (8192^2, P=40, the bench workload): one step of the vortex lattice with glibc noise
    against the real reference (about a minute on one host core), else against the C restatement."""
    w = h = 8192
    if refsim.available(w, h, 40, 5):
        ref = refsim.RefSim(w, h, 40, 5, seed=1)
    else:
        ref = OracleSim(w, h, 40, 5)
    gpu = fsb.FluidSim(w, h, 40, 5)
    d, u, v = vortex_lattice(w, h)
    for fid, f in ((0, d), (1, u), (2, v)):
        ref.write(fid, f); gpu.write(fid, f)
    del d, u, v
    ref.step(0.01); gpu.step(0.01)
    for fid, n in ((0, "d"), (1, "u"), (2, "v")):
        a, b = gpu.read(fid), ref.read(fid)
        assert_bits(a, b, f"8192^2 field {n}")
        assert rel_max_abs(a, b) <= TOL


def test_step_real_reference_wide():
    """A wide grid (35 strips, several CTAs in the chain) against the real reference, if prebuilt."""
    w, h, p, d = 96, 1100, 20, 5
    if not refsim.available(w, h, p, d):
        pytest.skip("oracle/_ref not present")
    ref = refsim.RefSim(w, h, p, d, seed=1)
    gpu = fsb.FluidSim(w, h, p, d)
    rng = np.random.default_rng(2)
    for fid in (0, 1, 2):
        f = random_field(rng, w, h, 0.5)
        ref.write(fid, f); gpu.write(fid, f)
    for _ in range(3):
        ref.step(0.01); gpu.step(0.01)
    for a, b, n in zip(gpu.fields(), ref.fields(), "duv"):
        assert_bits(a, b, f"wide field {n}")


def test_api_semantics():
    """Bounds behaviour and immediate visibility of adds (src/fluid_sim.cpp:424-457), reset (main.cpp:289)."""
    gpu = fsb.FluidSim(32, 24, noise="zero")
    gpu.addDensity(-1, 0, 5.0); gpu.addDensity(32, 0, 5.0); gpu.addDensity(0, 24, 5.0)   # ignored
    gpu.addVelocity(3, -1, 1.0, 1.0)                                                        # ignored
    assert gpu.getDensity(-1, 0) == 0 and gpu.getDensity(40, 2) == 0
    assert gpu.getVelocity(0, 99) == (0, 0)
    gpu.addDensity(3, 4, 1.5); gpu.addDensity(3, 4, 2.25)
    assert gpu.getDensity(3, 4) == np.float32(3.75)
    gpu.addVelocity(5, 6, 3.0, 4.0)
    assert gpu.getVelocity(5, 6) == (3.0, 4.0)
    assert gpu.getVelocityMagnitude(5, 6) == np.float32(5.0)
    assert gpu.getNormalizedVelocity(5, 6) == (np.float32(0.6), np.float32(0.8))
    gpu.addDensity(3, 4, 1.0)                     # after a read: mirror and device both see it
    assert gpu.getDensity(3, 4) == np.float32(4.75)
    assert gpu.read(fsb.DENSITY)[3, 4] == np.float32(4.75)
    assert gpu.read(fsb.DENSITY).sum() == np.float32(4.75)
    assert gpu.getWidth() == 32 and gpu.getHeight() == 24
    gpu.step(0.01)
    assert gpu.read(fsb.DENSITY).sum() > 0
    gpu.reset()
    assert not gpu.read(fsb.DENSITY).any() and not gpu.read(fsb.VELOCITY_X).any()
    assert gpu.launch_count() > 0


def test_getters_match_oracle():
    gpu, ora = make_pair(40, 40)
    for s in (gpu, ora):
        s.addVelocity(5, 6, 3.0, 4.0); s.addVelocity(7, 8, 0.0005, 0.0); s.addVelocity(9, 9, 30.0, 40.0)
        s.addVelocity(1, 1, 2.0, 1.0); s.addVelocity(2, 2, 5.5, 2.0)
    for (x, y) in [(5, 6), (7, 8), (9, 9), (1, 1), (2, 2), (0, 0), (-1, 3), (40, 0)]:
        assert gpu.getVelocityMagnitude(x, y) == ora.getVelocityMagnitude(x, y)
        assert gpu.getNormalizedVelocity(x, y) == ora.getNormalizedVelocity(x, y)
        assert gpu.getVelocityColor(x, y) == ora.getVelocityColor(x, y)


def test_two_sims_share_nothing():
    a = fsb.FluidSim(64, 64); b = fsb.FluidSim(64, 64)
    demo_injection(a, 64, 64); demo_injection(b, 64, 64)
    for _ in range(3):
        a.step(0.01)
    for _ in range(3):
        b.step(0.01)
    for x, y in zip(a.fields(), b.fields()):
        assert_bits(x, y, "independent sims")


def test_cpp_dropin_headless_main():
    """examples/headless_main.cpp (the reference's main loop without GL, src/main.cpp:350-396) linked against
    the drop-in class: same sums as the unmodified reference after 100 frames, and the reset idiom works."""
    import os
    import subprocess
    from tests.util import ROOT
    exe = os.path.join(ROOT, "examples", "headless_main")
    if not os.path.exists(exe):
        pytest.skip("examples/headless_main not built")
    out = subprocess.run([exe, "100"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    g = load_golden("config1_200x200_p20_d5")
    sd = float(np.sum(g["d_100"].astype(np.float64)))
    su = float(np.sum(g["u_100"].astype(np.float64)))
    sv = float(np.sum(g["v_100"].astype(np.float64)))
    want = "frames 100  sum(d) %.8f  sum(u) %.8f  sum(v) %.8f  d[100][110] %.9g" % (sd, su, sv, float(g["d_100"][100, 110]))
    lines = out.stdout.strip().splitlines()
    assert lines[0] == want, (lines[0], want)
    assert lines[1] == "after reset d[100][110] 0"


def test_cpp_dropin_object_semantics_against_reference(tmp_path):
    """tests/dropin_vs_ref.cpp: the drop-in class and the compiled reference side by side in one process - copy
    construction after three steps, two objects stepped alternately, the `sim = FluidSim()` reset idiom, copy
    assignment: bit-identical fields throughout (the reference's rand() is process-global, src/fluid_sim.cpp:378-379)."""
    import os
    import subprocess
    from oracle import refsim
    from tests.util import ROOT
    libdir = os.path.join(ROOT, "fluid_sim_b200")
    if not os.path.exists(os.path.join(libdir, "libfluidsim_dropin.so")) or not refsim.available(200, 200, 20, 5):
        pytest.skip("drop-in library or oracle/_ref/libref_200x200_p20_d5.so not built")
    exe = str(tmp_path / "dropin_vs_ref")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "dropin_vs_ref.cpp"),
                    "-o", exe, "-L", libdir, "-lfluidsim_dropin", "-lfluidsim_b200", "-ldl", f"-Wl,-rpath,{libdir}"], check=True)
    out = subprocess.run([exe, refsim.lib_path(200, 200, 20, 5)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "ALL IDENTICAL" in out.stdout, out.stdout + out.stderr


@pytest.mark.parametrize("args", ["256 512 20 5 2", "200 1024 40 5 1", "300 256 33 0 2", "130 512 100 5 1",
                                  "128 512 20 7 2 diff=2e-4", "128 512 10 5 2 visc=0", "8192 512 20 5 2 demo=1 halo=256",
                                  "256 512 20 5 2 chain=0", "130 512 100 5 1 chain=0", "512 1024 40 5 3 demo=1"])
def test_multi_gpu_matches_single_gpu(args):
    """Column blocks on 2 GPUs give the single-GPU bits: as one chained launch per rank and step (results pushed into
    the neighbour's halo through peer mappings) where a chained configuration exists, and pass by pass with NCCL halo
    exchanges otherwise (odd sweep counts, zero viscosity, density diffusion, chain=0)."""
    import os
    import subprocess
    import sys
    import torch
    from tests.util import ROOT
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "mgpu_check.py")] + args.split()
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert "MGPU-PARITY PASS" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
