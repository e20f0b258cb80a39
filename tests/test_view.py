"""Headless density view: the CUDA kernel behind fsb_view_density against the CPU restatement
of the reference viewer's quads + density shader (oracle/view_oracle.c; parity unpinned: the reference draws through
OpenGL, which is unavailable on this box)."""
import numpy as np
import pytest

from oracle import orasim


def _density(w, h, seed):
    rng = np.random.default_rng(seed)
    d = rng.random((w, h), dtype=np.float32) * np.float32(1.6) - np.float32(0.2)   # below 0.01, inside (0,1), above 1
    d[rng.random((w, h)) < 0.3] = 0.0
    d[0, 0] = np.float32(0.01)        # exactly the skip threshold: drawn
    d[1, 0] = np.float32(0.0099999)   # just below: skipped
    d[2, 0] = np.float32(0.4)         # alpha reaches the 0.8 cap exactly at 0.4
    return d


def test_oracle_view_known_pixels():
    """Hand-checked pixels of the restatement itself: one cell per pixel."""
    d = np.zeros((4, 3), np.float32)
    d[1, 2] = 0.25      # rgb = (0.25, 0, 0.75) * 0.5, alpha = 0.25 + 0.5
    d[3, 0] = 2.0       # clamps: rgb = (1, 0, 0) * 0.8, alpha = 0.64 + 0.2
    img = orasim.view_density(d, 4, 3)
    assert img.shape == (3, 4, 4)
    assert tuple(img[2, 1]) == (32, 0, 96, 191)      # 0.125*255 = 31.875 -> 32, 0.375*255 = 95.6 -> 96, 0.75*255 = 191.25
    assert tuple(img[0, 3]) == (204, 0, 0, 214)      # 0.8*255 = 204, 0.84*255 = 214.2
    assert tuple(img[0, 0]) == (0, 0, 0, 255)        # cleared
    # several pixels per cell / several cells per pixel keep the image piecewise constant / point sampled
    big = orasim.view_density(d, 8, 6)
    assert (big[4:6, 2:4] == img[2, 1]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,pw,ph", [(200, 200, 800, 600), (48, 40, 97, 53), (257, 131, 800, 600),
                                       (2048, 2048, 800, 600), (64, 96, 64, 96), (640, 512, 1920, 1080)])
def test_view_density_matches_restatement(w, h, pw, ph):
    import fluid_sim_b200 as fsb
    d = _density(w, h, seed=w * 7 + h)
    with fsb.FluidSim(w, h, noise="zero") as sim:
        sim.write(fsb.DENSITY, d)
        got = sim.view_density(pw, ph)
        again = sim.view_density(pw, ph)
    want = orasim.view_density(d, pw, ph)
    assert np.array_equal(got, again)
    bad = np.argwhere((got != want).any(axis=2))
    assert bad.size == 0, f"{len(bad)} pixels differ, first at (row, col) = {tuple(bad[0])}: {got[tuple(bad[0])]} vs {want[tuple(bad[0])]}"
    assert (got[..., 1] == 0).all()                       # the shader has no green
    assert got[..., 3].min() >= 191                       # alpha = a*a + 1 - a has its minimum 0.75 at a = 0.5


@pytest.mark.gpu
def test_view_density_after_steps_and_bad_arguments():
    import ctypes as C
    import fluid_sim_b200 as fsb
    from tests.util import demo_injection
    with fsb.FluidSim(200, 200) as sim:
        demo_injection(sim, 200, 200)
        for _ in range(3):
            sim.step(0.01)
        img = sim.view_density()                          # the reference's 800 x 600 window
        want = orasim.view_density(sim.read(fsb.DENSITY), 800, 600)
        assert np.array_equal(img, want)
        assert (img[..., 0] > 0).any()                    # something was drawn
        with pytest.raises(fsb.FsbError):
            sim.view_density(0, 600, out=np.empty(0, np.uint8))
        ptr = C.c_void_p()
        sim._check(sim._lib.fsb_view_density_device(sim._h, 320, 200, C.byref(ptr)))
        assert ptr.value
        sim.sync()


def test_oracle_glyph_known_pixels():
    """One horizontal glyph: 64 x 64 cells on 64 x 64 pixels, velocity (5, 0) at cell (10, 20)."""
    w = h = 64
    d = np.zeros((w, h), np.float32)
    u = np.zeros((w, h), np.float32); v = np.zeros((w, h), np.float32)
    u[10, 20] = 5.0                                  # |v| = 5: ramp at 0.5 -> (0, 1, 0) exactly at the green corner
    img = orasim.view_velocity(d, u, v, 64, 64)
    # start (10.5, 20.5) px, 0.05 NDC = 1.6 px long: column 10 (centre 10.5) is the only centre in [10.5, 12.1) besides 11
    lit = np.argwhere(img[..., 1] > 0)
    assert {tuple(p) for p in lit} == {(19, 10), (20, 10), (19, 11), (20, 11)}      # two rows per column: [-1, 1) around y
    assert tuple(img[20, 10]) == (0, 204, 0, 214)    # green * 0.8 = 204; alpha 0.64 + 0.2 -> 214.2
    u[11, 20] = 5.0                                  # odd cell: not sampled
    assert np.array_equal(orasim.view_velocity(d, u, v, 64, 64), img)


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,pw,ph", [(200, 200, 800, 600), (48, 40, 97, 53), (257, 131, 640, 480), (64, 96, 64, 96)])
def test_view_velocity_matches_restatement(w, h, pw, ph):
    import fluid_sim_b200 as fsb
    rng = np.random.default_rng(w + h)
    d = _density(w, h, seed=w)
    u = (rng.standard_normal((w, h)) * 3).astype(np.float32)
    v = (rng.standard_normal((w, h)) * 3).astype(np.float32)
    u[rng.random((w, h)) < 0.4] = 0.0; v[u == 0.0] = np.float32(0.05)       # below the 0.1 threshold: no line
    u[0, 0], v[0, 0] = 30.0, -40.0                   # speed above the ramp's end
    with fsb.FluidSim(w, h, noise="zero") as sim:
        sim.write(fsb.DENSITY, d); sim.write(fsb.VELOCITY_X, u); sim.write(fsb.VELOCITY_Y, v)
        got = sim.view_velocity(pw, ph)
        dens = sim.view_density(pw, ph)
    want = orasim.view_velocity(d, u, v, pw, ph)
    bad = np.argwhere((got != want).any(axis=2))
    assert bad.size == 0, f"{len(bad)} pixels differ, first at (row, col) = {tuple(bad[0])}: {got[tuple(bad[0])]} vs {want[tuple(bad[0])]}"
    assert (got != dens).any()                       # lines were drawn over the density image


@pytest.mark.gpu
def test_headless_view_example_writes_the_reference_frame(tmp_path):
    """examples/headless_view.cpp (drop-in class + C ABI): 3 frames of the demo, image equal to the restatement's."""
    import os
    import subprocess
    import fluid_sim_b200 as fsb
    from tests.util import ROOT, demo_injection
    exe = os.path.join(ROOT, "examples", "headless_view")
    if not os.path.exists(exe):
        pytest.skip("examples/headless_view not built")
    ppm = tmp_path / "f.ppm"
    out = subprocess.run([exe, "3", str(ppm), "1", "320", "240"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    raw = ppm.read_bytes()
    head = b"P6\n320 240\n255\n"
    assert raw.startswith(head) and len(raw) == len(head) + 320 * 240 * 3
    got = np.frombuffer(raw[len(head):], np.uint8).reshape(240, 320, 3)[::-1]      # back to bottom-up rows
    with fsb.FluidSim(200, 200) as sim:
        demo_injection(sim, 200, 200)
        for _ in range(3):
            sim.step(0.01)
        d, u, v = sim.read(fsb.DENSITY), sim.read(fsb.VELOCITY_X), sim.read(fsb.VELOCITY_Y)
    want = orasim.view_velocity(d, u, v, 320, 240)[..., :3]
    assert np.array_equal(got, want)
