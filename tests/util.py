"""Shared helpers for the parity tests."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def demo_injection(sim, w, h):
    """Start-up injection of the reference's main loop (src/main.cpp:351-363)."""
    cx, cy = w // 2, h // 2
    for i in range(-5, 6):
        for j in range(-5, 6):
            if i * i + j * j < 25:
                sim.addDensity(cx + i, cy + j, 10.0)
                sim.addVelocity(cx + i, cy + j, 0.0, 2.0)


def apply_golden_inputs(sim, g):
    w, h = int(g["w"]), int(g["h"])
    demo_injection(sim, w, h)
    for x, y, dens, ax, ay in g["calls"]:
        sim.addDensity(int(x), int(y), float(dens))
        sim.addVelocity(int(x), int(y), float(ax), float(ay))


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def bits_equal(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def describe_diff(a, b):
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    bad = a.view(np.uint32) != b.view(np.uint32)
    n = int(bad.sum())
    if n == 0:
        return "identical"
    idx = np.argwhere(bad)
    scale = float(np.max(np.abs(b))) or 1.0
    err = float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64)))) / scale
    first = tuple(idx[0])
    return (f"{n} of {a.size} cells differ (first at {first}: got {a[first]!r} want {b[first]!r}); "
            f"rel max-abs {err:.3e}; rows {idx[:,0].min()}..{idx[:,0].max()} cols {idx[:,1].min()}..{idx[:,1].max()}")


def assert_bits(a, b, what=""):
    assert bits_equal(a, b), f"{what}: {describe_diff(a, b)}"


def rel_max_abs(a, b):
    """north_star's metric: max|a-b| / max|b| per field."""
    scale = float(np.max(np.abs(b))) or 1.0
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)))) / scale


def random_field(rng, w, h, scale=1.0):
    return (rng.standard_normal((w, h)) * scale).astype(np.float32)


def vortex_lattice(w, h, dt=0.01, k=8):
    """Synthetic input I3 of SURVEY.md 8(d): a vortex lattice with a 4-cell displacement per step."""
    i = np.arange(w, dtype=np.float64)[:, None]
    j = np.arange(h, dtype=np.float64)[None, :]
    A = 4.0 / (dt * w)
    sx, cx = np.sin(2 * np.pi * k * i / w), np.cos(2 * np.pi * k * i / w)
    sy, cy = np.sin(2 * np.pi * k * j / h), np.cos(2 * np.pi * k * j / h)
    u = (A * sx * cy).astype(np.float32)
    v = (-A * cx * sy).astype(np.float32)
    d = (0.5 * (1 + sx * sy)).astype(np.float32)
    return d, u, v


def vortex_lattice_block(w, h, a, b, dt=0.01, k=8):
    """Columns [a, b) of vortex_lattice(w, h) (columns outside the grid are zero)."""
    i = np.arange(w, dtype=np.float64)[:, None]
    j = np.arange(a, b, dtype=np.float64)[None, :]
    A = 4.0 / (dt * w)
    sx, cx = np.sin(2 * np.pi * k * i / w), np.cos(2 * np.pi * k * i / w)
    sy, cy = np.sin(2 * np.pi * k * j / h), np.cos(2 * np.pi * k * j / h)
    inside = ((j >= 0) & (j < h)).astype(np.float64)
    u = (A * sx * cy * inside).astype(np.float32)
    v = (-A * cx * sy * inside).astype(np.float32)
    d = (0.5 * (1 + sx * sy) * inside).astype(np.float32)
    return d, u, v
