#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REAL reference (oracle/_ref, built from /root/reference by
oracle/build_ref.py).  Run in the build container, where /root/reference exists:

    python oracle/build_ref.py && python tests/golden/make_golden.py

Each file holds the inputs (as the list of add* calls) and the reference's d/u/v after N steps.
The reference has no tests or fixtures of its own; these are its outputs.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.refsim import RefSim, demo_injection  # noqa: E402


def extra_injection(sim, w, h, seed, n=40):
    rng = np.random.default_rng(seed)
    calls = []
    for _ in range(n):
        x = int(rng.integers(-2, w + 2)); y = int(rng.integers(-2, h + 2))
        a = float(np.float32(rng.normal())); b = float(np.float32(rng.normal()))
        sim.addDensity(x, y, abs(a)); sim.addVelocity(x, y, a, b)
        calls.append((x, y, abs(a), a, b))
    return np.array(calls, dtype=np.float64)


def make(name, w, h, p, d, steps, checkpoints, extra_seed=None):
    sim = RefSim(w, h, p, d, seed=1)
    demo_injection(sim, w, h)
    calls = extra_injection(sim, w, h, extra_seed) if extra_seed is not None else np.zeros((0, 5))
    out = {"w": w, "h": h, "p": p, "d": d, "dt": np.float32(0.01), "calls": calls}
    for s in range(1, steps + 1):
        sim.step(0.01)
        if s in checkpoints:
            dd, uu, vv = sim.fields()
            out[f"d_{s}"] = dd; out[f"u_{s}"] = uu; out[f"v_{s}"] = vv
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print("wrote", name)


if __name__ == "__main__":
    make("config1_200x200_p20_d5", 200, 200, 20, 5, 100, (1, 10, 100))
    make("small_48x40_p3_d2", 48, 40, 3, 2, 30, (1, 30), extra_seed=3)
    make("odd_257x131_p20_d5", 257, 131, 20, 5, 10, (1, 10), extra_seed=4)
    make("tall_131x257_p7_d3", 131, 257, 7, 3, 10, (1, 10), extra_seed=5)
