#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY -- builds the real reference into oracle/_ref/.

Compiles /root/reference/src/fluid_sim.cpp *where it lies* (nothing is copied into
the repository) with the reference's own flags (`-std=c++17 -O2`,
/root/reference/CMakeLists.txt:4,24; no -march, hence no FMA contraction) against
the GLM stand-in in oracle/glm_stub, together with oracle/ref_shim.cpp, into one
shared library per (W, H, P, D) variant:

    oracle/_ref/libref_<W>x<H>_p<P>_d<D>.so

The variant 200x200 / P=20 / D=5 is the verbatim source.  Every other variant is a
*generated* copy of include/fluid_sim.h and src/fluid_sim.cpp with exactly four
literal substitutions: GRID_SIZE_X, GRID_SIZE_Y
(include/fluid_sim.h:11-12), `k < 20` (src/fluid_sim.cpp:319) and `k < 5` (:50).
Generated sources are written under oracle/_ref/gen/ only (git-ignored), and each
substitution is asserted to hit exactly one site.

oracle/_ref/ is git-ignored but NOT gpurun-ignored: the built .so files travel to the
GPU box, where /root/reference does not exist.  Run this script here, in the build
container; `__graft_entry__.build()` calls it when /root/reference is present.
"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("FLUID_SIM_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")

# (W, H, P, D).  The first is the verbatim reference.
DEFAULT_VARIANTS = [
    (200, 200, 20, 5),     # config 1 (verbatim)
    (64, 64, 20, 5),       # small square
    (48, 40, 3, 2),        # tiny, odd sweep counts, W != H
    (257, 131, 20, 5),     # odd, non-square
    (131, 257, 7, 3),      # transposed aspect, sweep counts not in {20,5}
    (96, 1100, 20, 5),     # wide: many column strips
    (512, 512, 40, 5),     # P = 40 (two solver passes)
    (2048, 2048, 20, 5),   # config 2
    (128, 8192, 40, 5),    # bounded samples of config 3 (same columns, fewer rows) for the CPU baseline
    (256, 8192, 40, 5),
    (512, 8192, 40, 5),
    (1024, 8192, 40, 5),
    (8192, 8192, 40, 5),   # config 3 (one step costs about a minute of CPU)
    # bounded samples of the multi-GPU workloads (same columns, few rows) for bench.py's reference arm at N > 1
    (512, 4096, 100, 5),   # config 5 at 1 GPU (32768 x 4096)
    (256, 8192, 100, 5),   # config 5 at 2 GPUs
    (128, 16384, 100, 5),  # config 5 at 4 GPUs
    (64, 32768, 100, 5),   # config 5 at 8 GPUs
    (128, 16384, 20, 5),   # config 4 (16384^2, strong scaling)
]

CXXFLAGS = ["-std=c++17", "-O2", "-fPIC"]


def lib_path(w, h, p, d):
    return os.path.join(OUT, f"libref_{w}x{h}_p{p}_d{d}.so")


def sub_once(text, pattern, repl, what):
    new, n = re.subn(pattern, repl, text)
    if n != 1:
        raise RuntimeError(f"substitution '{what}' matched {n} sites, expected 1")
    return new


def generate(w, h, p, d):
    """Write the literal-substituted copy under oracle/_ref/gen/<tag>/ and return (incdir, cpp)."""
    tag = f"{w}x{h}_p{p}_d{d}"
    gdir = os.path.join(OUT, "gen", tag)
    os.makedirs(os.path.join(gdir, "include"), exist_ok=True)
    os.makedirs(os.path.join(gdir, "src"), exist_ok=True)
    hdr = open(os.path.join(REF, "include", "fluid_sim.h")).read()
    hdr = sub_once(hdr, r"GRID_SIZE_X = 200;", f"GRID_SIZE_X = {w};", "GRID_SIZE_X")
    hdr = sub_once(hdr, r"GRID_SIZE_Y = 200;", f"GRID_SIZE_Y = {h};", "GRID_SIZE_Y")
    src = open(os.path.join(REF, "src", "fluid_sim.cpp")).read()
    src = sub_once(src, r"k < 20;", f"k < {p};", "projection sweeps")
    src = sub_once(src, r"k < 5;", f"k < {d};", "diffusion sweeps")
    with open(os.path.join(gdir, "include", "fluid_sim.h"), "w") as f:
        f.write(hdr)
    with open(os.path.join(gdir, "src", "fluid_sim.cpp"), "w") as f:
        f.write(src)
    return os.path.join(gdir, "include"), os.path.join(gdir, "src", "fluid_sim.cpp")


def build(w, h, p, d, force=False):
    os.makedirs(OUT, exist_ok=True)
    out = lib_path(w, h, p, d)
    if os.path.exists(out) and not force:
        return out
    if (w, h, p, d) == (200, 200, 20, 5):
        incdir = os.path.join(REF, "include")
        cpp = os.path.join(REF, "src", "fluid_sim.cpp")
    else:
        incdir, cpp = generate(w, h, p, d)
    cmd = ["g++", *CXXFLAGS, "-shared", "-Wl,-Bsymbolic-functions",
           f"-DREF_PROJ_ITERS={p}", f"-DREF_DIFF_ITERS={d}",
           "-I", incdir, "-I", os.path.join(HERE, "glm_stub"),
           cpp, os.path.join(HERE, "ref_shim.cpp"), "-o", out, "-ldl"]
    subprocess.run(cmd, check=True)
    return out


def main(argv):
    if not os.path.isdir(REF):
        print(f"reference not present at {REF}; keeping prebuilt files in {OUT}")
        return 0
    force = "--force" in argv
    variants = DEFAULT_VARIANTS
    args = [a for a in argv if not a.startswith("--")]
    if args:
        variants = [tuple(int(x) for x in a.replace("x", ",").split(",")) for a in args]
    for v in variants:
        path = build(*v, force=force)
        print("built", os.path.relpath(path, HERE))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
