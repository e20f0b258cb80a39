"""TEST INFRASTRUCTURE ONLY -- ctypes front end to oracle/liboracle.so, the runtime-sized
C restatement of the reference solver (oracle/fluid_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")

DENSITY, VEL_X, VEL_Y, PREV_DENSITY, PREV_VEL_X, PREV_VEL_Y, TEMP1, TEMP2 = range(8)
SCHEME_RK4, SCHEME_MACCORMACK, SCHEME_SEMI_LAGRANGIAN = 0, 1, 2

_fp = C.POINTER(C.c_float)
_lib = None


def build(force=False):
    srcs = [os.path.join(HERE, "fluid_oracle.c"), os.path.join(HERE, "view_oracle.c")]
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(f) for f in srcs):
        subprocess.run(["gcc", "-std=c11", "-O2", "-ffp-contract=off", "-fPIC", "-shared",
                        *srcs, "-o", LIB, "-lm"], check=True)
    return LIB


def view_density(d, px_w=800, px_h=600):
    """oracle/view_oracle.c: the reference viewer's density image of a [W, H] float32 array, [px_h, px_w, 4] uint8."""
    L = lib()
    d = np.ascontiguousarray(d, np.float32)
    out = np.empty((px_h, px_w, 4), np.uint8)
    L.ora_view_density.argtypes = [_fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.ora_view_density.restype = None
    L.ora_view_density(d.ctypes.data_as(_fp), d.shape[0], d.shape[1], px_w, px_h, out.ctypes.data)
    return out


def view_velocity(d, u, v, px_w=800, px_h=600):
    """oracle/view_oracle.c: the viewer's image with velocity vectors on (density image, then the glyph lines)."""
    L = lib()
    img = view_density(d, px_w, px_h)
    u = np.ascontiguousarray(u, np.float32)
    v = np.ascontiguousarray(v, np.float32)
    L.ora_view_glyphs.argtypes = [_fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.ora_view_glyphs.restype = None
    L.ora_view_glyphs(u.ctypes.data_as(_fp), v.ctypes.data_as(_fp), u.shape[0], u.shape[1], px_w, px_h, img.ctypes.data)
    return img


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(LIB)
    L.ora_create.restype = C.c_void_p
    L.ora_create.argtypes = [C.c_int] * 4
    L.ora_destroy.argtypes = [C.c_void_p]
    L.ora_reset.argtypes = [C.c_void_p]
    L.ora_set_params.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int]
    L.ora_set_noise.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint64]
    L.ora_rand.argtypes = [C.c_void_p]
    L.ora_rand_count.argtypes = [C.c_void_p]
    L.ora_rand_count.restype = C.c_uint64
    for n in ("ora_step", "ora_velocity_step", "ora_density_step"):
        getattr(L, n).argtypes = [C.c_void_p, C.c_float]
    L.ora_add_density.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float]
    L.ora_add_velocity.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float]
    L.ora_read_field.argtypes = [C.c_void_p, C.c_int, _fp]
    L.ora_write_field.argtypes = [C.c_void_p, C.c_int, _fp]
    L.ora_set_boundary.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.ora_diffuse.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float]
    L.ora_project.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    L.ora_advect.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
    L.ora_add_noise.argtypes = [C.c_void_p]
    L.ora_bilinear.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float]
    L.ora_bilinear.restype = C.c_float
    L.ora_velocity_magnitude.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.ora_velocity_magnitude.restype = C.c_float
    L.ora_normalized_velocity.argtypes = [C.c_void_p, C.c_int, C.c_int, _fp]
    L.ora_velocity_color.argtypes = [C.c_void_p, C.c_int, C.c_int, _fp]
    _lib = L
    return L


class OracleSim:
    """Runtime-sized oracle with the reference's method names (include/fluid_sim.h:20-37)."""

    def __init__(self, w=200, h=200, p=20, d=5, visc=1e-4, diff=0.0, omega=1.5,
                 scheme=SCHEME_RK4, noise="glibc", seed=1, skip=0):
        self.L = lib()
        self.w, self.h, self.p, self.d = w, h, p, d
        self.h_ = self.L.ora_create(w, h, p, d)
        if not self.h_:
            raise MemoryError("oracle allocation failed")
        self.L.ora_set_params(self.h_, visc, diff, omega, scheme)
        self.L.ora_set_noise(self.h_, 0 if noise == "glibc" else 1, seed, skip)

    def close(self):
        if self.h_:
            self.L.ora_destroy(self.h_)
            self.h_ = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self, dt=0.01):
        self.L.ora_step(self.h_, dt)

    def velocityStep(self, dt):
        self.L.ora_velocity_step(self.h_, dt)

    def densityStep(self, dt):
        self.L.ora_density_step(self.h_, dt)

    def addDensity(self, x, y, a):
        self.L.ora_add_density(self.h_, x, y, a)

    def addVelocity(self, x, y, ax, ay):
        self.L.ora_add_velocity(self.h_, x, y, ax, ay)

    def reset(self):
        self.L.ora_reset(self.h_)

    def read(self, fid):
        out = np.empty((self.w, self.h), dtype=np.float32)
        self.L.ora_read_field(self.h_, fid, out.ctypes.data_as(_fp))
        return out

    def write(self, fid, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float32)
        assert arr.shape == (self.w, self.h)
        self.L.ora_write_field(self.h_, fid, arr.ctypes.data_as(_fp))

    def fields(self):
        return self.read(DENSITY), self.read(VEL_X), self.read(VEL_Y)

    def getDensity(self, x, y):
        if 0 <= x < self.w and 0 <= y < self.h:
            return self.read(DENSITY)[x, y]
        return np.float32(0)

    def setBoundary(self, b, fid):
        self.L.ora_set_boundary(self.h_, b, fid)

    def diffuse(self, b, dst, src, diff, dt):
        self.L.ora_diffuse(self.h_, b, dst, src, diff, dt)

    def project(self, u=VEL_X, v=VEL_Y, p=PREV_VEL_X, div=PREV_VEL_Y):
        self.L.ora_project(self.h_, u, v, p, div)

    def advect(self, b, dst, src, u, v, dt):
        self.L.ora_advect(self.h_, b, dst, src, u, v, dt)

    def addNoise(self):
        self.L.ora_add_noise(self.h_)

    def rand(self):
        return self.L.ora_rand(self.h_)

    def bilinear(self, fid, x, y):
        return np.float32(self.L.ora_bilinear(self.h_, fid, x, y))

    def getVelocityMagnitude(self, x, y):
        return np.float32(self.L.ora_velocity_magnitude(self.h_, x, y))

    def getNormalizedVelocity(self, x, y):
        out = (C.c_float * 2)()
        self.L.ora_normalized_velocity(self.h_, x, y, out)
        return tuple(np.float32(v) for v in out)

    def getVelocityColor(self, x, y):
        out = (C.c_float * 3)()
        self.L.ora_velocity_color(self.h_, x, y, out)
        return tuple(np.float32(v) for v in out)
