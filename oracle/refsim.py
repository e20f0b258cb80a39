"""TEST INFRASTRUCTURE ONLY -- ctypes front end to the real reference built by
oracle/build_ref.py (oracle/_ref/libref_<W>x<H>_p<P>_d<D>.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  It never touches /root/reference at run time; it only
loads prebuilt libraries, so it also works on the GPU box.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

# field ids of oracle/ref_shim.cpp::field_of
DENSITY, VEL_X, VEL_Y, PREV_DENSITY, PREV_VEL_X, PREV_VEL_Y, TEMP1, TEMP2 = range(8)

_fp = C.POINTER(C.c_float)


def lib_path(w, h, p=20, d=5):
    return os.path.join(REF_DIR, f"libref_{w}x{h}_p{p}_d{d}.so")


def available(w, h, p=20, d=5):
    return os.path.exists(lib_path(w, h, p, d))


_libs = {}


def _load(w, h, p, d):
    key = (w, h, p, d)
    if key in _libs:
        return _libs[key]
    lib = C.CDLL(lib_path(w, h, p, d))
    lib.ref_create.restype = C.c_void_p
    lib.ref_destroy.argtypes = [C.c_void_p]
    lib.ref_reset.argtypes = [C.c_void_p]
    lib.ref_srand.argtypes = [C.c_uint]
    lib.ref_noise_mode.argtypes = [C.c_int]
    lib.ref_rand.restype = C.c_int
    lib.ref_step.argtypes = [C.c_void_p, C.c_float]
    lib.ref_add_density.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float]
    lib.ref_add_velocity.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float]
    lib.ref_get_density.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.ref_get_density.restype = C.c_float
    for name in ("ref_get_velocity", "ref_get_normalized_velocity", "ref_get_velocity_color"):
        getattr(lib, name).argtypes = [C.c_void_p, C.c_int, C.c_int, _fp]
    lib.ref_get_velocity_magnitude.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.ref_get_velocity_magnitude.restype = C.c_float
    lib.ref_get_width.argtypes = [C.c_void_p]
    lib.ref_get_height.argtypes = [C.c_void_p]
    lib.ref_read_field.argtypes = [C.c_void_p, C.c_int, _fp]
    lib.ref_write_field.argtypes = [C.c_void_p, C.c_int, _fp]
    lib.ref_set_boundary.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.ref_diffuse.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float]
    lib.ref_project.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.ref_advect.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
    lib.ref_advect_scheme.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
    lib.ref_bilinear.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float]
    lib.ref_bilinear.restype = C.c_float
    lib.ref_velocity_step.argtypes = [C.c_void_p, C.c_float]
    lib.ref_density_step.argtypes = [C.c_void_p, C.c_float]
    assert (lib.ref_grid_w(), lib.ref_grid_h(), lib.ref_proj_iters(), lib.ref_diff_iters()) == key
    _libs[key] = lib
    return lib


class RefSim:
    """The reference's FluidSim (include/fluid_sim.h:17-37), same method names."""

    def __init__(self, w=200, h=200, p=20, d=5, seed=1, noise="glibc"):
        self.lib = _load(w, h, p, d)
        self.w, self.h, self.p, self.d = w, h, p, d
        self.h_ = self.lib.ref_create()
        # rand() state is process-global in the reference (never seeded: default seed 1).
        self.lib.ref_noise_mode(0 if noise == "glibc" else 1)
        if seed is not None:
            self.lib.ref_srand(seed)

    def close(self):
        if self.h_:
            self.lib.ref_destroy(self.h_)
            self.h_ = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # public API -------------------------------------------------------------
    def step(self, dt=0.01):
        self.lib.ref_step(self.h_, dt)

    def addDensity(self, x, y, a):
        self.lib.ref_add_density(self.h_, x, y, a)

    def addVelocity(self, x, y, ax, ay):
        self.lib.ref_add_velocity(self.h_, x, y, ax, ay)

    def getDensity(self, x, y):
        return self.lib.ref_get_density(self.h_, x, y)

    def _vec(self, fn, x, y, n):
        out = (C.c_float * n)()
        fn(self.h_, x, y, out)
        return tuple(np.float32(v) for v in out)

    def getVelocity(self, x, y):
        return self._vec(self.lib.ref_get_velocity, x, y, 2)

    def getNormalizedVelocity(self, x, y):
        return self._vec(self.lib.ref_get_normalized_velocity, x, y, 2)

    def getVelocityMagnitude(self, x, y):
        return np.float32(self.lib.ref_get_velocity_magnitude(self.h_, x, y))

    def getVelocityColor(self, x, y):
        return self._vec(self.lib.ref_get_velocity_color, x, y, 3)

    def getWidth(self):
        return self.lib.ref_get_width(self.h_)

    def getHeight(self):
        return self.lib.ref_get_height(self.h_)

    def reset(self):
        self.lib.ref_reset(self.h_)

    # raw access ---------------------------------------------------------------
    def read(self, fid):
        out = np.empty((self.w, self.h), dtype=np.float32)
        self.lib.ref_read_field(self.h_, fid, out.ctypes.data_as(_fp))
        return out

    def write(self, fid, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float32)
        assert arr.shape == (self.w, self.h)
        self.lib.ref_write_field(self.h_, fid, arr.ctypes.data_as(_fp))

    def fields(self):
        return self.read(DENSITY), self.read(VEL_X), self.read(VEL_Y)

    # private methods ------------------------------------------------------------
    def setBoundary(self, b, fid):
        self.lib.ref_set_boundary(self.h_, b, fid)

    def diffuse(self, b, dst, src, diff, dt):
        self.lib.ref_diffuse(self.h_, b, dst, src, diff, dt)

    def project(self, u=VEL_X, v=VEL_Y, p=PREV_VEL_X, div=PREV_VEL_Y):
        self.lib.ref_project(self.h_, u, v, p, div)

    def advect(self, b, dst, src, u, v, dt, scheme=None):
        if scheme is None:
            self.lib.ref_advect(self.h_, b, dst, src, u, v, dt)
        else:
            self.lib.ref_advect_scheme(self.h_, scheme, b, dst, src, u, v, dt)

    def bilinear(self, fid, x, y):
        return np.float32(self.lib.ref_bilinear(self.h_, fid, x, y))

    def velocityStep(self, dt):
        self.lib.ref_velocity_step(self.h_, dt)

    def densityStep(self, dt):
        self.lib.ref_density_step(self.h_, dt)


def demo_injection(sim, w, h):
    """Start-up injection of /root/reference/src/main.cpp:351-363 (69 cells at 200x200)."""
    cx, cy = w // 2, h // 2
    n = 0
    for i in range(-5, 6):
        for j in range(-5, 6):
            if i * i + j * j < 25:
                sim.addDensity(cx + i, cy + j, 10.0)
                sim.addVelocity(cx + i, cy + j, 0.0, 2.0)
                n += 1
    return n
