"""fluid_sim_b200 -- Python mirror of the reference's FluidSim class on top of the C ABI.

The product is the shared library ``libfluidsim_b200.so`` (hand-written sm_100a CUDA kernels
behind the ``extern "C"`` entry points of ``include/fluid_sim_b200.h``).  This module is a thin
ctypes binding whose method names, argument meaning and error behaviour follow the reference
class (``/root/reference/include/fluid_sim.h:20-37``): out-of-range adds are ignored, out-of-range
reads return 0.  There is no CPU fallback: if the library is missing or no CUDA device is present,
construction raises.
"""
import ctypes as C
import os

import numpy as np

__all__ = ["FluidSim", "FsbError", "load_library", "library_path", "build_library",
           "DENSITY", "VELOCITY_X", "VELOCITY_Y", "PREV_DENSITY", "PREV_VEL_X", "PREV_VEL_Y",
           "ADVECT_RK4", "ADVECT_MACCORMACK", "ADVECT_SEMI_LAGRANGIAN", "NOISE_GLIBC", "NOISE_ZERO"]

_HERE = os.path.dirname(os.path.abspath(__file__))

DENSITY, VELOCITY_X, VELOCITY_Y, PREV_DENSITY, PREV_VEL_X, PREV_VEL_Y = range(6)
ADVECT_RK4, ADVECT_MACCORMACK, ADVECT_SEMI_LAGRANGIAN = 0, 1, 2
NOISE_GLIBC, NOISE_ZERO = 0, 1

_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


class FsbError(RuntimeError):
    pass


class Params(C.Structure):
    """struct fsb_params of include/fluid_sim_b200.h"""
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("proj_iters", C.c_int), ("diff_iters", C.c_int),
                ("viscosity", C.c_float), ("diffusion", C.c_float), ("omega", C.c_float),
                ("advect_scheme", C.c_int), ("noise_mode", C.c_int), ("noise_seed", C.c_uint32),
                ("noise_pos", C.c_uint64), ("device", C.c_int), ("noise_process_stream", C.c_int),
                ("reserved", C.c_int * 6)]


def library_path():
    # FSB_LIBRARY selects an experimental build of the same ABI (development only)
    return os.environ.get("FSB_LIBRARY") or os.path.join(_HERE, "libfluidsim_b200.so")


def build_library(verbose=False):
    """Compile the CUDA sources in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    import subprocess
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "-j4"], capture_output=not verbose, text=True)
    if r.returncode != 0:
        raise FsbError("building libfluidsim_b200.so failed:\n" + (r.stdout or "") + (r.stderr or ""))
    return library_path()


_lib = None

# every extern "C" symbol include/fluid_sim_b200.h declares: (name, restype, argtypes)
_SIM = C.c_void_p
ABI = [
    ("fsb_default_params", None, [C.POINTER(Params)]),
    ("fsb_create", C.c_int, [C.POINTER(Params), C.POINTER(_SIM)]),
    ("fsb_destroy", C.c_int, [_SIM]),
    ("fsb_reset", C.c_int, [_SIM]),
    ("fsb_step", C.c_int, [_SIM, C.c_float]),
    ("fsb_sync", C.c_int, [_SIM]),
    ("fsb_add_density", C.c_int, [_SIM, C.c_int, C.c_int, C.c_float]),
    ("fsb_add_velocity", C.c_int, [_SIM, C.c_int, C.c_int, C.c_float, C.c_float]),
    ("fsb_add_density_n", C.c_int, [_SIM, C.c_int, _ip, _ip, _fp]),
    ("fsb_add_velocity_n", C.c_int, [_SIM, C.c_int, _ip, _ip, _fp, _fp]),
    ("fsb_get_density", C.c_int, [_SIM, C.c_int, C.c_int, _fp]),
    ("fsb_get_velocity", C.c_int, [_SIM, C.c_int, C.c_int, _fp]),
    ("fsb_get_normalized_velocity", C.c_int, [_SIM, C.c_int, C.c_int, _fp]),
    ("fsb_get_velocity_magnitude", C.c_int, [_SIM, C.c_int, C.c_int, _fp]),
    ("fsb_get_velocity_color", C.c_int, [_SIM, C.c_int, C.c_int, _fp]),
    ("fsb_get_width", C.c_int, [_SIM]),
    ("fsb_get_height", C.c_int, [_SIM]),
    ("fsb_read_field", C.c_int, [_SIM, C.c_int, C.c_void_p]),
    ("fsb_read_field_async", C.c_int, [_SIM, C.c_int, C.c_void_p]),
    ("fsb_read_wait", C.c_int, [_SIM]),
    ("fsb_view_density", C.c_int, [_SIM, C.c_int, C.c_int, C.c_void_p]),
    ("fsb_view_density_device", C.c_int, [_SIM, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    ("fsb_view_velocity", C.c_int, [_SIM, C.c_int, C.c_int, C.c_void_p]),
    ("fsb_write_field", C.c_int, [_SIM, C.c_int, C.c_void_p]),
    ("fsb_add_field", C.c_int, [_SIM, C.c_int, C.c_void_p]),
    ("fsb_host_alloc", C.c_void_p, [C.c_size_t]),
    ("fsb_host_free", None, [C.c_void_p]),
    ("fsb_device_field", C.c_int, [_SIM, C.c_int, C.POINTER(C.c_void_p)]),
    ("fsb_get_noise_pos", C.c_int, [_SIM, C.POINTER(C.c_uint64)]),
    ("fsb_set_noise", C.c_int, [_SIM, C.c_int, C.c_uint32, C.c_uint64]),
    ("fsb_op_add_noise", C.c_int, [_SIM]),
    ("fsb_op_set_boundary", C.c_int, [_SIM, C.c_int, C.c_int]),
    ("fsb_op_diffuse", C.c_int, [_SIM, C.c_int, C.c_int, C.c_float, C.c_float]),
    ("fsb_op_project", C.c_int, [_SIM]),
    ("fsb_op_advect", C.c_int, [_SIM, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]),
    ("fsb_op_velocity_step", C.c_int, [_SIM, C.c_float]),
    ("fsb_op_density_step", C.c_int, [_SIM, C.c_float]),
    ("fsb_get_stream", C.c_int, [_SIM, C.POINTER(C.c_void_p)]),
    ("fsb_profile_enable", C.c_int, [_SIM, C.c_int]),
    ("fsb_profile_read", C.c_int, [_SIM, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_int]),
    ("fsb_chain_trace", C.c_int, [_SIM, C.c_int, C.POINTER(C.c_uint64), C.c_int, C.c_int, _ip]),
    ("fsb_create_block", C.c_int, [C.POINTER(Params), C.c_int, C.c_int, C.c_int, C.POINTER(_SIM)]),
    ("fsb_block_info", C.c_int, [_SIM, _ip, _ip, _ip, _ip]),
    ("fsb_link_handle", C.c_int, [_SIM, C.c_void_p]),
    ("fsb_link_connect", C.c_int, [_SIM, C.c_void_p, C.c_void_p]),
    ("fsb_field_handles", C.c_int, [_SIM, C.c_void_p]),
    ("fsb_link_connect_fields", C.c_int, [_SIM, C.c_void_p, C.c_int]),
    ("fsb_blk_wait_returns", C.c_int, [_SIM, C.c_int, C.c_uint64]),
    ("fsb_blk_sweep", C.c_int, [_SIM, C.c_int, C.c_int, _ip, _ip, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_uint64, _ip]),
    ("fsb_blk_add_zero", C.c_int, [_SIM, C.c_int, _ip]),
    ("fsb_blk_set_boundary", C.c_int, [_SIM, C.c_int, _ip, _ip, _ip]),
    ("fsb_blk_divergence", C.c_int, [_SIM]),
    ("fsb_blk_gradient", C.c_int, [_SIM]),
    ("fsb_blk_vmax_word", C.c_int, [_SIM, C.POINTER(C.c_void_p)]),
    ("fsb_blk_chain_prepare", C.c_int, [_SIM, C.c_float]),
    ("fsb_blk_chain_handles", C.c_int, [_SIM, C.c_void_p]),
    ("fsb_blk_chain_connect", C.c_int, [_SIM, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("fsb_blk_step", C.c_int, [_SIM, C.c_float]),
    ("fsb_blk_chain_error", C.c_int, [_SIM, C.POINTER(C.c_uint64)]),
    ("fsb_blk_advect_velocity", C.c_int, [_SIM, C.c_float, C.c_int]),
    ("fsb_blk_advect_density", C.c_int, [_SIM, C.c_float, C.c_int]),
    ("fsb_pressure_sweeps_noalloc", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    ("fsb_launch_count", C.c_int, [_SIM, C.POINTER(C.c_uint64)]),
    ("fsb_set_tuning", C.c_int, [_SIM, C.c_int, C.c_int, C.c_int]),
    ("fsb_set_chain_level", C.c_int, [_SIM, C.c_int]),
    ("fsb_set_chain_overlap", C.c_int, [_SIM, C.c_int]),
    ("fsb_dbg_chain_describe", C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.POINTER(C.c_longlong), C.c_int,
                                         C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    ("fsb_last_error", C.c_char_p, []),
    ("fsb_version", C.c_char_p, []),
]


def load_library():
    """dlopen libfluidsim_b200.so and type every entry point.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise FsbError(f"{path} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(there is no CPU fallback)")
    lib = C.CDLL(path)
    for name, res, args in ABI:
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class FluidSim:
    """Drop-in for the reference's ``FluidSim`` (same method names), running on one B200."""

    def __init__(self, width=200, height=200, proj_iters=20, diff_iters=5, viscosity=1e-4, diffusion=0.0,
                 omega=1.5, advect_scheme=ADVECT_RK4, noise="glibc", noise_seed=1, noise_pos=0, device=-1,
                 strips_per_cta=0, sweeps_per_warp=0, max_stages=0, process_stream=False, chain=None, overlap=None):
        self._lib = load_library()
        p = Params()
        self._lib.fsb_default_params(C.byref(p))
        p.width, p.height, p.proj_iters, p.diff_iters = width, height, proj_iters, diff_iters
        p.viscosity, p.diffusion, p.omega = viscosity, diffusion, omega
        p.advect_scheme = advect_scheme
        p.noise_mode = NOISE_GLIBC if noise == "glibc" else NOISE_ZERO
        p.noise_seed, p.noise_pos, p.device = noise_seed, noise_pos, device
        p.noise_process_stream = 1 if process_stream else 0
        self._h = _SIM()
        self.width, self.height = width, height
        self._check(self._lib.fsb_create(C.byref(p), C.byref(self._h)))
        if strips_per_cta or sweeps_per_warp or max_stages:
            self._check(self._lib.fsb_set_tuning(self._h, strips_per_cta, sweeps_per_warp, max_stages))
        if chain is not None:
            self._check(self._lib.fsb_set_chain_level(self._h, chain))
        if overlap is not None:
            self._check(self._lib.fsb_set_chain_overlap(self._h, 1 if overlap else 0))

    # -- plumbing ---------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise FsbError(f"fluid_sim_b200 error {rc}: {self._lib.fsb_last_error().decode()}")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.fsb_destroy(self._h)
            self._h = _SIM()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- the reference's public interface (include/fluid_sim.h:20-37) ---------------------------
    def step(self, dt=0.01):
        self._check(self._lib.fsb_step(self._h, dt))

    def addDensity(self, x, y, amount):
        self._check(self._lib.fsb_add_density(self._h, x, y, amount))

    def addVelocity(self, x, y, amountX, amountY):
        self._check(self._lib.fsb_add_velocity(self._h, x, y, amountX, amountY))

    def getDensity(self, x, y):
        out = C.c_float()
        self._check(self._lib.fsb_get_density(self._h, x, y, C.byref(out)))
        return np.float32(out.value)

    def _vec(self, fn, x, y, n):
        out = (C.c_float * n)()
        self._check(fn(self._h, x, y, out))
        return tuple(np.float32(v) for v in out)

    def getVelocity(self, x, y):
        return self._vec(self._lib.fsb_get_velocity, x, y, 2)

    def getNormalizedVelocity(self, x, y):
        return self._vec(self._lib.fsb_get_normalized_velocity, x, y, 2)

    def getVelocityMagnitude(self, x, y):
        out = C.c_float()
        self._check(self._lib.fsb_get_velocity_magnitude(self._h, x, y, C.byref(out)))
        return np.float32(out.value)

    def getVelocityColor(self, x, y):
        return self._vec(self._lib.fsb_get_velocity_color, x, y, 3)

    def getWidth(self):
        return self._lib.fsb_get_width(self._h)

    def getHeight(self):
        return self._lib.fsb_get_height(self._h)

    def reset(self):
        """``sim = FluidSim();`` of src/main.cpp:289."""
        self._check(self._lib.fsb_reset(self._h))

    # -- batched / bulk forms ------------------------------------------------------------------
    def addDensityN(self, xs, ys, amounts):
        xs = np.ascontiguousarray(xs, np.int32); ys = np.ascontiguousarray(ys, np.int32)
        am = np.ascontiguousarray(amounts, np.float32)
        self._check(self._lib.fsb_add_density_n(self._h, len(xs), xs.ctypes.data_as(_ip), ys.ctypes.data_as(_ip),
                                                am.ctypes.data_as(_fp)))

    def addVelocityN(self, xs, ys, axs, ays):
        xs = np.ascontiguousarray(xs, np.int32); ys = np.ascontiguousarray(ys, np.int32)
        ax = np.ascontiguousarray(axs, np.float32); ay = np.ascontiguousarray(ays, np.float32)
        self._check(self._lib.fsb_add_velocity_n(self._h, len(xs), xs.ctypes.data_as(_ip), ys.ctypes.data_as(_ip),
                                                 ax.ctypes.data_as(_fp), ay.ctypes.data_as(_fp)))

    def sync(self):
        self._check(self._lib.fsb_sync(self._h))

    def read(self, field, out=None):
        if out is None:
            out = np.empty((self.width, self.height), np.float32)
        assert out.dtype == np.float32 and out.size == self.width * self.height and out.flags.c_contiguous
        self._check(self._lib.fsb_read_field(self._h, field, out.ctypes.data))
        return out

    def write(self, field, arr):
        arr = np.ascontiguousarray(arr, np.float32)
        assert arr.size == self.width * self.height
        self._check(self._lib.fsb_write_field(self._h, field, arr.ctypes.data))

    def add_field(self, field, arr):
        arr = np.ascontiguousarray(arr, np.float32)
        assert arr.size == self.width * self.height
        self._check(self._lib.fsb_add_field(self._h, field, arr.ctypes.data))

    def read_async(self, field, out):
        """Queue the device-to-host copy of a field into `out` (a pinned_array()); later steps overlap with it."""
        assert out.dtype == np.float32 and out.size == self.width * self.height and out.flags.c_contiguous
        self._check(self._lib.fsb_read_field_async(self._h, field, out.ctypes.data))

    def read_wait(self):
        self._check(self._lib.fsb_read_wait(self._h))

    def view_density(self, px_w=800, px_h=600, out=None):
        """The reference viewer's density image (src/main.cpp:388-427 + density shader) as [px_h, px_w, 4] uint8,
        row 0 = bottom window row."""
        if out is None:
            out = np.empty((px_h, px_w, 4), np.uint8)
        assert out.dtype == np.uint8 and out.size == px_w * px_h * 4 and out.flags.c_contiguous
        self._check(self._lib.fsb_view_density(self._h, px_w, px_h, out.ctypes.data))
        return out

    def view_velocity(self, px_w=800, px_h=600, out=None):
        """The viewer's image with velocity vectors on (src/main.cpp:452-514): density image + glyph lines."""
        if out is None:
            out = np.empty((px_h, px_w, 4), np.uint8)
        assert out.dtype == np.uint8 and out.size == px_w * px_h * 4 and out.flags.c_contiguous
        self._check(self._lib.fsb_view_velocity(self._h, px_w, px_h, out.ctypes.data))
        return out

    def fields(self):
        return self.read(DENSITY), self.read(VELOCITY_X), self.read(VELOCITY_Y)

    def pinned_array(self, shape=None):
        """float32 numpy array backed by page-locked host memory (freed with the process)."""
        shape = shape or (self.width, self.height)
        n = int(np.prod(shape))
        p = self._lib.fsb_host_alloc(n * 4)
        if not p:
            raise FsbError(self._lib.fsb_last_error().decode())
        buf = (C.c_float * n).from_address(p)
        return np.frombuffer(buf, np.float32).reshape(shape)

    def device_field(self, field):
        p = C.c_void_p()
        self._check(self._lib.fsb_device_field(self._h, field, C.byref(p)))
        return p.value

    def noise_pos(self):
        v = C.c_uint64()
        self._check(self._lib.fsb_get_noise_pos(self._h, C.byref(v)))
        return v.value

    def set_noise(self, mode="glibc", seed=1, pos=0):
        self._check(self._lib.fsb_set_noise(self._h, NOISE_GLIBC if mode == "glibc" else NOISE_ZERO, seed, pos))

    def launch_count(self):
        v = C.c_uint64()
        self._check(self._lib.fsb_launch_count(self._h, C.byref(v)))
        return v.value

    def stream(self):
        p = C.c_void_p()
        self._check(self._lib.fsb_get_stream(self._h, C.byref(p)))
        return p.value

    PROF_NAMES = ("noise", "sor", "gs", "divergence", "gradient", "advect_velocity", "advect_scalar", "boundary", "chain")

    def profile_enable(self, on=True):
        self._check(self._lib.fsb_profile_enable(self._h, 1 if on else 0))

    def profile_read(self, reset=True):
        """{class: (total_ms, launches)} accumulated since the last reset; synchronises."""
        n = len(self.PROF_NAMES)
        ms = (C.c_double * n)()
        cnt = (C.c_uint64 * n)()
        self._check(self._lib.fsb_profile_read(self._h, ms, cnt, 1 if reset else 0))
        return {name: (ms[i], cnt[i]) for i, name in enumerate(self.PROF_NAMES)}

    def chain_trace(self, which=0, max_phases=12, max_items=160):
        """[phases, max_items, 3] uint64 ns (wait begins, work begins, results published) of the last chained launch."""
        out = np.zeros((max_phases, max_items, 3), np.uint64)
        n = C.c_int()
        self._check(self._lib.fsb_chain_trace(self._h, which, out.ctypes.data_as(C.POINTER(C.c_uint64)), max_phases, max_items,
                                              C.byref(n)))
        return out[:n.value]

    def set_tuning(self, strips_per_cta=0, sweeps_per_warp=0, max_stages=0):
        self._check(self._lib.fsb_set_tuning(self._h, strips_per_cta, sweeps_per_warp, max_stages))

    # -- the reference's private passes (per-kernel parity) -------------------------------------------
    def addNoise(self):
        self._check(self._lib.fsb_op_add_noise(self._h))

    def setBoundary(self, b, field):
        self._check(self._lib.fsb_op_set_boundary(self._h, b, field))

    def diffuse(self, b, field, diff, dt):
        self._check(self._lib.fsb_op_diffuse(self._h, b, field, diff, dt))

    def project(self):
        self._check(self._lib.fsb_op_project(self._h))

    def advect(self, b, dst, src, u, v, dt):
        self._check(self._lib.fsb_op_advect(self._h, b, dst, src, u, v, dt))

    def velocityStep(self, dt):
        self._check(self._lib.fsb_op_velocity_step(self._h, dt))

    def densityStep(self, dt):
        self._check(self._lib.fsb_op_density_step(self._h, dt))
