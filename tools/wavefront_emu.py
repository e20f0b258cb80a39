"""Development aid: numpy emulation of the skewed, time-pipelined strip sweep that
fluid_sim_b200/csrc/wavefront.cuh implements on the GPU.

Global lane g owns skewed columns: at global time t, level k of lane g sits on cell
(i, j) = (t - k - g, g - k).  Every level of every lane advances one row per step and
all updates of a step read only values of step t-1 (and t-2), which is what makes the
lexicographic Gauss-Seidel / SOR order (src/fluid_sim.cpp:52-59, :321-329) reproducible
bit for bit with S sweeps in flight.  Used by tests/test_wavefront_emu.py.
"""
import numpy as np

f32 = np.float32


def set_bnd(b, x):
    """src/fluid_sim.cpp:347-367 on a numpy array (in place)."""
    W, H = x.shape
    x[1:W - 1, 0] = -x[1:W - 1, 1] if b == 2 else x[1:W - 1, 1]
    x[1:W - 1, H - 1] = -x[1:W - 1, H - 2] if b == 2 else x[1:W - 1, H - 2]
    x[0, 1:H - 1] = -x[1, 1:H - 1] if b == 1 else x[1, 1:H - 1]
    x[W - 1, 1:H - 1] = -x[W - 2, 1:H - 1] if b == 1 else x[W - 2, 1:H - 1]
    x[0, 0] = f32(0.5) * (x[1, 0] + x[0, 1])
    x[0, H - 1] = f32(0.5) * (x[1, H - 1] + x[0, H - 2])
    x[W - 1, 0] = f32(0.5) * (x[W - 2, 0] + x[W - 1, 1])
    x[W - 1, H - 1] = f32(0.5) * (x[W - 2, H - 1] + x[W - 1, H - 2])


def sweep_pass(x0, rhs, S, b, mode, a=f32(0), c_recip=f32(1), omega=f32(1.5)):
    """Apply S lexicographic sweeps (mode 'gs' or 'sor') to the interior of x0, boundary
    cells handled as the reference's per-sweep setBoundary would; returns the new array
    with walls/corners still to be finalised by set_bnd()."""
    W, H = rhs.shape
    if x0 is None:
        x0 = np.zeros((W, H), f32)
    out = x0.copy()
    G = ((H + S + 31) // 32) * 32           # global lanes (strips of 32)
    si = f32(-1) if b == 1 else f32(1)      # sign on the i-walls (rows 0, W-1)
    sj = f32(-1) if b == 2 else f32(1)      # sign on the j-walls (cols 0, H-1)
    x = np.zeros((S + 1, G), f32)           # x[k, g] after the previous step
    Uprev = np.zeros((S + 1, G), f32)       # shift(x) of two steps ago
    Dm1 = np.zeros((S + 1, G), f32)         # D of step t-1
    Dm2 = np.zeros((S + 1, G), f32)         # D of step t-2
    g = np.arange(G)
    kk = np.arange(S + 1)[:, None]
    jj = g[None, :] - kk                    # column of (k, g), fixed over time
    T = (G - 1) + (W - 1) + S + 1
    for t in range(T + 1):
        ii = t - kk - g[None, :]            # row of (k, g) at this step
        U = np.zeros_like(x)
        U[:, 1:] = x[:, :-1]
        D = np.zeros_like(x)
        D[1:, 1:] = Dm2[:-1, :-1]
        # level 0: load inputs
        i0 = ii[0]
        j0 = jj[0]
        valid0 = (i0 >= 0) & (i0 < W) & (j0 >= 0) & (j0 < H)
        ic = np.clip(i0, 0, W - 1)
        jc = np.clip(j0, 0, H - 1)
        newx = np.zeros_like(x)
        newx[0] = np.where(valid0, x0[ic, jc], f32(0))
        D[0] = np.where(valid0, rhs[ic, jc], f32(0))
        # levels 1..S
        down = U[:-1]       # (i+1, j) level k-1
        up = x[1:]          # (i-1, j) level k
        right = x[:-1]      # (i, j+1) level k-1
        left = U[1:]        # (i, j-1) level k
        cold = Uprev[:-1]   # (i, j)   level k-1
        d = D[1:]
        with np.errstate(all="ignore"):
            if mode == "gs":
                val = ((((d + down) + up) + right) + left) * f32(0.25)
            else:
                nb = ((down + up) + right) + left
                nv = (d + a * nb) * c_recip
                val = cold + omega * (nv - cold)
        i = ii[1:]
        j = np.broadcast_to(jj[1:], i.shape)
        k = np.broadcast_to(kk[1:], i.shape)
        inrow = (i >= 1) & (i <= W - 2)
        incol = (j >= 1) & (j <= H - 2)
        res = np.zeros_like(val)
        res = np.where(inrow & incol, val, res)
        first = (k == 1)
        res = np.where((i == 0) & incol, np.where(first, cold, si * down), res)        # top wall
        res = np.where((i == W - 1) & incol, si * up, res)                             # bottom wall
        res = np.where((j == 0) & inrow, np.where(first, cold, sj * right), res)       # left wall
        res = np.where((j == H - 1) & inrow, sj * left, res)                           # right wall
        newx[1:] = res
        # output: level S on interior cells
        iS = ii[S]
        jS = jj[S]
        ok = (iS >= 1) & (iS <= W - 2) & (jS >= 1) & (jS <= H - 2)
        out[iS[ok], jS[ok]] = newx[S][ok]
        Uprev = U
        x = newx
        Dm2 = Dm1
        Dm1 = D
    return out


def diffuse(x, b, diff, dt, D, omega=f32(1.5), chunks=None):
    """src/fluid_sim.cpp:43-62 with the sweeps executed by sweep_pass (src == initial x)."""
    W, H = x.shape
    a = f32(f32(f32(dt) * f32(diff)) * f32(W)) * f32(H)
    c_recip = f32(1.0) / (f32(1) + f32(4) * a)
    src = x.copy()
    cur = x.copy()
    for S in (chunks or [D]):
        cur = sweep_pass(cur, src, S, b, "sor", a, c_recip, f32(omega))
        set_bnd(b, cur)
    return cur


def project(u, v, P, chunks=None):
    """src/fluid_sim.cpp:301-344 with the Gauss-Seidel sweeps executed by sweep_pass."""
    W, H = u.shape
    h = f32(1.0) / f32(W)
    div = np.zeros((W, H), f32)
    div[1:-1, 1:-1] = (f32(-0.5) * h) * (((u[2:, 1:-1] - u[:-2, 1:-1]) + v[1:-1, 2:]) - v[1:-1, :-2])
    p = None
    for S in (chunks or [P]):
        p = sweep_pass(p, div, S, 0, "gs")
        set_bnd(0, p)
    u = u.copy()
    v = v.copy()
    u[1:-1, 1:-1] -= (f32(0.5) * (p[2:, 1:-1] - p[:-2, 1:-1])) / h
    v[1:-1, 1:-1] -= (f32(0.5) * (p[1:-1, 2:] - p[1:-1, :-2])) / h
    set_bnd(1, u)
    set_bnd(2, v)
    return u, v, p
