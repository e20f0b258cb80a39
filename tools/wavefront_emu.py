"""Development aid: numpy emulation of the skewed, time-pipelined strip sweep that
fluid_sim_b200/csrc/wavefront.cu implements on the GPU.

Global lane g owns skewed columns: at global time t, level k of lane g sits on cell
(i, j) = (t - k - g, g - k).  Every level of every lane advances one row per step and
all updates of a step read only values of step t-1 (and t-2), which is what makes the
lexicographic Gauss-Seidel / SOR order (src/fluid_sim.cpp:52-59, :321-329) reproducible
bit for bit with S sweeps in flight.  Checked by tests/test_wavefront_emu.py against plain lexicographic loops.
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


def sweep_pass_2col(x0, rhs, S, b, mode, a=f32(0), c_recip=f32(1), omega=f32(1.5)):
    """The same S sweeps with TWO columns per lane (groundwork for the next kernel generation, not yet on the GPU).

    Lane g, level k owns the column pair J = g - k (columns A = 2J and B = 2J + 1) on row i = t - k - g: a cell is
    updated at time i + J + 2k, so the wave crosses two columns per step and the chain of hand-overs along the top row
    is half as long.  Neighbours of A: up = own A (t-1), left = lane g-1's B (t-1), down = lane g-1's A of level k-1
    (t-1), right = lane g-1's B of level k-1 (t-2).  Neighbours of B: up = own B (t-1), left = the A just computed,
    down = lane g-1's B of level k-1 (t-1), right = this lane's own A of level k-1 before this step's update."""
    W, H = rhs.shape
    if x0 is None:
        x0 = np.zeros((W, H), f32)
    out = x0.copy()
    G = (((H + 1) // 2 + S + 31) // 32) * 32          # lanes = column pairs
    si = f32(-1) if b == 1 else f32(1)
    sj = f32(-1) if b == 2 else f32(1)
    z = lambda: np.zeros((S + 1, G), f32)
    xA, xB, UpA, UpB = z(), z(), z(), z()
    DA1, DB1, DA2, DB2 = z(), z(), z(), z()
    g = np.arange(G)
    kk = np.arange(S + 1)[:, None]
    J = g[None, :] - kk
    jA, jB = 2 * J, 2 * J + 1
    T = (G - 1) + (W - 1) + S + 1

    def shift(v):
        u = np.zeros_like(v)
        u[:, 1:] = v[:, :-1]
        return u

    def load(arr, i, j):
        ok = (i >= 0) & (i < W) & (j >= 0) & (j < H)
        return np.where(ok, arr[np.clip(i, 0, W - 1), np.clip(j, 0, H - 1)], f32(0))

    def update(d, down, up, right, left, cold, i, j):
        with np.errstate(all="ignore"):
            if mode == "gs":
                val = ((((d + down) + up) + right) + left) * f32(0.25)
            else:
                nv = (d + a * (((down + up) + right) + left)) * c_recip
                val = cold + omega * (nv - cold)
        k = np.broadcast_to(kk[1:], i.shape)
        inrow, incol, first = (i >= 1) & (i <= W - 2), (j >= 1) & (j <= H - 2), (k == 1)
        res = np.where(inrow & incol, val, f32(0))
        res = np.where((i == 0) & incol, np.where(first, cold, si * down), res)
        res = np.where((i == W - 1) & incol, si * up, res)
        res = np.where((j == 0) & inrow, np.where(first, cold, sj * right), res)
        res = np.where((j == H - 1) & inrow, sj * left, res)
        return res

    for t in range(T + 1):
        ii = t - kk - g[None, :]
        UA, UB = shift(xA), shift(xB)
        DA, DB = z(), z()
        DA[1:, 1:] = DA2[:-1, :-1]
        DB[1:, 1:] = DB2[:-1, :-1]
        nA, nB = z(), z()
        nA[0], nB[0] = load(x0, ii[0], jA[0]), load(x0, ii[0], jB[0])
        DA[0], DB[0] = load(rhs, ii[0], jA[0]), load(rhs, ii[0], jB[0])
        i = ii[1:]
        resA = update(DA[1:], UA[:-1], xA[1:], UpB[:-1], UB[1:], UpA[:-1], i, np.broadcast_to(jA[1:], i.shape))
        resB = update(DB[1:], UB[:-1], xB[1:], xA[:-1], resA, UpB[:-1], i, np.broadcast_to(jB[1:], i.shape))
        nA[1:], nB[1:] = resA, resB
        for n_, j_ in ((nA, jA), (nB, jB)):
            ok = (ii[S] >= 1) & (ii[S] <= W - 2) & (j_[S] >= 1) & (j_[S] <= H - 2)
            out[ii[S][ok], j_[S][ok]] = n_[S][ok]
        UpA, UpB = UA, UB
        xA, xB = nA, nB
        DA2, DB2, DA1, DB1 = DA1, DB1, DA, DB
    return out


def diffuse(x, b, diff, dt, D, omega=f32(1.5), chunks=None, sweep=sweep_pass):
    """src/fluid_sim.cpp:43-62 with the sweeps executed by sweep_pass (src == initial x)."""
    W, H = x.shape
    a = f32(f32(f32(dt) * f32(diff)) * f32(W)) * f32(H)
    c_recip = f32(1.0) / (f32(1) + f32(4) * a)
    src = x.copy()
    cur = x.copy()
    for S in (chunks or [D]):
        cur = sweep(cur, src, S, b, "sor", a, c_recip, f32(omega))
        set_bnd(b, cur)
    return cur


def project(u, v, P, chunks=None, sweep=sweep_pass):
    """src/fluid_sim.cpp:301-344 with the Gauss-Seidel sweeps executed by sweep_pass."""
    W, H = u.shape
    h = f32(1.0) / f32(W)
    div = np.zeros((W, H), f32)
    div[1:-1, 1:-1] = (f32(-0.5) * h) * (((u[2:, 1:-1] - u[:-2, 1:-1]) + v[1:-1, 2:]) - v[1:-1, :-2])
    p = None
    for S in (chunks or [P]):
        p = sweep(p, div, S, 0, "gs")
        set_bnd(0, p)
    u = u.copy()
    v = v.copy()
    u[1:-1, 1:-1] -= (f32(0.5) * (p[2:, 1:-1] - p[:-2, 1:-1])) / h
    v[1:-1, 1:-1] -= (f32(0.5) * (p[1:-1, 2:] - p[1:-1, :-2])) / h
    set_bnd(1, u)
    set_bnd(2, v)
    return u, v, p
