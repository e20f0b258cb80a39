"""One FluidSim::step sharded over the GPUs of a box: every rank owns all rows of a block of columns.

The exact-order relaxations (diffuse / project, reference src/fluid_sim.cpp:50-61, :319-331) stay ONE chain of
column strips across all ranks: inside the sweep kernel the last CTA of a rank hands the strip boundary to the
first CTA of the next rank through peer memory (CUDA IPC over NVLink) with the same ring + progress-word protocol
that links CTAs on one GPU.  Everything else needs halo columns only; this module moves them between passes with
``torch.distributed`` (NCCL send/recv), which is plumbing: all arithmetic runs in libfluidsim_b200.so.

The result is bit-identical to the single-GPU run (tests/mgpu_check.py).
"""
import ctypes as C
import math

import numpy as np
import torch
import torch.distributed as dist

from . import (DENSITY, VELOCITY_X, VELOCITY_Y, PREV_DENSITY, PREV_VEL_X, PREV_VEL_Y, NOISE_GLIBC, NOISE_ZERO,
               FsbError, Params, load_library, _SIM, _ip)

__all__ = ["BlockFluidSim", "column_range", "plan_projection", "plan_diffusion", "exchange_halo", "return_columns"]


def column_range(height, rank, world, align=64):
    """Columns [begin, end) owned by `rank`; blocks are equal and a multiple of `align` (CTA granularity)."""
    if height % world:
        raise ValueError("height must be divisible by the number of ranks")
    hl = height // world
    if hl % align:
        raise ValueError(f"columns per rank ({hl}) must be a multiple of {align}")
    return rank * hl, (rank + 1) * hl


def plan_projection(sweeps, levels=10, max_stages=5):
    """Launches (levels per warp, stages) covering `sweeps` pressure sweeps; the same on every rank."""
    out, left = [], sweeps
    while left >= levels:
        st = min(left // levels, max_stages)
        out.append((levels, st))
        left -= levels * st
    for lv in (5, 2, 1):
        while left >= lv:
            out.append((lv, 1))
            left -= lv
    return out


def plan_diffusion(sweeps):
    """Launches (levels per warp, stages) covering `sweeps` diffusion sweeps: one launch for the counts that are an
    instantiated sweeps-per-warp size (the reference's 5 among them), otherwise the largest sizes first."""
    if sweeps in (1, 2, 5, 10, 20):
        return [(sweeps, 1)]
    out, left = [], sweeps
    for lv in (10, 5, 2, 1):
        while left >= lv:
            out.append((lv, 1))
            left -= lv
    return out


# ---- halo movement on plain tensors (runs on CPU tensors with gloo as well: tests/test_multi_host.py) -------
def _neighbours(rank, world):
    return (rank - 1 if rank > 0 else None), (rank + 1 if rank < world - 1 else None)


def exchange_halo(tensors, halo, ncols, width, rank, world, group=None):
    """Fill `width` halo columns on both sides of every [rows, pitch] tensor from the neighbours' owned columns.
    Owned columns are local [halo, halo + ncols)."""
    if world == 1 or width == 0:
        return
    left, right = _neighbours(rank, world)
    # Two separate send/recv groups, the left neighbour's first: the pair (r, r+1) is then r's second group and
    # r+1's first, so a rank waits only for the neighbour it really depends on.  One group holding both neighbours
    # would chain every rank to the slowest one (the sweep chain leaves the ranks staggered in time).
    for peer, send0, recv0 in ((left, halo, halo - width), (right, halo + ncols - width, halo + ncols)):
        if peer is None:
            continue
        ops, recvs = [], []
        for t in tensors:
            send = t[:, send0:send0 + width].contiguous()
            buf = torch.empty_like(send)
            ops += [dist.P2POp(dist.isend, send, peer, group), dist.P2POp(dist.irecv, buf, peer, group)]
            recvs.append((t, buf))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        for t, buf in recvs:
            t[:, recv0:recv0 + width] = buf


def return_columns(tensors, halo, ncols, width, rank, world, group=None):
    """After a sweep pass: the `width` result columns a rank computed for its LEFT neighbour (they sit in its left
    halo) go back to that neighbour's last owned columns.  Interior rows only; walls are redone afterwards."""
    if world == 1 or width == 0:
        return
    left, right = _neighbours(rank, world)
    ops, recvs = [], []
    for t in tensors:
        if left is not None:
            ops.append(dist.P2POp(dist.isend, t[1:-1, halo - width:halo].contiguous(), left, group))
        if right is not None:
            buf = torch.empty((t.shape[0] - 2, width), dtype=t.dtype, device=t.device)
            ops.append(dist.P2POp(dist.irecv, buf, right, group))
            recvs.append((t, buf))
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    for t, buf in recvs:
        t[1:-1, halo + ncols - width:halo + ncols] = buf


class _DevicePointer:
    """Zero-copy view of library-owned device memory for torch (CUDA array interface)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class BlockFluidSim:
    """This rank's column block of a width x height simulation spread over `dist.get_world_size()` GPUs."""

    def __init__(self, width, height, proj_iters=20, diff_iters=5, viscosity=1e-4, diffusion=0.0, omega=1.5,
                 noise="glibc", noise_seed=1, noise_pos=0, device=0, halo=128, levels=10, strips_per_cta=1, reach=None, group=None,
                 chain=True):
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.group = group
        self.width, self.height = width, height
        self.P, self.D = proj_iters, diff_iters
        self.visc, self.diff = viscosity, diffusion
        self.halo, self.levels = halo, levels
        self.j0, self.j1 = column_range(height, self.rank, self.world, 32 * strips_per_cta)
        self.ncols = self.j1 - self.j0
        self.plan = plan_projection(proj_iters, levels)
        self.dplan = plan_diffusion(diff_iters)
        if max([lv * st for lv, st in self.plan + self.dplan] + [0]) > halo:
            raise FsbError("BlockFluidSim: halo narrower than the sweeps fused into one launch")
        self._lib = load_library()
        p = Params()
        self._lib.fsb_default_params(C.byref(p))
        p.width, p.height, p.proj_iters, p.diff_iters = width, height, proj_iters, diff_iters
        p.viscosity, p.diffusion, p.omega = viscosity, diffusion, omega
        p.noise_mode = NOISE_GLIBC if noise == "glibc" else NOISE_ZERO
        p.noise_seed, p.noise_pos, p.device = noise_seed, noise_pos, device
        self._h = _SIM()
        self._check(self._lib.fsb_create_block(C.byref(p), self.j0, self.j1, halo, C.byref(self._h)))
        self._check(self._lib.fsb_set_tuning(self._h, strips_per_cta, levels, 0))
        pitch, jbase = C.c_int(), C.c_int()
        self._check(self._lib.fsb_block_info(self._h, C.byref(pitch), C.byref(jbase), None, None))
        self.pitch, self.jbase = pitch.value, jbase.value
        self.device = torch.device("cuda", device)
        st = C.c_void_p()
        self._check(self._lib.fsb_get_stream(self._h, C.byref(st)))
        self.stream = torch.cuda.ExternalStream(st.value, device=self.device)
        self.seq = 0
        # columns exchanged before every advection: a neighbour can only supply columns it owns
        # Columns exchanged before every advection.  A back-trace moves up to dt * width * max|v| columns per stage
        # (src/fluid_sim.cpp:264, :275), so the width follows the velocity field: `reach=None` (default) exchanges what
        # the all-reduced maximum of the previous step asks for, with a margin, up to `max_reach`; a number fixes it.
        # Either way every rank checks on the device that the window it received was wide enough (sync() raises).
        narrowest = min(b - a for a, b in (column_range(height, r, self.world, 32 * strips_per_cta) for r in range(self.world)))
        self.max_reach = min(halo, narrowest)
        self.fixed_reach = reach
        if reach is not None and not 3 <= reach <= self.max_reach:
            raise FsbError("BlockFluidSim: reach must be between 3 and min(halo, narrowest column block)")
        self.reach = self.max_reach if reach is None else reach
        self._overshoot = torch.zeros((), dtype=torch.float32, device=self.device)
        self._need = torch.zeros((), dtype=torch.float32, device=self.device)      # columns the last traces needed
        self._need_hosts = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(4)]
        self._pending = []      # (event, pinned host scalar, width used) of the steps whose all-reduced need is not read yet
        self._known_need = None
        self._step_no = 0
        self._side = torch.cuda.Stream(device=self.device)
        vp = C.c_void_p()
        self._check(self._lib.fsb_blk_vmax_word(self._h, C.byref(vp)))
        self._vmax = torch.as_tensor(_DevicePointer(vp.value, (1,)), device=self.device)
        self._used = 0          # widest exchange of the step being run
        self.trace = None      # set to {} to collect GPU time per phase (CUDA events on the simulation's stream)
        self._ev = []
        self._connect()
        self.chained = False
        # chain="auto": the chained launch where it was measured faster than pass-by-pass launches (blocks of at least
        # 4096 columns: 32768 x 4096 per rank at 8 ranks 71.7 against 78.9 ms per step; with 2048 columns per rank,
        # config 4 at 8 ranks, a rank has too few blocks for it: 20.1 against 19.1 ms)
        if chain is True or (chain == "auto" and self.ncols >= 4096):
            self._connect_chain()

    # -- plumbing ----------------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise FsbError(f"fluid_sim_b200 error {rc}: {self._lib.fsb_last_error().decode()}")

    def _connect(self):
        """Exchange the CUDA IPC handles of the link buffers and map the neighbours'."""
        mine = (C.c_ubyte * 64)()
        self._check(self._lib.fsb_link_handle(self._h, mine))
        t = torch.tensor(list(mine), dtype=torch.uint8, device=self.device)
        allh = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(allh, t, group=self.group)
        left, right = _neighbours(self.rank, self.world)
        lh = (C.c_ubyte * 64)(*allh[left].cpu().tolist()) if left is not None else None
        rh = (C.c_ubyte * 64)(*allh[right].cpu().tolist()) if right is not None else None
        self._check(self._lib.fsb_link_connect(self._h, lh, rh))
        # the rank on the right stores the result columns it owes this rank straight into these field buffers
        fh = (C.c_ubyte * 384)()
        self._check(self._lib.fsb_field_handles(self._h, fh))
        t = torch.tensor(list(fh), dtype=torch.uint8, device=self.device)
        allf = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(allf, t, group=self.group)
        if left is not None:
            lf = (C.c_ubyte * 384)(*allf[left].cpu().tolist())
            self._check(self._lib.fsb_link_connect_fields(self._h, lf, self.ncols))
        self.returned = [0, 0, 0]    # strips' worth of columns the right rank has stored here so far, per sweep slot
        dist.barrier(group=self.group)

    def _connect_chain(self):
        """The whole step as one chained launch per rank (csrc/chain.cu): possible when every rank has a chained
        configuration (sweep counts that are multiples of 10 / 5, non-zero viscosity, zero density diffusion).  The
        ranks then map each other's fields, second-projection scratch and chain buffers (CUDA IPC); a step is
        `fsb_blk_step` and nothing else - no halo exchange, no collective, no tensor operation."""
        rc = self._lib.fsb_blk_chain_prepare(self._h, C.c_float(0.01))
        ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) != 1:
            return
        ch = (C.c_ubyte * 192)()
        self._check(self._lib.fsb_blk_chain_handles(self._h, ch))
        fh = (C.c_ubyte * 384)()
        self._check(self._lib.fsb_field_handles(self._h, fh))
        t = torch.tensor(list(ch) + list(fh), dtype=torch.uint8, device=self.device)
        allh = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(allh, t, group=self.group)
        left, right = _neighbours(self.rank, self.world)

        def pack(r):
            if r is None:
                return None, None
            raw = allh[r].cpu().tolist()
            return (C.c_ubyte * 192)(*raw[:192]), (C.c_ubyte * 384)(*raw[192:])
        lc, lf = pack(left)
        rc_, rf = pack(right)
        self._check(self._lib.fsb_blk_chain_connect(self._h, lc, lf, rc_, rf))
        dist.barrier(group=self.group)
        self.chained = True

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.fsb_destroy(self._h)
            self._h = _SIM()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tensor(self, field):
        """The local [width, pitch] array of a field as a torch tensor (valid until the next advection swaps buffers)."""
        ptr = C.c_void_p()
        self._check(self._lib.fsb_device_field(self._h, field, C.byref(ptr)))
        return torch.as_tensor(_DevicePointer(ptr.value, (self.width, self.pitch)), device=self.device)

    def owned(self, field):
        return self.tensor(field)[:, self.halo:self.halo + self.ncols]

    def sync(self):
        self._check(self._lib.fsb_sync(self._h))
        if self.chained:
            w = C.c_uint64()
            self._check(self._lib.fsb_blk_chain_error(self._h, C.byref(w)))
            if w.value:
                raise FsbError("BlockFluidSim: a back-trace left the halo columns (or a wait timed out) in a chained step, "
                               f"error word {w.value:#x}; results are invalid (create the blocks with a wider `halo`)")
            return
        over = float(self._overshoot.item())
        if over > 0:
            raise FsbError(f"BlockFluidSim: an advection reached {int(over)} columns past the {self.reach} exchanged; "
                           "results since the last sync() are invalid (raise `reach`/`halo` or lower dt)")

    def _mark(self, name):
        if self.trace is None:
            return
        e = torch.cuda.Event(enable_timing=True)
        e.record(self.stream)
        self._ev.append((name, e))

    def trace_timeline(self):
        """[(phase, ms since the first mark)] since the last report (synchronises); leaves the marks in place."""
        self.sync()
        return [(n, self._ev[0][1].elapsed_time(e)) for n, e in self._ev]

    def trace_report(self):
        """ms per phase since the last report (synchronises)."""
        self.sync()
        out = {}
        for (n0, e0), (n1, e1) in zip(self._ev, self._ev[1:]):
            out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1)
        self._ev = []
        return out

    def _ids(self, vals):
        return (C.c_int * len(vals))(*vals)

    def _bnd(self, fields, bs, pz=None):
        self._check(self._lib.fsb_blk_set_boundary(self._h, len(fields), self._ids(fields), self._ids(bs),
                                                   self._ids(pz) if pz else None))

    def _halo(self, fields, width):
        exchange_halo([self.tensor(f) for f in fields], self.halo, self.ncols, width, self.rank, self.world, self.group)

    def _return(self, fields, width):
        return_columns([self.tensor(f) for f in fields], self.halo, self.ncols, width, self.rank, self.world, self.group)

    def _await_returns(self, nfields, sweeps):
        """Hold the stream until the rank on the right has stored this rank's last `sweeps` columns (and its first
        halo column) of every field of the launch; no host wait, no collective."""
        if self.rank == self.world - 1:
            return
        for slot in range(nfields):
            self.returned[slot] += sweeps // 32 + 1
            self._check(self._lib.fsb_blk_wait_returns(self._h, slot, self.returned[slot]))

    def _check_reach(self, dt, width):
        """Columns an RK4 back-trace can move: every stage moves at most dt*width*max|velocityY| cells (bilinear samples
        never exceed the largest tap).  The gradient kernel of the projection that has just run left the largest
        |velocityY| of this rank's columns in one device word (atomicMax fused into the kernel); this turns it into
        the window the traces need, records - on the device, without a host wait - by how much that exceeds the
        `width` columns just exchanged, and clears the word.  The maxima of all ranks are all-reduced at the end of
        the step (lagging one step: _publish_need), which is also how an overshoot on another rank's columns, which
        could reach into this rank's window, is noticed everywhere.  `sync()` raises if any step overshot."""
        m = self._vmax[0]
        need = torch.ceil(m * (dt * self.width)) + 3.0      # + bilinear tap + rounding + the extra advected halo column
        self._overshoot = torch.maximum(self._overshoot, need - float(width))
        self._need = torch.maximum(self._need, need)
        self._vmax.zero_()

    def _publish_need(self):
        """End of a step: all-reduce (max) the widest window any rank's traces needed and bring it to the host, on a
        side stream.  The result is read two steps later (_choose_reach), so the host never waits for the GPU."""
        ready = torch.cuda.Event()
        ready.record(self.stream)
        self._side.wait_event(ready)
        host = self._need_hosts[self._step_no % len(self._need_hosts)]
        with torch.cuda.stream(self._side):
            t = self._need.clone()
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            host.copy_(t, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._side)
        self._pending.append((ev, host, self._used))
        self._need = torch.zeros((), dtype=torch.float32, device=self.device)
        self._step_no += 1

    def _choose_reach(self):
        """Exchange width of this step, the same number on every rank: 2 x what the step before the previous one needed
        + 4 columns (velocities change little in two steps; the device-side check catches the rest), within
        [8, max_reach] and a multiple of 4; the full width until that is known.  Using the step before the previous
        one keeps the host a full step ahead of the GPU."""
        while len(self._pending) > 1:                        # everything but the previous step's record
            ev, host, used = self._pending.pop(0)
            ev.synchronize()                                 # recorded two steps ago: long finished
            needed = float(host.item())
            if needed > used > 0:
                raise FsbError(f"BlockFluidSim: a step's traces needed {int(needed)} columns, {used} were exchanged; the "
                               "results since then are invalid (raise `halo`, fix `reach`, or lower dt)")
            self._known_need = needed
        if self.fixed_reach is not None:
            return self.fixed_reach
        if self._known_need is None:
            return self.max_reach
        want = int(math.ceil(2.0 * self._known_need + 4.0))
        want = (max(want, 8) + 3) // 4 * 4
        return min(want, self.max_reach)

    # -- inputs / outputs -----------------------------------------------------------------------------
    def set_from_global(self, field, full):
        """Load this rank's columns (and halos, where the grid has them) from a full [width, height] host array."""
        loc = np.zeros((self.width, self.pitch), np.float32)
        a, b = max(self.jbase, 0), min(self.jbase + self.pitch, self.height)
        loc[:, a - self.jbase:b - self.jbase] = full[:, a:b]
        self._check(self._lib.fsb_write_field(self._h, field, loc.ctypes.data))

    def read_owned(self, field):
        with torch.cuda.stream(self.stream):
            out = self.owned(field).contiguous().cpu()
        return out.numpy()

    def read_owned_async(self, field, host):
        """Queue the copy of this rank's owned columns of a field into `host` (a pinned [width, ncols] float32 tensor)
        and return; the copy runs on a second stream from a staging buffer, so later steps overlap with it."""
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._staging = torch.empty((self.width, self.ncols), dtype=torch.float32, device=self.device)
            self._copy_done = None
        with torch.cuda.stream(self.stream):
            if self._copy_done is not None:
                self.stream.wait_event(self._copy_done)        # the staging buffer is free again
            self._staging.copy_(self.owned(field))
            ready = torch.cuda.Event()
            ready.record(self.stream)
        self._copy_stream.wait_event(ready)
        with torch.cuda.stream(self._copy_stream):
            host.copy_(self._staging, non_blocking=True)
            self._copy_done = torch.cuda.Event()
            self._copy_done.record(self._copy_stream)

    def read_wait(self):
        """Block until the copy queued by read_owned_async has arrived."""
        if getattr(self, "_copy_done", None) is not None:
            self._copy_done.synchronize()

    def gather(self, field, dst=0):
        """Full [width, height] array on rank `dst` (None elsewhere)."""
        with torch.cuda.stream(self.stream):
            mine = self.owned(field).contiguous()
            parts = [torch.empty_like(mine) for _ in range(self.world)] if self.rank == dst else None
            dist.gather(mine, parts, dst=dst, group=self.group)
            if self.rank != dst:
                return None
            return torch.cat(parts, dim=1).cpu().numpy()

    def addDensity(self, x, y, amount):
        self._check(self._lib.fsb_add_density(self._h, x, y, amount))   # ignored unless this rank owns column y

    def addVelocity(self, x, y, ax, ay):
        self._check(self._lib.fsb_add_velocity(self._h, x, y, ax, ay))

    def _coefficient(self, diff, dt):
        """a = ((dt*diff)*width)*height in float32, src/fluid_sim.cpp:45"""
        return np.float32(np.float32(np.float32(dt) * np.float32(diff)) * np.float32(self.width)) * np.float32(self.height)

    def _diffuse(self, fields, bs, scratch, diff, dt, fold):
        """diffuse() of src/fluid_sim.cpp:43-62 on `fields` (sharing the coefficient): any sweep count (several launches
        keep the pre-diffusion fields in `scratch` as right-hand side), a == 0 (x + 0.0f and the boundary update).
        Returns True if sweep launches ran (the one-column halos are then fresh)."""
        if self.D <= 0:
            return False
        n = len(fields)
        if self._coefficient(diff, dt) == 0:
            self._check(self._lib.fsb_blk_add_zero(self._h, n, self._ids(fields)))
            self._bnd(fields, bs)
            return False
        rhs = None
        if len(self.dplan) > 1:
            for f, sc in zip(fields, scratch):
                self.tensor(sc).copy_(self.tensor(f))
            rhs = self._ids(scratch)
        for lv, st in self.dplan:
            self.seq += 1
            self._check(self._lib.fsb_blk_sweep(self._h, 1, n, self._ids(fields), self._ids(bs), lv, st, 0, diff, dt,
                                                self.seq, rhs))
            self._mark("sweep")
            self._await_returns(n, lv * st)
            self._mark("return")
            self._bnd(fields, bs)
        return True

    # -- the step (same order as velocityStep / densityStep, src/fluid_sim.cpp:370-422) ---------------------
    def _project(self, fresh_halo):
        if not fresh_halo:
            self._halo([VELOCITY_X, VELOCITY_Y], 1)
        self._mark("halo1")
        self._check(self._lib.fsb_blk_divergence(self._h))
        self._mark("pointwise")
        for k, (lv, st) in enumerate(self.plan):
            self.seq += 1
            self._check(self._lib.fsb_blk_sweep(self._h, 0, 1, None, None, lv, st, 1 if k == 0 else 0, 0.0, 0.0, self.seq, None))
            self._mark("sweep")
            self._await_returns(1, lv * st)
            self._mark("return")
            self._bnd([PREV_VEL_X], [0])
        if not self.plan:
            self.tensor(PREV_VEL_X).zero_()
            self._halo([PREV_VEL_X], 1)
        self._mark("halo1")
        self._check(self._lib.fsb_blk_gradient(self._h))
        self._bnd([VELOCITY_X, VELOCITY_Y], [1, 2])
        self._mark("pointwise")

    def step(self, dt=0.01):
        if self.chained:
            self._check(self._lib.fsb_blk_step(self._h, C.c_float(dt)))
            return
        with torch.cuda.stream(self.stream):
            self._mark("start")
            self._check(self._lib.fsb_op_add_noise(self._h))
            self._mark("pointwise")
            self.reach = self._choose_reach()
            self._used = self.reach
            swept = self._diffuse([VELOCITY_X, VELOCITY_Y], [1, 2], [PREV_VEL_X, PREV_VEL_Y], self.visc, dt, fold=False)
            self._project(fresh_halo=swept)
            self._halo([VELOCITY_X, VELOCITY_Y], self.reach)
            self._check_reach(dt, self.reach)
            self._mark("haloR")
            # the first halo column on each side is advected here as well: the divergence of the second projection
            # needs it, and the R-column halo just received covers its traces too (one exchange less per step)
            self._check(self._lib.fsb_blk_advect_velocity(self._h, dt, 1))
            self._bnd([VELOCITY_X, VELOCITY_Y], [1, 2])
            self._mark("pointwise")
            self._project(fresh_halo=True)
            # density: diffusion 0 is the identity d + 0.0f plus the boundary update, folded into the gather; with no
            # diffusion sweeps at all (D == 0) the reference's loop body, boundary update included, never runs
            fold = 0
            if self._coefficient(self.diff, dt) == 0:
                fold = 1 if self.D > 0 else 0
                if fold:
                    self._bnd([DENSITY], [0], [1])
            else:
                self._diffuse([DENSITY], [0], [PREV_DENSITY], self.diff, dt, fold=False)
            self._halo([DENSITY, VELOCITY_X, VELOCITY_Y], self.reach)
            self._check_reach(dt, self.reach)
            self._mark("haloR")
            self._check(self._lib.fsb_blk_advect_density(self._h, dt, fold))
            self._bnd([DENSITY], [0])
            self._mark("pointwise")
            self._publish_need()
