#!/usr/bin/env python3
"""Benchmark of the B200-native FluidSim::step (contract: see the task description / DESIGN.md).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]

A "step" is one FluidSim::step (velocityStep + densityStep, src/fluid_sim.cpp:23-28) over the whole
grid.  At N=1 the workload is BASELINE.json configs[2]: 8192 x 8192, 40 projection sweeps, 5 diffusion
sweeps, dt 0.01, vortex-lattice input, glibc rand() noise.  The metric is
cell-updates/s = W*H*steps / time.  One JSON line is printed by rank 0.

Keys beyond the base contract:
  roofline      dominant kernel (one fused Gauss-Seidel pass): effective GB/s = algorithmic bytes of
                the reference's pass structure (12 B per cell per sweep) / measured
                launch time, against MEASURED_PEAKS.json; `compulsory_gbs` is the same launch counted
                at its own compulsory traffic (inputs read once, result written once).
  step_roofline whole step at B(P) = 356 + 24 P bytes per cell.
  cpu_baseline  the real reference (oracle/_ref) timed on this host on a bounded sample.
  e2e           the reference's frame loop through the C ABI with host buffers: mouse-patch injection
                (src/main.cpp:216-252), step, full density read-back (src/main.cpp:388-396).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cell_updates_per_sec"
UNIT = "cell-updates/s"


def algorithmic_bytes_per_cell(P, D=5):
    """B(P) of SURVEY.md 8(d): noise 16 + copies 48 + diffuse 3*12*D + 2*(36 + 12 P) + advect 40."""
    return 16 + 48 + 3 * 12 * D + 2 * (36 + 12 * P) + 40


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the real reference built by oracle/build_ref.py
# --------------------------------------------------------------------------------------------
REF_RATE_GUESS = 1.0e6   # cell-updates/s of one host core at P = 40, to size the sample


def reference_samples(H, P):
    """Row counts of the prebuilt real-reference libraries (oracle/_ref) with H columns and P sweeps, largest first."""
    import re
    from oracle import refsim
    rows = []
    if os.path.isdir(refsim.REF_DIR):
        for name in os.listdir(refsim.REF_DIR):
            m = re.fullmatch(r"libref_(\d+)x(\d+)_p(\d+)_d5\.so", name)
            if m and int(m.group(2)) == H and int(m.group(3)) == P:
                rows.append(int(m.group(1)))
    return sorted(rows, reverse=True)


def pick_reference_sample(H, P, nsteps, budget_s):
    rows = reference_samples(H, P)
    rate = REF_RATE_GUESS * (356 + 24 * 40) / (356 + 24 * P)     # the step cost grows with the sweep count
    for w in rows:
        if w * H * nsteps / rate <= budget_s:
            return w, H
    return (rows[-1], H) if rows else None


def run_reference_sample(w, h, P, nsteps, warmup):
    """Times the real reference's step() on a w x h grid (same columns as the workload)."""
    from oracle import refsim
    from tests.util import vortex_lattice
    sim = refsim.RefSim(w, h, P, 5, seed=1)
    d, u, v = vortex_lattice(w, h)
    sim.write(refsim.DENSITY, d); sim.write(refsim.VEL_X, u); sim.write(refsim.VEL_Y, v)
    for _ in range(warmup):
        sim.step(0.01)
    t0 = time.perf_counter()
    for _ in range(nsteps):
        sim.step(0.01)
    dt = time.perf_counter() - t0
    sim.close()
    return w * h * nsteps / dt, dt / nsteps


def run_oracle_sample(w, h, P, nsteps, warmup):
    from oracle.orasim import OracleSim
    from tests.util import vortex_lattice
    sim = OracleSim(w, h, P, 5)
    d, u, v = vortex_lattice(w, h)
    sim.write(0, d); sim.write(1, u); sim.write(2, v)
    for _ in range(warmup):
        sim.step(0.01)
    t0 = time.perf_counter()
    for _ in range(nsteps):
        sim.step(0.01)
    dt = time.perf_counter() - t0
    return w * h * nsteps / dt, dt / nsteps


def cpu_reference(H, P, nsteps, warmup, budget_s):
    """(value, ms_per_step, kind, sample description, [rows, cols] of the sample)"""
    pick = pick_reference_sample(H, P, nsteps + warmup, budget_s)
    if pick is not None:
        w, h = pick
        v, spt = run_reference_sample(w, h, P, nsteps, warmup)
        return v, spt * 1e3, "reference", f"{w}x{h} rows x cols of the {H}-column workload, P={P}, {nsteps} step(s), real reference (oracle/_ref)", [w, h]
    w = max(16, int(budget_s * REF_RATE_GUESS / max(1, (nsteps + warmup) * H)))
    v, spt = run_oracle_sample(w, H, P, nsteps, warmup)
    return v, spt * 1e3, "port", f"{w}x{H} slab, P={P}, {nsteps} step(s), C restatement (oracle/fluid_oracle.c)", [w, H]


def workload_for(n_gpus, args):
    if args.width and args.height:
        return args.width, args.height, args.proj or 40, "custom"
    if args.workload == "config4":
        # config 4 of BASELINE.json: one fixed 16384^2 grid, P = 20, cut into column blocks (strong scaling)
        return 16384, 16384, args.proj or 20, "config4: 16384x16384 grid, 20 projection sweeps, strong-scaled"
    if n_gpus == 1 and args.workload != "config5":
        return 8192, 8192, args.proj or 40, "config3: 8192x8192 grid, 40 projection sweeps, 1xB200"
    # config 5 of BASELINE.json, weak-scaled: 32768 rows x 4096 columns per GPU (32768^2 at 8 GPUs), P = 100
    return 32768, 4096 * n_gpus, args.proj or 100, "config5: 32768-row grid, 4096 columns per GPU (32768^2 at 8 GPUs), 100 projection sweeps"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--proj", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-scaling-base", action="store_true", help="skip the extra single-GPU run of the weak-scaling workload")
    ap.add_argument("--workload", default="", help="config5: run the weak-scaling per-GPU workload at N=1 too; config4: 16384^2, P=20, strong scaling; "
                    "view: time the headless density / velocity-glyph views")
    ap.add_argument("--strips-per-cta", type=int, default=0)
    ap.add_argument("--sweeps-per-warp", type=int, default=0)
    ap.add_argument("--max-stages", type=int, default=0)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n_gpus = max(args.gpus, world)
    K, Wm = args.steps, max(args.warmup, 0)
    W, H, P, workload = workload_for(n_gpus, args)

    # ------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        if rank != 0:
            return 0
        value, ms, kind, sample, sample_grid = cpu_reference(H, P, K, Wm, budget_s=150.0)
        line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K,
                "warmup": Wm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload, "grid": [W, H], "proj_iters": P, "diff_iters": 5, "dt": 0.01,
                           "input": "vortex lattice (SURVEY.md 8(d) I3) + glibc rand() noise",
                           "reference_sample": sample_grid},
                "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample},
                "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    # ------------------------------------------------------------------ native arm
    import torch
    import torch.distributed as dist
    import fluid_sim_b200 as fsb
    from tests.util import vortex_lattice

    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device: the native arm has no CPU fallback"}))
        return 1
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.workload == "view":
        # the headless views (what the reference's viewer draws, src/main.cpp:383-514) at the reference's
        # 800 x 600 window.  Not the north-star metric: a separate line, pixels per second per view.
        import ctypes as C
        pw, ph = 800, 600
        out = {"metric": "view_pixels_per_sec", "unit": "pixels/s", "n_gpus": 1, "steps": K, "warmup": max(Wm, 3),
               "higher_is_better": True, "dtype": "f32 -> rgba8", "data": "synthetic", "vs_baseline": None,
               "config": {"workload": "headless density and velocity-glyph views, 800x600 window", "grids": []}, "views": []}
        peak, _ = measured_peak_gbs()
        for n in (200, 2048):
            sim = fsb.FluidSim(n, n, 20, 5, device=local_rank)
            d, u, v = vortex_lattice(n, n)
            sim.write(fsb.DENSITY, d); sim.write(fsb.VELOCITY_X, u); sim.write(fsb.VELOCITY_Y, v)
            for _ in range(2):
                sim.step(0.01)
            sim.sync()
            st = torch.cuda.ExternalStream(sim.stream(), device=torch.device("cuda", local_rank))
            img = np.empty((ph, pw, 4), np.uint8)
            dptr = C.c_void_p()
            res = {}
            for name, call in (("density", lambda: sim._check(sim._lib.fsb_view_density_device(sim._h, pw, ph, C.byref(dptr)))),
                               ("density+glyphs (with the 1.9 MB read-back)", lambda: sim.view_velocity(pw, ph, img))):
                for _ in range(max(Wm, 3)):
                    call()
                sim.sync()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0 = time.perf_counter()
                e0.record(st)
                for _ in range(K):
                    call()
                e1.record(st)
                sim.sync()
                ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / K
                res[name] = {"ms_per_frame": ms, "pixels_per_sec": pw * ph / (ms * 1e-3),
                             "gbs_at_8_bytes_per_pixel": pw * ph * 8 / (ms * 1e-3) / 1e9}
            out["config"]["grids"].append([n, n])
            out["views"].append({"grid": [n, n], **res})
            sim.close()
        out["value"] = out["views"][-1]["density"]["pixels_per_sec"]
        out["ms_per_step"] = out["views"][-1]["density"]["ms_per_frame"]
        out["roofline"] = {"bound": "launch latency (480 000 pixels x 8 bytes = 3.8 MB per frame)", "achieved":
                           out["views"][-1]["density"]["gbs_at_8_bytes_per_pixel"], "peak": peak, "unit": "GB/s",
                           "frac": out["views"][-1]["density"]["gbs_at_8_bytes_per_pixel"] / peak, "traffic": None}
        print(json.dumps(out))
        return 0

    multi = world > 1
    parity = None
    if multi:
        from fluid_sim_b200.multi import BlockFluidSim
        from tests.util import vortex_lattice_block
        # ---- gate before anything is timed: the sharded step must give the single-GPU bits.  Small grid, same launch
        # plan as the timed workload (same sweeps per warp, stages and strips per CTA, 2 steps, glibc noise); rank 0
        # steps the same problem on one GPU and compares the gathered fields as uint32.
        pw, ph = 512, 256 * world
        spc = 2 if P <= 20 else 1
        use_chain = (H // world) >= 4096          # the timed workload's launch structure (BlockFluidSim chain="auto")
        chk = BlockFluidSim(pw, ph, P, 5, device=local_rank, strips_per_cta=spc, chain=use_chain)
        cd, cu, cv = vortex_lattice(pw, ph)
        cu = (cu + np.random.default_rng(7).standard_normal((pw, ph)).astype(np.float32) * 0.05).astype(np.float32)
        for fid, arr in ((fsb.DENSITY, cd), (fsb.VELOCITY_X, cu), (fsb.VELOCITY_Y, cv)):
            chk.set_from_global(fid, arr)
        for _ in range(2):
            chk.step(0.01)
        chk.sync()
        got = [chk.gather(f) for f in (fsb.DENSITY, fsb.VELOCITY_X, fsb.VELOCITY_Y)]
        plan = chk.plan
        chk.close()
        ok = torch.ones((), dtype=torch.int32, device="cuda")
        if rank == 0:
            one = fsb.FluidSim(pw, ph, P, 5, device=local_rank)
            one.write(fsb.DENSITY, cd); one.write(fsb.VELOCITY_X, cu); one.write(fsb.VELOCITY_Y, cv)
            for _ in range(2):
                one.step(0.01)
            same = [bool(np.array_equal(g.view(np.uint32), w_.view(np.uint32))) for g, w_ in zip(got, one.fields())]
            one.close()
            parity = {"ok": all(same), "world": world, "grid": [pw, ph], "steps": 2, "proj_iters": P, "diff_iters": 5,
                      "plan": [list(x) for x in plan], "strips_per_cta": spc, "fields_identical": same,
                      "launch_structure": "one chained launch per rank and step" if use_chain else "one launch per pass, NCCL halo exchanges",
                      "compared": "uint32 bits of density, velX, velY gathered from all ranks vs one GPU"}
            ok.fill_(1 if all(same) else 0)
        dist.broadcast(ok, 0)
        if int(ok.item()) != 1:
            if rank == 0:
                print(json.dumps({"error": "multi-GPU fields differ from the single-GPU fields", "parity_vs_single_gpu": parity}))
            dist.destroy_process_group()
            return 1
        del got, cd, cu, cv
        # (10, 2) passes leave room for two strips per CTA; the (10, 5) passes of config 5 need one strip per CTA
        sim = BlockFluidSim(W, H, P, 5, device=local_rank, strips_per_cta=2 if P <= 20 else 1, chain="auto")
        a0, a1 = sim.jbase, sim.jbase + sim.pitch
        d, u, v = vortex_lattice_block(W, H, a0, a1)
        for fid, arr in ((fsb.DENSITY, d), (fsb.VELOCITY_X, u), (fsb.VELOCITY_Y, v)):
            sim._check(sim._lib.fsb_write_field(sim._h, fid, np.ascontiguousarray(arr).ctypes.data))
        del d, u, v
        stream = sim.stream
        lib_sim = sim

        class _Shim:   # the few calls the timing code below makes, on the block simulation
            def step(self, dt): lib_sim.step(dt)
            def sync(self): lib_sim.sync()
            def profile_enable(self, on): lib_sim._check(lib_sim._lib.fsb_profile_enable(lib_sim._h, 1 if on else 0))
            def profile_read(self, reset=True):
                import ctypes as C
                names = fsb.FluidSim.PROF_NAMES
                ms = (C.c_double * len(names))(); cnt = (C.c_uint64 * len(names))()
                lib_sim._check(lib_sim._lib.fsb_profile_read(lib_sim._h, ms, cnt, 1 if reset else 0))
                return {n: (ms[i], cnt[i]) for i, n in enumerate(names)}
            def launch_count(self):
                import ctypes as C
                v_ = C.c_uint64(); lib_sim._check(lib_sim._lib.fsb_launch_count(lib_sim._h, C.byref(v_))); return v_.value
        sim = _Shim()
    else:
        sim = fsb.FluidSim(W, H, P, 5, device=local_rank, strips_per_cta=args.strips_per_cta,
                           sweeps_per_warp=args.sweeps_per_warp, max_stages=args.max_stages)
        stream = torch.cuda.ExternalStream(sim.stream(), device=torch.device("cuda", local_rank))

        # synthetic input, built on the host in page-locked memory and injected through the C ABI
        d, u, v = vortex_lattice(W, H)
        buf = sim.pinned_array()
        for fid, arr in ((fsb.DENSITY, d), (fsb.VELOCITY_X, u), (fsb.VELOCITY_Y, v)):
            buf[...] = arr
            sim.add_field(fid, buf)
        del d, u, v

    for _ in range(max(Wm, 3) if K > 0 else 0):
        sim.step(0.01)
    sim.sync()

    # -------- value: K steps, state resident in HBM, CUDA events on the launching stream at both ends of the region.
    # Nothing else is recorded inside it: on one GPU the launches of consecutive steps overlap (the first blocks of step
    # n+1 run while the last blocks of step n finish), and a timing event between two launches would serialise them.
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = sim.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(K):
        sim.step(0.01)
    ev1.record(stream)
    sim.sync()
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    launches = sim.launch_count() - l0
    # -------- per kernel class: the same K steps once more, untimed for `value`, with timing events around every launch
    # (kernel_ms_per_step).  Those events serialise the launches: a class's time is its launches alone on the GPU.
    sim.profile_enable(True)
    sim.profile_read(reset=True)
    for _ in range(K):
        sim.step(0.01)
    sim.sync()
    prof = sim.profile_read(reset=True)
    sim.profile_enable(False)
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_per_step = ms_total / K
    cells_total = W * H            # H is the global column count at N > 1
    value = cells_total * K / (ms_total * 1e-3)

    # -------- e2e: the reference's frame loop with host buffers (inject patch, step, read density back)
    patch = [(i, j) for i in range(-3, 4) for j in range(-3, 4)]
    xs = np.array([W // 2 + i for i, _ in patch], np.int32)
    ys = np.array([H // 3 + j for _, j in patch], np.int32)
    ones = np.full(len(patch), 1.0, np.float32)
    dx = np.full(len(patch), 0.05, np.float32)
    dy = np.full(len(patch), -0.02, np.float32)
    h2d = xs.nbytes * 2 + ys.nbytes * 2 + ones.nbytes + dx.nbytes + dy.nbytes
    if multi:
        host_ds = [torch.empty((W, lib_sim.ncols), dtype=torch.float32).pin_memory() for _ in range(2)]
        d2h = W * lib_sim.ncols * 4
        nframe = [0]

        def frame():   # every rank reads its own columns back every step; the copy overlaps with the next step
            for x_, y_, o_, a_, b_ in zip(xs, ys, ones, dx, dy):
                lib_sim.addDensity(int(x_), int(y_), float(o_)); lib_sim.addVelocity(int(x_), int(y_), float(a_), float(b_))
            lib_sim.step(0.01)
            lib_sim.read_wait()
            lib_sim.read_owned_async(fsb.DENSITY, host_ds[nframe[0] % 2])
            nframe[0] += 1

        def checksum_of():
            lib_sim.read_wait()
            return float(host_ds[(nframe[0] - 1) % 2].double().sum())
    else:
        d2h = W * H * 4

        # every step's density goes to pinned host memory, as the reference's loop reads every cell after every step;
        # the copy of step k (fsb_read_field_async, second stream) overlaps with the velocity half of step k+1
        bufs = [buf, sim.pinned_array()]
        nframe = [0]

        def frame():
            sim.addDensityN(xs, ys, ones); sim.addVelocityN(xs, ys, dx, dy); sim.step(0.01)
            sim.read_wait()                                   # the previous step's image is complete on the host
            sim.read_async(fsb.DENSITY, bufs[nframe[0] % 2])
            nframe[0] += 1

        def checksum_of():
            sim.read_wait()
            return float(bufs[(nframe[0] - 1) % 2].astype(np.float64).sum())
    for _ in range(2):
        frame()
    (lib_sim if multi else sim).read_wait()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    t_host0 = time.perf_counter()
    for _ in range(K):
        frame()
    (lib_sim if multi else sim).read_wait()                   # the last image has arrived: end of the timed region
    e1.record(stream)
    sim.sync()
    e2e_host_ms = (time.perf_counter() - t_host0) * 1e3
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    e2e_ms = max(e2e_ms, e2e_host_ms)                         # the copies run on a second stream: take the host clock too
    if world > 1:
        t = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
        d2h *= world
    e2e_value = cells_total * K / (e2e_ms * 1e-3)
    checksum = checksum_of()

    # -------- the same per-GPU workload on ONE GPU (rank 0, the other ranks idle): the N = 1 point of the weak-scaling
    # series.  `bench.py --gpus 1` measures the 8192^2, P=40 headline instead, whose steps cost 2.5x less per cell.
    base = None
    c5 = (32768, 4096, 100)
    if not args.no_scaling_base and args.workload != "config4" and (multi or (args.workload != "config5" and not (args.width and args.height))):
        barrier()
        if rank == 0:
            bw, bh, bp = (W, H // world, P) if multi else c5
            if not multi:
                sim.close()                                       # free the headline workload's fields first
            one = fsb.FluidSim(bw, bh, bp, 5, device=local_rank)
            d, u, v = vortex_lattice(bw, bh)
            buf1 = one.pinned_array()
            for fid, arr in ((fsb.DENSITY, d), (fsb.VELOCITY_X, u), (fsb.VELOCITY_Y, v)):
                buf1[...] = arr
                one.add_field(fid, buf1)
            del d, u, v
            st1 = torch.cuda.ExternalStream(one.stream(), device=torch.device("cuda", local_rank))
            for _ in range(3):
                one.step(0.01)
            one.sync()
            b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b0.record(st1)
            for _ in range(K):
                one.step(0.01)
            b1.record(st1)
            one.sync()
            base_ms = b0.elapsed_time(b1) / K
            base = {"n_gpus": 1, "grid": [bw, bh], "proj_iters": bp, "ms_per_step": base_ms,
                    "value": bw * bh / (base_ms * 1e-3), "unit": UNIT,
                    "command": "python bench.py --gpus 1 --workload config5",
                    "note": "N = 1 point of the weak-scaling series that `--gpus N` runs (config 5: 32768 rows x 4096 "
                            "columns per GPU, P = 100); the headline `value` of a 1-GPU run is config 3 (8192^2, P = 40), "
                            "whose steps cost 2.5x less per cell"}
            one.close()
        barrier()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # -------- roofline of the dominant kernel and of the whole step
    peak, peak_src = measured_peak_gbs()
    cells = W * (H // world)                                  # cells of one rank's block
    tr = {}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            tr = json.load(open(traffic_file))
        except Exception:
            tr = {}
    chain_ms, chain_n = prof.get("chain", (0.0, 0))
    gs_ms, gs_n = prof["gs"]
    levels = float(P)
    if chain_n > 0 and gs_n == 0:
        # One chained launch per step: diffuse(u, v), project, advect(u, v), project and the density step as phases of
        # one cooperative kernel (fluid_sim_b200/csrc/chain.cu).
        launch_alone_ms = chain_ms / chain_n
        per_step = chain_n / float(K)
        whole = abs(per_step - 1.0) < 1e-6
        # the step is this one launch (noise included) and consecutive launches overlap: the launch's share of the timed
        # region is the region / launches (CUDA events at its ends); `launch_alone_ms` is the same launch with the GPU to
        # itself (profiled pass)
        everything = whole and int(launches) == K and prof.get("noise", (0.0, 0))[1] == 0
        launch_ms = ms_per_step if everything else launch_alone_ms
        # SURVEY.md 8(d) fusion rule: inputs not produced inside the kernel + outputs consumed outside it
        fused_bpc = 24.0 if whole else 20.0                   # u, v, d in and out / u, v in and out + the pressure
        # what the phases really move: SOR 2 x 8, per projection div 12 + sweeps 8 + gradient 20, advections 16 + 16
        moved_bpc = (16.0 + 2 * 40.0 + 32.0) if whole else 56.0
        eff_bpc = (algorithmic_bytes_per_cell(P) - 16.0) if whole else (36.0 + 12.0 * P + 120.0)
        if everything:
            moved_bpc += 16.0                                     # the noise phase: u, v read and written
            eff_bpc += 16.0
        fused_gbs = fused_bpc * cells / (launch_ms * 1e-3) / 1e9
        roofline = {"kernel": ("chain_kernel<5, 10>: noise, diffuse(u), diffuse(v), project, advect(u, v), project and the density step: "
                               "ALL of FluidSim::step as phases of ONE launch; launches of consecutive steps overlap" if everything else
                               "chain_kernel<5, 10>: diffuse(u), diffuse(v), project, advect(u, v), project and the density step of "
                               "one FluidSim::step as phases of ONE cooperative launch (everything after the noise)") if whole else
                              "chain_kernel: diffuse + project / project as phases of one launch",
                    "bound": ("the serial work of one thread block on its block of columns (every SM that holds a block is busy "
                              "from launch to launch: the sweep phases run at the step latency of the exact-order pipeline); " if everything else
                              "dependency chain of the exact-order sweeps (stagger between column blocks + step latency of the "
                              "sweep pipeline); ") + "not HBM: see `frac`, and `ceilings` for issue slots / shared-memory wavefronts",
                    "achieved": fused_gbs, "peak": peak, "unit": "GB/s", "frac": fused_gbs / peak,
                    "moved_gbs": moved_bpc * cells / (launch_ms * 1e-3) / 1e9,
                    "effective_gbs": eff_bpc * cells / (launch_ms * 1e-3) / 1e9,
                    "effective_frac": eff_bpc * cells / (launch_ms * 1e-3) / 1e9 / peak,
                    "traffic": None, "peak_source": peak_src, "launch_ms": launch_ms, "launches_timed": int(launches) if everything else chain_n,
                    "launch_alone_ms": launch_alone_ms,
                    "launch_timing": ("timed region (CUDA events at both ends, K back-to-back launches that overlap) / K; "
                                      "launch_alone_ms: one launch with the GPU to itself (events around every launch)") if everything
                                     else "CUDA events around every launch",
                    "algorithmic_bytes_per_launch": fused_bpc * cells, "moved_bytes_per_launch": moved_bpc * cells,
                    "sweep_cell_updates_per_sec": (2 * P + 10) * cells / (launch_ms * 1e-3),
                    "note": "achieved / frac = SURVEY.md 8(d)'s fusion rule (fields read once, results written once: "
                            f"{fused_bpc:g} B per cell per launch) / launch time.  moved_gbs = the bytes the phases really move "
                            f"({moved_bpc:g} B per cell: every phase streams its fields through HBM once; `traffic` = ncu DRAM bytes "
                            "of one launch).  effective_gbs / effective_frac = the same launch counted at the reference's pass "
                            "structure (B(P) minus the noise pass), which fusion makes exceed the real traffic"}
        key = f"{W}x{H}_p{P}_chain"
        if isinstance(tr.get(key), dict):
            t = tr[key]
            roofline["traffic"] = t.get("dram_bytes")
            clk = (clocks.get("sm_mhz") or 1965.0) * 1e6
            sm_cycles = t.get("grid", 130) * clk * launch_ms * 1e-3          # on the SMs the launch occupies
            roofline["ceilings"] = {"issue_slot_frac": t["inst_executed"] / (4.0 * sm_cycles),
                                    "smem_wavefront_frac": t["smem_wavefronts"] / sm_cycles,
                                    "source": "instruction and shared-memory wavefront counts of one launch (ncu, profiles/) / "
                                              "(SMs used x SM clock x event-timed launch duration): neither is the bound"}
    else:
        gs_launch_ms = gs_ms / max(gs_n, 1)
        launches_per_proj = max(1.0, gs_n / (2.0 * K))          # two projections per step
        levels = P / launches_per_proj                            # sweeps fused into one launch
        algo_bytes = 12.0 * levels * cells                        # reference pass structure: 12 B per cell per sweep
        compulsory = 8.0 * cells if launches_per_proj <= 1.0 else (8.0 + 12.0 * (launches_per_proj - 1)) / launches_per_proj * cells
        achieved = algo_bytes / (gs_launch_ms * 1e-3) / 1e9 if gs_launch_ms > 0 else 0.0
        compulsory_gbs = compulsory / (gs_launch_ms * 1e-3) / 1e9 if gs_launch_ms > 0 else 0.0
        fused_gbs = 12.0 * cells / (gs_launch_ms * 1e-3) / 1e9 if gs_launch_ms > 0 else 0.0   # 12 B per cell per fused launch
        roofline = {"kernel": f"wf_kernel<S, GaussSeidel>: one launch fuses {levels:g} sweeps of a projection",
                    "bound": "dependency chain of the exact-order sweeps / instruction issue and shared-memory bandwidth "
                             "(not HBM: see `frac`)",
                    "achieved": fused_gbs, "peak": peak, "unit": "GB/s", "frac": fused_gbs / peak,
                    "compulsory_gbs": compulsory_gbs, "effective_gbs": achieved, "effective_frac": achieved / peak,
                    "traffic": None, "peak_source": peak_src, "launch_ms": gs_launch_ms, "launches_timed": gs_n,
                    "algorithmic_bytes_per_launch": algo_bytes,
                    "sweep_cell_updates_per_sec": levels * cells / (gs_launch_ms * 1e-3) if gs_launch_ms > 0 else 0.0,
                    "compulsory_bytes_per_launch": compulsory,
                    "note": "achieved / frac = SURVEY.md 8(d)'s fusion rule, 12 B per cell per fused launch, / launch time (the launch "
                            "really moves the compulsory 8 B per cell: `compulsory_gbs`): the kernel is nowhere near HBM, it is bound "
                            "by the dependency chain of the exact-order sweeps.  effective_gbs / effective_frac = the same launch counted "
                            "at the reference's pass structure (12 B per cell per SWEEP), which fusion makes exceed the real "
                            "traffic; `traffic` = ncu DRAM bytes per launch (profiles/)"}
        key = f"{W}x{H}_p{P}_gs"
        if key in tr and not isinstance(tr[key], dict):
            roofline["traffic"] = tr[key]
    Bp = algorithmic_bytes_per_cell(P)
    step_gbs = Bp * cells / (ms_per_step * 1e-3) / 1e9
    step_roofline = {"bytes_per_cell": Bp, "achieved": step_gbs, "peak": peak, "unit": "GB/s", "frac": step_gbs / peak,
                     "note": "the whole step counted at the reference's pass structure, B(P) = 356 + 24 P bytes per cell (the figure "
                             "north_star's 0.70 target is stated in): an effective rate, not DRAM utilisation -- it exceeds 1 once "
                             "fused passes no longer move those bytes (real traffic: roofline.traffic)"}
    shares = {k: {"ms_per_step": v[0] / K, "launches_per_step": v[1] / K} for k, v in prof.items()}
    # per kernel class: effective GB/s at SURVEY.md 8(d)'s algorithmic bytes per cell and launch
    per_launch_bytes = {"noise": 16.0, "divergence": 16.0, "gradient": 20.0, "advect_velocity": 16.0, "advect_scalar": 16.0,
                        "sor": 12.0 * 5 * 2, "gs": 12.0 * levels,
                        "chain": algorithmic_bytes_per_cell(P) - (16.0 if prof.get("noise", (0.0, 0))[1] > 0 else 0.0)}
    for k, bpc in per_launch_bytes.items():
        if k in shares and prof[k][1] > 0 and prof[k][0] > 0:
            gbs = bpc * cells / (prof[k][0] / prof[k][1] * 1e-3) / 1e9
            shares[k]["effective_gbs"] = gbs
            shares[k]["frac_of_peak"] = gbs / peak

    # -------- CPU baseline: the real reference on this host, bounded sample
    if args.no_cpu_baseline or world > 1:
        cpu = None
    else:
        cv, cms, kind, sample, sample_grid = cpu_reference(H, P, 1, 0, budget_s=25.0)
        cpu = {"value": cv, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample, "sample_grid": sample_grid,
               "ms_per_step_of_sample": cms,
               "host_cores_visible": os.cpu_count()}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(Wm, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if args.workload == "config4" else "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload, "grid": [W, H], "grid_per_gpu": [W, H // world], "proj_iters": P, "diff_iters": 5, "dt": 0.01,
                       "input": "vortex lattice (SURVEY.md 8(d) I3) + glibc rand() noise",
                       "l2": "fields are larger than L2 (no flush needed)" if cells * 4 > 126e6 else "fields fit in L2",
                       "decomposition": "single GPU" if world == 1 else
                       ("column blocks, one per GPU; each rank runs its blocks of the step as ONE chained launch: the exact-order "
                        "sweep chain crosses GPUs through CUDA-IPC peer rings inside the kernel, blocks near a rank boundary store "
                        "their results into the neighbour's halo and publish their done-words there (no halo exchange, no collective)"
                        if (H // world) >= 4096 else
                        "column blocks, one per GPU: the exact-order sweep chain crosses GPUs through CUDA-IPC peer rings "
                        "inside the kernel, halo columns move with NCCL send/recv between passes")},
            "roofline": roofline, "step_roofline": step_roofline, "kernel_ms_per_step": shares,
            "kernel_ms_per_step_note": "from a second, untimed pass of K steps with CUDA timing events around every launch; the events "
                                       "serialise launches that otherwise overlap, so an entry is that class alone on the GPU and the "
                                       "entries may add up to more than ms_per_step",
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_ms / K, "density_checksum": checksum},
            "gpu_launches": int(launches), "clocks": clocks}
    if base is not None:
        line["weak_scaling_base"] = base
    if parity is not None:
        line["parity_vs_single_gpu"] = parity
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
