"""CPU check of the protocol that lets the launches of consecutive steps overlap (fluid_sim_b200/csrc/chain.cu,
DESIGN.md section 4 "Overlapping steps").

fsb_dbg_chain_describe returns the phase table the library builds for the one-launch step -- items, dependencies inside
the launch, dependencies on the previous launch, the arrays every phase touches -- without a device.  This test turns two
consecutive launches into a happens-before graph (those dependencies, the advection phases' wait for the gradient blocks
in reach, program order of the items one CTA runs) and checks that EVERY pair of accesses to the same array that
overlap in columns, at least one of them a write, is ordered: within a launch across phases, and between launch n and
launch n+1 with the arrays rotated as chained_step rotates them.  (Launch n+2 starts only after launch n has finished
altogether, which every CTA checks at its start; items of one sweep phase synchronise through their rings.)
"""
import ctypes as C

import numpy as np
import pytest

import fluid_sim_b200 as fsb

SOR, GS, DIV, GRAD, ADV, GSX0, BNDP, NOISE = 1, 2, 3, 4, 5, 6, 7, 8
WORDS = 53


def describe(w, h, d, p, nsm=148, variant=0, block=(0, 0, 0)):
    lib = fsb.load_library()
    out = (C.c_longlong * (24 * WORDS))()
    n, grid, cw = C.c_int(), C.c_int(), C.c_int()
    rc = lib.fsb_dbg_chain_describe(variant, w, h, d, p, nsm, block[0], block[1], block[2], out, 24 * WORDS, C.byref(n),
                                    C.byref(grid), C.byref(cw))
    if rc != 0:
        return None
    a = np.array(out[:n.value * WORDS], dtype=np.int64).reshape(n.value, WORDS)
    phases = []
    for r in a:
        phases.append(dict(kind=int(r[0]), nitems=int(r[1]), ncl=int(r[2]), nf=int(r[3]), cpv=int(r[4]), shift=int(r[5]),
                           deps=[tuple(int(v) for v in r[8 + 4 * k:12 + 4 * k]) for k in range(int(r[6]))],
                           xdeps=[tuple(int(v) for v in r[24 + 4 * k:28 + 4 * k]) for k in range(int(r[7]))],
                           ptr=[int(v) for v in r[36:44]], wf=[[int(v) for v in r[44:47]], [int(v) for v in r[47:50]]],
                           reach=int(r[50]), item0=int(r[51]), push_cols=int(r[52])))
    return phases, grid.value, cw.value


# the arrays behind the roles of a launch, rotated as chained_step (fluid_sim_b200.cu) rotates them
def physical(launch):
    vel = [("Q0", "Q2", "Q3"), ("Q3", "Q0", "Q1"), ("Q1", "Q3", "Q2")][launch % 3 if launch < 3 else 0]
    den = [("Da", "Db"), ("Db", "Dc"), ("Dc", "Da")][launch % 3]
    m = {1: vel[0] + ".u", 2: vel[0] + ".v", 3: vel[1] + ".u", 4: vel[1] + ".v", 5: vel[2] + ".u", 6: vel[2] + ".v",
         7: "P1.p", 8: "P1.div", 9: den[0], 10: den[1], 11: "P2.p", 12: "P2.div"}
    return m


def dep_items(ph, cl, dep, phases):
    """items of phase dep[0] an item with column range index `cl` of phase `ph` waits for (chain_wait in chain.cu)"""
    q = phases[dep[0]]
    b0 = cl * ph["cpv"] - (ph["shift"] if dep[1] else 0)
    b1 = b0 + ph["cpv"] - 1
    a = (b0 - dep[2] + q["shift"]) // q["cpv"]
    b = (b1 + dep[3] + q["shift"]) // q["cpv"]
    a, b = max(a, 0), min(b, q["ncl"] - 1)
    return [x for x in range(a * q["nf"], (b + 1) * q["nf"])]


def accesses(ph, item, H, reach_cols, density_sources):
    """[(role, lo, hi, is_write)] of one item; rows are not modelled (every item covers all rows of its columns)"""
    cl, fi = item // ph["nf"], item % ph["nf"]
    k = ph["kind"]
    b0, b1 = cl * ph["cpv"], min(cl * ph["cpv"] + ph["cpv"] - 1, H - 1)
    acc = []

    def walls(lo, hi):      # columns whose wall cells the block writes: its own, and a wall column whose interior neighbour it owns
        cols = [(lo, hi)]
        if lo <= 1 <= hi:
            cols.append((0, 0))
        if lo <= H - 2 <= hi:
            cols.append((H - 1, H - 1))
        return cols

    u, v, q, dst_a, dst_b, src_a, src_b, bnd_extra = ph["ptr"]
    if k == NOISE:
        acc += [(src_a, b0, b1, False), (src_b, b0, b1, False), (u, b0, b1, True), (v, b0, b1, True)]
        if density_sources:
            acc += [(9, b0, b1, True)]
    elif k in (SOR, GS, GSX0):
        x0, rhs, out = ph["wf"][fi]
        if b0 <= H - 1:
            for r in (x0, rhs):
                if r:
                    acc.append((r, b0, b1, False))
        lo, hi = max(b0 - ph["shift"], 1), min(b0 - ph["shift"] + ph["cpv"] - 1, H - 2)
        if lo <= hi:
            acc.append((out, lo, hi, True))
        acc.append((("ring", id(ph)), cl, cl, True))                 # to the item on the right
        if cl > 0:
            acc.append((("ring", id(ph)), cl - 1, cl - 1, False))    # from the item on the left
    elif k == DIV:
        acc += [(u, max(b0 - 1, 0), min(b1 + 1, H - 1), False), (v, max(b0 - 1, 0), min(b1 + 1, H - 1), False), (q, b0, b1, True)]
    elif k == BNDP:
        acc += [(q, b0, b1, False)] + [(q, lo, hi, True) for lo, hi in walls(b0, b1)]
    elif k == GRAD:
        acc += [(q, max(b0 - 1, 0), min(b1 + 1, H - 1), False)]
        for f in (u, v):
            acc += [(f, b0, b1, False)] + [(f, lo, hi, True) for lo, hi in walls(b0, b1)]
        if bnd_extra:
            acc += [(bnd_extra, b0, b1, False)] + [(bnd_extra, lo, hi, True) for lo, hi in walls(b0, b1)]
    elif k == ADV:
        lo, hi = max(b0 - reach_cols, 0), min(b1 + reach_cols, H - 1)
        for f in (u, v, src_a, src_b):
            if f:
                acc.append((f, lo, hi, False))
        for f in (dst_a, dst_b):
            if f:
                acc += [(f, a, b, True) for a, b in walls(b0, b1)]
    return acc


def check(phases, grid, h, reach_cols, density_sources, launches=(0, 1), rotate=True):
    # nodes: (launch, phase, item) for two consecutive launches (or one)
    index, nodes = {}, []
    for L in launches:
        for pi, ph in enumerate(phases):
            for it in range(ph["nitems"]):
                index[(L, pi, it)] = len(nodes)
                nodes.append((L, pi, it))
    n = len(nodes)
    pred = [set() for _ in range(n)]
    for (L, pi, it), me in index.items():
        ph = phases[pi]
        cl = it // ph["nf"]
        for dep in ph["deps"]:
            for x in dep_items(ph, cl, dep, phases):
                pred[me].add(index[(L, dep[0], x)])
        if L == 1:
            for dep in ph["xdeps"]:
                for x in dep_items(ph, cl, dep, phases):
                    pred[me].add(index[(0, dep[0], x)])
        if ph["kind"] == ADV:
            # chain_advect_wait: the gradient blocks [b - k, b + k] with k blocks covering the reach
            gr = phases[ph["reach"]]
            bg = cl * ph["cpv"] // gr["cpv"]
            k = max(1, -(-reach_cols // gr["cpv"]))
            for x in range(max(bg - k, 0), min(bg + k, gr["ncl"] - 1) + 1):
                pred[me].add(index[(L, ph["reach"], x)])
    # Sweep phases: item c stores into the last `shift` columns of item c-1's range (the result shift) -- values that
    # reach it through the ring from item c-1, which produced them after its own dependencies were met.  So whatever
    # item c-1 of the same field waited for also precedes item c's stores.
    for L in launches:
        for pi, ph in enumerate(phases):
            if ph["kind"] in (SOR, GS, GSX0):
                for it in range(ph["nf"], ph["nitems"]):
                    pred[index[(L, pi, it)]] |= pred[index[(L, pi, it - ph["nf"])]]
    # program order of the items one CTA runs (item i of every phase on CTA i mod grid, phases in order)
    for L in launches:
        last = {}
        for pi, ph in enumerate(phases):
            for it in range(ph["nitems"]):
                cta = it % grid
                if cta in last:
                    pred[index[(L, pi, it)]].add(last[cta])
                last[cta] = index[(L, pi, it)]
    # reachability (nodes are in a topological order: launch, phase, item; same-phase items never depend on each other)
    reach = np.zeros((n, n), dtype=bool)
    for me in range(n):
        for q in pred[me]:
            assert q < me or nodes[q][:2] == nodes[me][:2], "dependency on a later item"
            reach[me] |= reach[q]
            reach[me, q] = True
    # accesses by physical array
    by_array = {}
    for (L, pi, it), me in index.items():
        roles = physical(L) if rotate else {k: f"A{k}" for k in range(1, 13)}
        for role, lo, hi, wr in accesses(phases[pi], it, h, reach_cols, density_sources):
            arr = (roles[role] if isinstance(role, int) else role)
            by_array.setdefault(arr, []).append((me, lo, hi, wr))
    bad = []
    for arr, acc in by_array.items():
        acc.sort()
        for i, (a, alo, ahi, awr) in enumerate(acc):
            for (b, blo, bhi, bwr) in acc[i + 1:]:
                if not (awr or bwr) or ahi < blo or bhi < alo or a == b:
                    continue
                (La, pa, _), (Lb, pb, _) = nodes[a], nodes[b]
                if La == Lb and pa == pb:
                    continue            # items of one phase: independent columns, or a sweep phase's own pipeline
                if not reach[b, a]:
                    bad.append((arr, nodes[a], (alo, ahi, "W" if awr else "R"), nodes[b], (blo, bhi, "W" if bwr else "R")))
    return bad


@pytest.mark.parametrize("w,h,d,p", [(512, 2048, 5, 40), (256, 1024, 10, 100), (128, 640, 5, 20), (64, 7232, 5, 20)])
@pytest.mark.parametrize("reach_cols", [3, 70, 100000])
@pytest.mark.parametrize("density_sources", [False, True])
def test_every_conflicting_access_is_ordered(w, h, d, p, reach_cols, density_sources):
    # (a launch that carries density sources has one more dependency: the noise phase, which adds them to the previous
    # step's density where it lies, waits for the previous launch's density advection of the block)
    got = describe(w, h, d, p, variant=3 if density_sources else 0)
    assert got is not None, "no chained configuration"
    phases, grid, cw = got
    assert len(phases[0]["xdeps"]) == (3 if density_sources else 2)
    bad = check(phases, grid, h, reach_cols, density_sources)
    assert not bad, f"{len(bad)} unordered pairs, first: {bad[:3]}"


def test_the_check_sees_a_missing_dependency():
    """Without the second gradient phase's wait for the previous launch's density advection the density walls it
    rewrites (and the next density advection's source) would race with that advection: the check must say so."""
    phases, grid, cw = describe(512, 2048, 5, 40)
    g2 = [q for q in phases if q["kind"] == GRAD][-1]
    assert g2["xdeps"]
    g2["xdeps"] = []
    assert check(phases, grid, 2048, 3, False)
    # ... and without the noise phase's wait for the previous launch's second gradient phase the step would start from
    # a velocity that is still being written
    phases, grid, cw = describe(512, 2048, 5, 40)
    phases[0]["xdeps"] = []
    assert check(phases, grid, 2048, 3, False)


@pytest.mark.parametrize("w,h,d,p", [(512, 2048, 5, 40), (256, 1024, 10, 100), (64, 7232, 5, 20)])
@pytest.mark.parametrize("reach_cols", [3, 100000])
def test_one_launch_after_the_noise(w, h, d, p, reach_cols):
    """The whole step after the noise as one launch (chain level 2 without the overlap, and what column blocks of a
    multi-GPU run do): the advected velocity is written where the first projection had its pressure and divergence, so
    the advection phase also waits for the gradient blocks and sweep items that still read them."""
    got = describe(w, h, d, p, variant=1)
    assert got is not None
    phases, grid, cw = got
    assert phases[0]["kind"] == SOR
    bad = check(phases, grid, h, reach_cols, False, launches=(0,), rotate=False)
    assert not bad, f"{len(bad)} unordered pairs, first: {bad[:3]}"


def test_plan_of_the_headline_workload():
    """8192^2, 40 projection sweeps on 148 SMs: ten phases (noise, diffusion, 2 x (divergence, one sweep pass, gradient),
    two advections), 64-column blocks, 130 CTAs; 100 sweeps take two passes per projection with a wall update of the
    pressure between them."""
    phases, grid, cw = describe(8192, 8192, 5, 40)
    assert [q["kind"] for q in phases] == [NOISE, SOR, DIV, GS, GRAD, ADV, DIV, GS, GRAD, ADV]
    assert (grid, cw) == (130, 64)
    assert phases[1]["nitems"] == 130 and phases[3]["nitems"] == 129 and phases[0]["nitems"] == 128
    phases, grid, cw = describe(32768, 4096, 5, 100)
    kinds = [q["kind"] for q in phases]
    assert kinds.count(GS) == 2 and kinds.count(GSX0) == 2 and kinds.count(BNDP) == 2 and cw == 32
    assert describe(100, 100, 5, 33) is None          # sweep counts without a chained configuration


def publishes_to(owner, ph, item, side):
    """Does the rank `owner` (its block description) publish the done-word of its `item` of phase `ph` in the arrays of
    its left (-1) / right (+1) neighbour?  (the publish rule at the end of chain_kernel's item loop)"""
    j0, j1, halo = owner["block"]
    cl = item // ph["nf"]
    lo = cl * ph["cpv"] - ph["shift"]
    hi = lo + ph["cpv"] - 1
    reach = halo + 2 * ph["cpv"]
    return lo < j0 + reach if side < 0 else hi >= j1 - reach


@pytest.mark.parametrize("h,ncols,halo,p,d", [(32768, 4096, 128, 100, 5), (16384, 4096, 128, 40, 5), (2048, 512, 128, 20, 5),
                                              (1024, 512, 256, 20, 10), (8192, 4096, 64, 20, 5)])
def test_column_blocks_only_wait_for_items_that_publish_there(h, ncols, halo, p, d):
    """Multi-GPU: a rank's items are numbered along the whole grid and wait for done-words in the rank's own array, which
    its own items and the neighbours' items near the boundary set.  Every item a dependency (or an advection phase's
    wait for the gradient blocks it can see) names must be one of those, or the wait could never end; and the global
    item numbers must fit the done[] array."""
    w = 256
    ranks = []
    for r in range(h // ncols):
        got = describe(w, h, d, p, variant=2, block=(r * ncols, (r + 1) * ncols, halo))
        assert got is not None
        ranks.append(dict(phases=got[0], cw=got[2], block=(r * ncols, (r + 1) * ncols, halo)))

    def owner_of(pi, x):
        for r, R in enumerate(ranks):
            q = R["phases"][pi]
            if q["item0"] <= x < q["item0"] + q["nitems"]:
                return r
        return None

    for r, R in enumerate(ranks):
        phases = R["phases"]
        j0, j1, _ = R["block"]
        for pi, ph in enumerate(phases):
            assert ph["item0"] + ph["nitems"] <= 1088
            for li in range(ph["nitems"]):
                item = ph["item0"] + li
                cl = item // ph["nf"]
                awaited = [(dep[0], x) for dep in ph["deps"] for x in dep_items(ph, cl, dep, phases)]
                if ph["kind"] == ADV:
                    gr = phases[ph["reach"]]
                    cw = R["cw"]
                    vis_lo = max((j0 - halo) // cw, 0) if r > 0 else 0
                    vis_hi = min((j1 + halo) // cw - 1, gr["ncl"] - 1) if r < len(ranks) - 1 else gr["ncl"] - 1
                    awaited += [(ph["reach"], x) for x in range(vis_lo, vis_hi + 1)]      # the widest wait: every visible block
                for (qi, x) in awaited:
                    o = owner_of(qi, x)
                    assert o is not None, f"rank {r} phase {pi} item {item} waits for item {x} of phase {qi}, which no rank runs"
                    if o == r:
                        continue
                    assert abs(o - r) == 1, f"rank {r} phase {pi} item {item} waits for item {x} of phase {qi}, owned by rank {o}"
                    assert publishes_to(ranks[o], ranks[o]["phases"][qi], x, r - o), \
                        f"rank {r} phase {pi} item {item} waits for item {x} of phase {qi}: rank {o} does not publish it there"


# ---- column blocks of a multi-GPU run: the same check across ranks ------------------------------------------------------
# Every rank runs the one launch after the noise on its block of columns (stored with `halo` columns either side) and
# ONLY READS ITS OWN ARRAYS: items near a rank boundary store their results in the neighbour's array as well (pointwise
# phases: within `push_cols` columns of the boundary; sweep phases: the result columns up to the left rank's first halo
# column) and publish their done-word there; an item that writes into a neighbour first waits for that neighbour to have
# finished the previous step; a rank's own launches follow one another.  Arrays swap roles from step to step as
# fsb_blk_step swaps them.

def mr_accesses(R, ph, item, H, reach_cols, step):
    """[(array rank, array name, lo, hi, is_write)] of one item of rank R (dict with index, block, neighbours)"""
    r, (j0, j1, halo) = R["r"], R["block"]
    nranks = R["nranks"]
    has_left, has_right = r > 0, r < nranks - 1
    lo_st, hi_st = max(j0 - halo, 0), min(j1 + halo, H) - 1           # columns this rank stores
    name = {3: f"F{step % 2}.u", 4: f"F{step % 2}.v", 7: f"F{1 - step % 2}.u", 8: f"F{1 - step % 2}.v", 9: f"D{step % 2}",
            10: f"D{1 - step % 2}", 11: "P2.p", 12: "P2.div"}
    out = []
    push_l = push_r = False
    cl = item // ph["nf"]
    if ph["kind"] in (SOR, GS, GSX0):
        push_l = has_left and cl * ph["cpv"] - ph["shift"] < j0 + 1
    elif ph["push_cols"] > 0:
        b0 = cl * ph["cpv"]
        push_l = has_left and b0 < j0 + ph["push_cols"]
        push_r = has_right and b0 + ph["cpv"] - 1 >= j1 - ph["push_cols"]
    for role, lo, hi, wr in accesses(ph, item, H, reach_cols, False):
        if not isinstance(role, int):
            out.append((-1, role, lo, hi, wr))                        # a ring of the sweep chain: one object for all ranks
            continue
        lo2, hi2 = max(lo, lo_st), min(hi, hi_st)
        if lo2 <= hi2:
            out.append((r, name[role], lo2, hi2, wr))
        if wr and ph["kind"] in (SOR, GS, GSX0):
            if push_l and lo <= j0:                                   # result columns < j0 + 1 also go to the left rank
                out.append((r - 1, name[role], lo, min(hi, j0), True))
        elif wr:
            if push_l:
                out.append((r - 1, name[role], lo, hi, True))
            if push_r:
                out.append((r + 1, name[role], lo, hi, True))
    return out, push_l, push_r


def mr_check(h, ncols, halo, p, d, reach_cols, step_words=True, cross_rank_done=True):
    w = 256
    nranks = h // ncols
    ranks = []
    for r in range(nranks):
        got = describe(w, h, d, p, variant=2, block=(r * ncols, (r + 1) * ncols, halo))
        assert got is not None
        ranks.append(dict(r=r, nranks=nranks, phases=got[0], grid=got[1], cw=got[2], block=(r * ncols, (r + 1) * ncols, halo)))
    nph = len(ranks[0]["phases"])
    index, nodes = {}, []
    for s in (0, 1):
        for R in ranks:
            for pi, ph in enumerate(R["phases"]):
                for li in range(ph["nitems"]):
                    key = (s, pi, ph["item0"] + li)                    # global item numbers: one owner each
                    assert key not in index
                    index[key] = len(nodes)
                    nodes.append((s, R["r"], pi, ph["item0"] + li))
        for R in ranks:                                               # "rank r has finished step s"
            index[("end", s, R["r"])] = len(nodes)
            nodes.append((s, R["r"], nph, -1))
    n = len(nodes)
    pred = [set() for _ in range(n)]
    acc_of = {}
    for s in (0, 1):
        for R in ranks:
            phases, (j0, j1, _) = R["phases"], R["block"]
            for pi, ph in enumerate(phases):
                for li in range(ph["nitems"]):
                    item = ph["item0"] + li
                    me = index[(s, pi, item)]
                    cl = item // ph["nf"]
                    for dep in ph["deps"]:
                        for x in dep_items(ph, cl, dep, phases):
                            if cross_rank_done or nodes[index[(s, dep[0], x)]][1] == R["r"]:
                                pred[me].add(index[(s, dep[0], x)])
                    if ph["kind"] == ADV:
                        gr = phases[ph["reach"]]
                        cw = R["cw"]
                        vis_lo = max((j0 - halo) // cw, 0) if R["r"] > 0 else 0
                        vis_hi = min((j1 + halo) // cw - 1, gr["ncl"] - 1) if R["r"] < nranks - 1 else gr["ncl"] - 1
                        bg = cl * ph["cpv"] // gr["cpv"]
                        k = max(1, -(-reach_cols // gr["cpv"]))
                        for x in range(max(bg - k, vis_lo), min(bg + k, vis_hi) + 1):
                            pred[me].add(index[(s, ph["reach"], x)])
                    acc, push_l, push_r = mr_accesses(R, ph, item, h, reach_cols, s)
                    acc_of[me] = acc
                    pred[index[("end", s, R["r"])]].add(me)
                    if s == 1:
                        pred[me].add(index[("end", 0, R["r"])])        # a rank's launches follow one another
                        if push_l and step_words:
                            pred[me].add(index[("end", 0, R["r"] - 1)])   # step word of the neighbour it writes into
                        if push_r and step_words:
                            pred[me].add(index[("end", 0, R["r"] + 1)])
    # the ring order inside a sweep phase, across ranks as well (items are numbered along the whole grid)
    for s in (0, 1):
        for pi, ph in enumerate(ranks[0]["phases"]):
            if ph["kind"] in (SOR, GS, GSX0):
                total = max(R["phases"][pi]["item0"] + R["phases"][pi]["nitems"] for R in ranks)
                for it in range(ph["nf"], total):
                    pred[index[(s, pi, it)]] |= {q for q in pred[index[(s, pi, it - ph["nf"])]] if nodes[q][2] != nph}
    # program order of the items one CTA of a rank runs
    for s in (0, 1):
        for R in ranks:
            last = {}
            for pi, ph in enumerate(R["phases"]):
                for li in range(ph["nitems"]):
                    cta = li % R["grid"]
                    me = index[(s, pi, ph["item0"] + li)]
                    if cta in last:
                        pred[me].add(last[cta])
                    last[cta] = me
    # reachability by depth-first search from every node (the node order is not topological across ranks)
    order, state = [], [0] * n
    for root in range(n):
        if state[root]:
            continue
        stack = [(root, iter(pred[root]))]
        state[root] = 1
        while stack:
            v, it = stack[-1]
            nxt = next(it, None)
            if nxt is None:
                state[v] = 2
                order.append(v)
                stack.pop()
            elif state[nxt] == 0:
                state[nxt] = 1
                stack.append((nxt, iter(pred[nxt])))
            else:
                assert state[nxt] == 2, "cyclic dependencies"
    reach = np.zeros((n, n), dtype=bool)
    for v in order:                                                   # predecessors first
        for q in pred[v]:
            reach[v] |= reach[q]
            reach[v, q] = True
    by_array = {}
    for me, acc in acc_of.items():
        for arr_rank, arr, lo, hi, wr in acc:
            by_array.setdefault((arr_rank, arr), []).append((me, lo, hi, wr))
    bad = []
    for arr, acc in by_array.items():
        for i, (a, alo, ahi, awr) in enumerate(acc):
            for (b, blo, bhi, bwr) in acc[i + 1:]:
                if not (awr or bwr) or ahi < blo or bhi < alo or a == b:
                    continue
                (sa, _, pa, _), (sb, _, pb, _) = nodes[a], nodes[b]
                if sa == sb and pa == pb:
                    continue
                if not (reach[b, a] or reach[a, b]):
                    bad.append((arr, nodes[a], (alo, ahi, "W" if awr else "R"), nodes[b], (blo, bhi, "W" if bwr else "R")))
    return bad


@pytest.mark.parametrize("h,ncols,halo,p,d", [(1536, 512, 128, 20, 5), (3072, 1024, 128, 40, 10)])
@pytest.mark.parametrize("reach_cols", [3, 70])
def test_column_blocks_conflicting_accesses_are_ordered(h, ncols, halo, p, d, reach_cols):
    bad = mr_check(h, ncols, halo, p, d, reach_cols)
    assert not bad, f"{len(bad)} unordered pairs, first: {bad[:4]}"


def test_the_cross_rank_check_sees_missing_synchronisation():
    """Without the step words a rank's next step would store into a neighbour's array that the neighbour's current step
    still reads; without the done-words published across the boundary nothing orders a rank's reads after the
    neighbour's stores."""
    assert mr_check(1536, 512, 128, 20, 5, 3, step_words=False)
    assert mr_check(1536, 512, 128, 20, 5, 3, cross_rank_done=False)
