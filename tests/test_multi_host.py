"""CPU tests (gloo, world_size 2) of the host-side logic of the multi-GPU path: column ownership, the launch plan all
ranks must agree on, and the halo / result-column movement between neighbouring ranks (fluid_sim_b200/multi.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fluid_sim_b200.multi import column_range, exchange_halo, plan_diffusion, plan_projection, return_columns


def test_column_ranges_tile_the_grid():
    for world in (1, 2, 4, 8):
        edges = [column_range(32768, r, world) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == 32768
        assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
    with pytest.raises(ValueError):
        column_range(1000, 0, 3)
    with pytest.raises(ValueError):
        column_range(96 * 2, 0, 2, align=64)


def test_projection_plan_covers_all_sweeps():
    for p in (0, 1, 7, 20, 40, 100, 133):
        plan = plan_projection(p)
        assert sum(lv * st for lv, st in plan) == p
        assert all(lv in (1, 2, 5, 10) and 1 <= st <= 5 for lv, st in plan)
    assert plan_projection(40) == [(10, 4)]
    assert plan_projection(100) == [(10, 5), (10, 5)]


def test_diffusion_plan_covers_all_sweeps():
    for d in range(0, 41):
        plan = plan_diffusion(d)
        assert sum(lv * st for lv, st in plan) == d
        assert all(lv in (1, 2, 5, 10, 20) and st == 1 for lv, st in plan)
    assert plan_diffusion(5) == [(5, 1)] and plan_diffusion(7) == [(5, 1), (2, 1)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, rows, ncols, halo, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        H = ncols * world
        full = torch.arange(rows * H, dtype=torch.float32).reshape(rows, H)
        j0 = rank * ncols
        pitch = ncols + 2 * halo
        loc = torch.full((rows, pitch), -1.0)
        loc[:, halo:halo + ncols] = full[:, j0:j0 + ncols]
        # halo exchange of width 3
        exchange_halo([loc], halo, ncols, 3, rank, world)
        ok = True
        if rank > 0:
            ok &= torch.equal(loc[:, halo - 3:halo], full[:, j0 - 3:j0])
        else:
            ok &= bool((loc[:, :halo] == -1).all())
        if rank < world - 1:
            ok &= torch.equal(loc[:, halo + ncols:halo + ncols + 3], full[:, j0 + ncols:j0 + ncols + 3])
        else:
            ok &= bool((loc[:, halo + ncols:] == -1).all())
        # result columns computed by the right neighbour (stored in ITS left halo) come back, interior rows only
        marker = 1000.0 * (rank + 1)
        loc[:, :halo] = marker
        return_columns([loc], halo, ncols, 2, rank, world)
        if rank < world - 1:
            ok &= bool((loc[1:-1, halo + ncols - 2:halo + ncols] == 1000.0 * (rank + 2)).all())
            ok &= torch.equal(loc[0, halo + ncols - 2:halo + ncols], full[0, j0 + ncols - 2:j0 + ncols])
        else:
            ok &= torch.equal(loc[:, halo + ncols - 2:halo + ncols], full[:, j0 + ncols - 2:j0 + ncols])
        out[rank] = 1 if ok else 0
    finally:
        dist.destroy_process_group()


def test_halo_movement_two_ranks_gloo():
    world = 2
    out = mp.Array("i", [0] * world)
    port = _free_port()
    procs = [mp.Process(target=_worker, args=(r, world, port, 6, 8, 4, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert list(out) == [1] * world
