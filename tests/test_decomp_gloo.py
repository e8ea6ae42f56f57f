"""-m "not gpu": host-side logic of the N > 1 path on CPU -- world_size-2 `gloo`
processes check that the Cartesian decomposition mirrored from
mpi_routines.F90:44-593 is consistent across ranks (cell ranges tile the global
grid, neighbour tables are mutual, particle ownership partitions a global load)."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, nprocdir, n_global, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from epoch_b200 import decomp
    coords = decomp.rank_to_coords(rank, nprocdir)
    assert decomp.coords_to_rank(coords, nprocdir) == rank
    rng_cells = [decomp.cell_range(n_global[d], nprocdir[d], coords[d]) for d in range(3)]
    nbr = {(d, s): decomp.neighbour(coords, nprocdir, d, s, True) for d in range(3) for s in (-1, 1)}
    # a global seeded load; ownership by the reference's rule x_min_local <= x < x_max_local
    dx = 1.0
    rng = np.random.default_rng(5)
    P = rng.uniform(0, 1, (20000, 3)) * np.array(n_global) * dx
    own = np.ones(len(P), bool)
    for d in range(3):
        lo, hi = rng_cells[d]
        own &= (P[:, d] >= (lo - 1) * dx) & (P[:, d] < hi * dx)
    info = [None] * world
    dist.all_gather_object(info, (rank, coords, rng_cells, nbr, np.flatnonzero(own)))
    if rank == 0:
        covered = np.zeros(n_global, int)
        owners = np.zeros(len(P), int)
        for (r, c, cells, nb, idx) in info:
            covered[tuple(slice(cells[d][0] - 1, cells[d][1]) for d in range(3))] += 1
            owners[idx] += 1
            for (d, s), other in nb.items():
                # my +s neighbour's -s neighbour is me
                assert info[other][3][(d, -s)] == r
            for d in range(3):
                n = cells[d][1] - cells[d][0] + 1
                assert abs(n - n_global[d] / nprocdir[d]) < 1.0    # near-even split
        q.put((bool((covered == 1).all()), bool((owners == 1).all())))
    dist.barrier()
    dist.destroy_process_group()


def _run(nprocdir, n_global, port):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world = int(np.prod(nprocdir))
    procs = [ctx.Process(target=_worker, args=(r, world, port, nprocdir, n_global, q)) for r in range(world)]
    [p.start() for p in procs]
    res = q.get(timeout=120)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert res == (True, True)


def test_slab_world2():
    _run((1, 1, 2), (16, 12, 21), 29701)


def test_x_split_world2():
    _run((2, 1, 1), (17, 8, 8), 29702)
