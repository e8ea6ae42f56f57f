"""Load-balancing hooks (housekeeping/balance.F90): the host decision logic
(calculate_breaks, get_load weighting) and the oracle's part_load_func; the device parity test is
in test_gpu_parity.py and the two-GPU redistribution in test_multi_gpu.py."""
import numpy as np
import pytest

from helpers import make_plasma
from epoch_b200 import balance


def loads_of(mins, maxs, load, ng):
    return [int(load[a + ng - 1:b + ng].sum()) for a, b in zip(mins, maxs)]


def test_calculate_breaks_uniform_load():
    ng, ncm = 5, 3
    load = np.zeros(64 + 2 * ng, dtype=np.int64)
    load[ng:-ng] = 100
    mins, maxs = balance.calculate_breaks(load, 4, ng, ncm)
    assert mins == [1, 17, 33, 49] and maxs == [16, 32, 48, 64]
    assert balance.calculate_breaks(load, 1, ng, ncm) == ([1], [64])


def test_calculate_breaks_hand_example():
    """12 cells, load 9 in the first four and 1 elsewhere, 2 processors, worked by hand from
    balance.F90:2526-2669: ideal = 22; the running sum passes it in cell 3 (18 -> 27; 18 is
    closer, so maxs(1) = 2, widened to 4 by nextra); the -1 sweeps take it to 3 (27|17, spread
    10) and 2 (18|26, spread 8); +1 back to 3 is no better, so the sweeps stop at 2; the final
    forward check enforces ncell_min = 3 cells and leaves the break after cell 3"""
    ng, ncm = 2, 3
    load = np.zeros(12 + 2 * ng, dtype=np.int64)
    load[ng:ng + 4] = 9
    load[ng + 4:ng + 12] = 1
    mins, maxs = balance.calculate_breaks(load, 2, ng, ncm)
    assert mins[0] == 1 and maxs[1] == 12 and mins[1] == maxs[0] + 1
    assert maxs[0] == 3
    assert loads_of(mins, maxs, load, ng) == [27, 17]


@pytest.mark.parametrize("seed", range(8))
def test_calculate_breaks_invariants(seed):
    """the ranges tile 1..n, every domain keeps ncell_min cells, and the split is never worse
    than the even one the run started with"""
    rng = np.random.default_rng(seed)
    ng, ncm = 5, 3
    n, nproc = int(rng.integers(40, 200)), int(rng.integers(2, 9))
    load = np.zeros(n + 2 * ng, dtype=np.int64)
    centre, width = rng.uniform(0, n), rng.uniform(3, n / 2)
    x = np.arange(n)
    load[ng:-ng] = (5 * 1000 * np.exp(-((x - centre) / width) ** 2)).astype(np.int64) + 50
    mins, maxs = balance.calculate_breaks(load, nproc, ng, ncm)
    assert mins[0] == 1 and maxs[-1] == n
    assert all(mins[p] == maxs[p - 1] + 1 for p in range(1, nproc))
    assert all(b - a + 1 >= ncm for a, b in zip(mins, maxs))
    even = [(p * n // nproc + 1, (p + 1) * n // nproc) for p in range(nproc)]
    worst_even = max(loads_of([e[0] for e in even], [e[1] for e in even], load, ng))
    assert max(loads_of(mins, maxs, load, ng)) <= worst_even


def test_global_load_weighting():
    ng = 4
    local = [np.arange(6 + 2 * ng, dtype=np.int64), np.ones(3 + 2 * ng, dtype=np.int64)]
    out = balance.global_load(local, [6, 3], ng, allreduce=lambda a: 2 * a)
    assert np.array_equal(out[0][:ng], 10 * np.arange(ng)) and out[0][ng] == 10 * ng + 3
    assert out[1][ng] == 10 + 6 and out[1][-1] == 10


@pytest.mark.parametrize("ndims,shape", [(1, "triangle"), (2, "tophat"), (3, "triangle"), (3, "bspline3")])
def test_get_load_oracle(oracle_abi, ndims, shape):
    """part_load_func against a numpy histogram; the total per axis is the particle count"""
    sim = make_plasma(oracle_abi, ndims, shape, [12, 8, 6], seed=2, ppc=5)
    loads = sim.ctx.get_load(sim.grid_min, sim.n_global)
    ng = sim.ctx.ng
    P = np.concatenate(sim.loads)
    for a in range(ndims):
        r = (P[:, a] - sim.grid_min[a]) / sim.d[a]
        cell = np.floor(r).astype(int) + 1 if shape == "tophat" else np.floor(r + 1.5).astype(int)
        want = np.bincount(cell + ng - 1, minlength=sim.n_global[a] + 2 * ng)
        assert np.array_equal(loads[a], want)
        assert loads[a].sum() == len(P)


def test_single_rank_balance_is_a_no_op(oracle_abi):
    sim = make_plasma(oracle_abi, 2, "triangle", [12, 8], ppc=2)
    assert sim.balance_workload() is False
    with pytest.raises(Exception):
        sim.ctx.redistribute([[1], [1]], [[12], [8]], [[1], [2]], [[12], [8]], sim.grid_min, sim.grid_min_local,
                             sim.min_local, sim.max_local)
