"""Moving window (housekeeping/window.F90): oracle known answers; the CUDA parity test is in
test_gpu_parity.py (it imports the deck from here)."""
import numpy as np
import pytest

from helpers import make_plasma, FIELDS6, ALL9
from epoch_b200.driver import C


def window_insert(ppc=2, seed=77):
    """stand-in for insert_particles: ppc particles per cell of the incoming column, seeded per shift"""
    count = [0]

    def insert(sim, x0):
        count[0] += 1
        rng = np.random.default_rng([seed, count[0], sim.rank])
        nd = sim.ndims
        ncell = int(np.prod(sim.n_local[1:])) if nd > 1 else 1
        out = []
        for isp, sp in enumerate(sim.species):
            n = ppc * ncell
            P = np.empty((n, nd + 4))
            P[:, 0] = x0 + rng.random(n) * sim.d[0]
            for d in range(1, nd):
                P[:, d] = rng.uniform(sim.min_local[d], sim.max_local[d], n)
            P[:, nd:nd + 3] = rng.normal(0.0, 0.05, (n, 3)) * sp.mass * C
            P[:, nd + 3] = rng.uniform(0.5, 1.5, n) * 1e25 * 1e-6 ** nd / ppc    # unique: the tests pair by weight
            out.append(P)
        return out
    return insert


def window_deck(abi, ndims, shape="triangle", planes=True, fbc="open", **kw):
    n = [20, 12, 8]
    bc_field = [fbc, fbc] + ["periodic"] * (2 * ndims - 2)
    bcs = ["open", "open"] + ["periodic"] * (2 * ndims - 2)
    sim = make_plasma(abi, ndims, shape, n, seed=12, ppc=2, u_th=0.2, bc_field=bc_field, bcs=bcs, **kw)
    pl = None
    if planes:
        rng = np.random.default_rng(5)
        shp = sim.ctx._np_shape()[:-1]           # the y-z extents with ghosts ([z, y] order)
        pl = {k: rng.normal(0, 1e9 if k[0] == "e" else 3.0, shp if shp else (1,)) for k in FIELDS6}
    sim.enable_window(v_x=0.9 * C, insert=window_insert(), planes=pl)
    return sim


@pytest.mark.parametrize("ndims", [1, 2, 3])
def test_shift_fields_known_answer(oracle_abi, ndims):
    """after one shift every array holds its old right neighbour (away from the x_max fix-up), the
    incoming cells follow window.F90:147-165, and particles left of the new x_min are gone"""
    sim = window_deck(oracle_abi, ndims)
    sim.ctx.set_time(sim.dt, 0.0, 0)
    before = sim.ctx.download_fields(list(ALL9))
    n_before = [sim.ctx.particle_count(i) for i in range(len(sim.species))]
    x_before = [sim.ctx.download_particles(i)[:, 0].copy() for i in range(len(sim.species))]
    sim.window["insert"] = None
    sim._shift_window(1)
    after = sim.ctx.download_fields(list(ALL9))
    ng = sim.ctx.ng
    nx = sim.n_local[0]
    for k in ALL9:
        # interior rows of the other axes (their ghosts are refreshed by field_bc after the shift)
        inner = tuple(slice(ng, -ng) for _ in range(ndims - 1))
        a, b = after[k][inner], before[k][inner]
        # x indices 1-ng .. nx-3 (array positions 0 .. nx+ng-4) are plain copies of i+1
        assert np.array_equal(a[..., :nx + ng - 3], b[..., 1:nx + ng - 2]), k
    for isp in range(len(sim.species)):
        keep = x_before[isp] >= sim.mins[0]
        assert sim.ctx.particle_count(isp) == int(keep.sum()) < n_before[isp]
    pl = sim.ctx  # incoming cells: ex(nx) = ex(nx+1) = plane, ey(nx) = mean of its neighbours
    ex, ey = after["ex"], after["ey"]
    i0 = ng - 1                                   # array position of x index 0
    assert np.array_equal(ex[..., i0 + nx], ex[..., i0 + nx + 1])
    assert np.allclose(ey[..., i0 + nx], 0.5 * (ey[..., i0 + nx - 1] + ey[..., i0 + nx + 1]), rtol=0, atol=0)


def test_window_runs_and_keeps_plasma(oracle_abi):
    """a window at 0.9 c over 30 steps: shifts happen, the particle count stays near the load
    (what leaves on the left is replaced on the right), energy stays finite"""
    sim = window_deck(oracle_abi, 2)
    n0 = sum(sim.ctx.particle_count(i) for i in range(2))
    for _ in range(30):
        sim.step_once()
    assert sim.window_offset > 5 * sim.d[0]
    n1 = sum(sim.ctx.particle_count(i) for i in range(2))
    assert 0.7 * n0 < n1 < 1.3 * n0
    fe, ke = sim.ctx.energy()
    assert np.isfinite(fe) and np.all(np.isfinite(ke))
