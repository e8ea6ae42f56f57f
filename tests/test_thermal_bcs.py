"""Thermal particle boundaries (row a11: particle_bcs, boundary.F90:1405-1461 and the five other faces;
flux_momentum_from_temperature / momentum_from_temperature, particle_temperature.F90:396-469).

A particle beyond the OUTER edge of a thermal face is re-emitted: position mirrored at the outer edge (deterministic:
compared exactly), momenta drawn from the wall temperature interpolated at its transverse position (random: the
reference's generator is one sequential host stream, so draws depend on the list order; compared through the
moments of the deviates normalised by an independent numpy restatement of the interpolation).
  not gpu: the oracle (KISS + polar Box-Muller as the reference) against the numpy restatement
  gpu:     the CUDA library against the oracle (exact part) and against the same moments
heat_bath is accepted as an open boundary (setup_domain_dependent_boundaries, boundary.F90:75-93)."""
import math

import numpy as np
import pytest

from helpers import Simulation, SpeciesSpec, Q0, M0

KB = 1.3806488e-23
NG = 5            # triangle: png + 2


def wall_tables(sim, face, t0):
    """ext_temp_<face>: (3, e2, e1) with a different smooth profile per component"""
    nd = sim.ndims
    d = (face - 1) // 2
    ta = [a for a in range(nd) if a != d]
    ext = [sim.n_local[a] + 2 * NG for a in ta]
    if not ta:
        return np.array([t0, 2.0 * t0, 0.5 * t0])
    j1 = np.arange(ext[0])[None, :]
    j2 = np.arange(ext[1])[:, None] if len(ta) == 2 else np.zeros((1, 1))
    out = np.empty((3, ext[1] if len(ta) == 2 else 1, ext[0]))
    for c in range(3):
        out[c] = t0 * (1.0 + 0.5 * c) * (1.0 + 0.3 * np.sin(0.4 * j1 + c) * np.cos(0.3 * j2))
    return out if len(ta) == 2 else out[:, 0, :]


def interp_temperature(sim, face, table, P):
    """numpy restatement of the triangle-weighted wall temperature at the particles' transverse positions"""
    nd = sim.ndims
    d = (face - 1) // 2
    ta = [a for a in range(nd) if a != d]
    n = len(P)
    if not ta:
        return np.tile(table, (n, 1))
    cells, gs = [], []
    for a in ta:
        cell_r = (P[:, a] - sim.grid_min_local[a]) / sim.d[a]
        c0 = np.floor(cell_r + 0.5)
        cf = c0 - cell_r
        cells.append(c0.astype(np.int64) + 1)
        gs.append(np.stack([0.5 * (0.25 + cf * cf + cf), 0.75 - cf * cf, 0.5 * (0.25 + cf * cf - cf)], axis=1))
    temp = np.zeros((n, 3))
    tab = table if len(ta) == 2 else table[:, None, :]
    for c in range(3):
        for i2 in ((-1, 0, 1) if len(ta) == 2 else (0,)):
            for i1 in (-1, 0, 1):
                j1 = cells[0] + i1 - (1 - NG)
                j2 = cells[1] + i2 - (1 - NG) if len(ta) == 2 else np.zeros(n, dtype=np.int64)
                w2 = gs[1][:, i2 + 1] if len(ta) == 2 else 1.0
                temp[:, c] += gs[0][:, i1 + 1] * w2 * tab[c, j2, j1]
    return temp


def build(abi, nd, faces, seed=5, n_out=4000, pbc_other="periodic"):
    """a box with thermal walls on `faces` for species 0, reflecting walls for species 1; n_out particles of each
    species are placed beyond each outer edge (moving outwards), the rest inside"""
    n = [24, 16, 12][:nd]
    cell = 1.0e-6
    bcs0 = ["periodic"] * (2 * nd)
    bcf = ["periodic"] * (2 * nd)
    for f in faces:
        d = (f - 1) // 2
        bcs0[2 * d] = bcs0[2 * d + 1] = "open"
        bcf[2 * d] = bcf[2 * d + 1] = "clamp"
    for f in faces:
        bcs0[f - 1] = "thermal"
    bcs1 = [("reflect" if b != "periodic" else b) for b in bcs0]
    species = [SpeciesSpec("electron", -Q0, M0, bcs=bcs0), SpeciesSpec("proton", Q0, 1836.2 * M0, bcs=bcs1)]
    sim = Simulation(nd, n, [0.0] * nd, [cell * x for x in n], shape="triangle", species=species, abi=abi, bc_field=bcf)
    rng = np.random.default_rng(seed)
    tables = {}
    for f in faces:
        tables[f] = wall_tables(sim, f, 2.0e6)
        sim.ctx.set_thermal_boundary(0, f, tables[f], seed=1234 + f)
    # boundary_shift = dx * REAL((1 + png + cpml_thickness) / 2) = 2 dx; x_min_outer = x_min - boundary_shift (boundary.F90:1336-1338)
    sim.outer = {}
    for d in range(nd):
        bs = sim.d[d] * float((1 + 3 + 0) // 2)
        sim.outer[2 * d + 1], sim.outer[2 * d + 2] = sim.mins[d] - bs, sim.maxs[d] + bs
    shift = cell * 2.0
    loads = []
    for isp, sp in enumerate(species):
        rows = []
        m_in = 3000
        Pin = np.empty((m_in, nd + 4))
        for a in range(nd):
            Pin[:, a] = rng.uniform(sim.min_local[a], sim.max_local[a], m_in)
        Pin[:, nd:nd + 3] = rng.normal(0.0, 0.1, (m_in, 3)) * sp.mass * 2.99792458e8
        rows.append(Pin)
        for f in faces:
            d, side = (f - 1) // 2, (f - 1) % 2
            Q = np.empty((n_out, nd + 4))
            for a in range(nd):
                Q[:, a] = rng.uniform(sim.min_local[a], sim.max_local[a], n_out)
            edge = sim.outer[f]
            depth = rng.uniform(0.05, 0.9, n_out) * cell
            if isp == 1:       # the reflecting species stays inside the outer edge (beyond it a reflecting wall removes, :1404-1461)
                edge = sim.mins[d] if side == 0 else sim.maxs[d]
            Q[:, d] = edge - depth if side == 0 else edge + depth
            Q[:, nd:nd + 3] = rng.normal(0.0, 0.1, (n_out, 3)) * sp.mass * 2.99792458e8
            rows.append(Q)
        P = np.concatenate(rows)
        P[:, nd + 3] = 1.0 + np.arange(len(P)) * 1e-6 + isp       # unique weights: the particle's identity
        sim.ctx.upload_particles(isp, P)
        loads.append(P)
    sim.loads, sim.tables, sim.shift = loads, tables, shift
    sim.ctx.set_time(sim.dt, 0.0, 0)
    return sim


def by_weight(P):
    return P[np.argsort(P[:, -1], kind="stable")]


def check_reemission(sim, faces, got, tol_sigma=5.0):
    """positions mirrored exactly, untouched particles untouched, moments of the normalised deviates"""
    nd = sim.ndims
    P0 = by_weight(sim.loads[0])
    P1 = by_weight(got)
    assert len(P1) == len(P0)                      # a thermal wall never removes a particle
    touched = np.zeros(len(P0), dtype=bool)
    for f in faces:
        d, side = (f - 1) // 2, (f - 1) % 2
        edge = sim.outer[f]
        sel = (P0[:, d] < edge) if side == 0 else (P0[:, d] >= edge)
        assert sel.sum() > 1000
        touched |= sel
        assert np.array_equal(P1[sel, d], 2.0 * edge - P0[sel, d])
        for a in range(nd):
            if a != d:
                assert np.array_equal(P1[sel, a], P0[sel, a])
        temp = interp_temperature(sim, f, sim.tables[f], P0[sel])
        sig = np.sqrt(temp * KB * M0)
        u = P1[sel, nd:nd + 3] / sig
        m = sel.sum()
        # normal component: Rayleigh, heading inwards; E z = sqrt(pi/2), E z^2 = 2, var z = 2 - pi/2
        z = u[:, d] * (1.0 if side == 0 else -1.0)
        assert np.all(z > 0.0)
        assert abs(z.mean() - math.sqrt(math.pi / 2)) < tol_sigma * math.sqrt((2 - math.pi / 2) / m)
        assert abs((z * z).mean() - 2.0) < tol_sigma * 2.0 / math.sqrt(m)          # var(z^2) = 4
        for c in range(3):
            if c == d:
                continue
            assert abs(u[:, c].mean()) < tol_sigma / math.sqrt(m)
            assert abs((u[:, c] ** 2).mean() - 1.0) < tol_sigma * math.sqrt(2.0 / m)
            assert abs((u[:, c] ** 4).mean() - 3.0) < tol_sigma * math.sqrt(96.0 / m)
    assert np.array_equal(P1[~touched], P0[~touched])
    return touched


CASES = [(1, (1, 2)), (2, (1, 2)), (2, (3,)), (3, (1, 6)), (3, (4,))]


@pytest.mark.parametrize("nd,faces", CASES)
def test_oracle_thermal_reemission(oracle_abi, nd, faces):
    oracle_abi.lib.orc_random_init(4711)
    sim = build(oracle_abi, nd, faces)
    sim.ctx.particle_bcs()
    check_reemission(sim, faces, sim.ctx.download_particles(0))
    # the reflecting species beside it: untouched by the thermal code, mirrored at the boundary itself
    P0, P1 = by_weight(sim.loads[1]), by_weight(sim.ctx.download_particles(1))
    assert len(P0) == len(P1)
    for f in faces:
        d, side = (f - 1) // 2, (f - 1) % 2
        wall = sim.mins[d] if side == 0 else sim.maxs[d]
        sel = (P0[:, d] < sim.min_local[d]) if side == 0 else (P0[:, d] >= wall)
        assert np.array_equal(P1[sel, d], 2.0 * wall - P0[sel, d])


def test_oracle_thermal_needs_table(oracle_abi):
    from epoch_b200 import EpochError
    sp = SpeciesSpec("electron", -Q0, M0, bcs=["thermal", "thermal"])
    sim = Simulation(1, [16], [0.0], [16e-6], species=[sp], abi=oracle_abi, bc_field=["clamp", "clamp"])
    sim.ctx.upload_particles(0, np.array([[4e-6, 0.0, 0.0, 0.0, 1.0]]))
    sim.ctx.set_time(sim.dt, 0.0, 0)
    with pytest.raises(EpochError, match="thermal"):
        sim.ctx.particle_bcs()


def test_heat_bath_is_open(oracle_abi):
    """heat_bath = open boundary (+ a host-side injector): particles beyond the outer edge are removed"""
    sp = SpeciesSpec("electron", -Q0, M0, bcs=["heat_bath", "heat_bath"])
    sim = Simulation(1, [16], [0.0], [16e-6], species=[sp], abi=oracle_abi, bc_field=["clamp", "clamp"])
    P = np.array([[4e-6, 0.0, 0.0, 0.0, 1.0], [-2.5e-6, 0.0, 0.0, 0.0, 2.0], [18.5e-6, 0.0, 0.0, 0.0, 3.0]])
    sim.ctx.upload_particles(0, P)
    sim.ctx.set_time(sim.dt, 0.0, 0)
    sim.ctx.particle_bcs()
    assert sim.ctx.particle_count(0) == 1


@pytest.mark.gpu
@pytest.mark.parametrize("nd,faces", CASES)
def test_cuda_thermal_reemission(product_abi, oracle_abi, nd, faces):
    oracle_abi.lib.orc_random_init(4711)
    sims = [build(abi, nd, faces) for abi in (product_abi, oracle_abi)]
    for s in sims:
        s.ctx.particle_bcs()
    got = [by_weight(s.ctx.download_particles(0)) for s in sims]
    touched = check_reemission(sims[0], faces, got[0])
    # against the oracle: the deterministic part bit for bit ...
    assert np.array_equal(got[0][:, :nd], got[1][:, :nd])
    assert np.array_equal(got[0][~touched], got[1][~touched])
    # ... the random part through its first two moments per component
    for c in range(3):
        a, b = got[0][touched, nd + c], got[1][touched, nd + c]
        s = b.std()
        assert abs(a.mean() - b.mean()) < 6.0 * s * math.sqrt(2.0 / len(a))
        assert abs(a.std() / s - 1.0) < 6.0 / math.sqrt(len(a))
    # the reflecting species equals the oracle's exactly
    assert np.array_equal(by_weight(sims[0].ctx.download_particles(1)), by_weight(sims[1].ctx.download_particles(1)))


@pytest.mark.gpu
@pytest.mark.parametrize("nd", [2, 3])
def test_cuda_thermal_walls_steps(product_abi, nd):
    """a hot plasma between thermal walls in x for 12 full steps (tiled kernels, deferred permute): nobody is lost,
    nobody ends beyond the outer edge, the walls re-emit at their own temperature"""
    n = [16, 16, 8][:nd]
    cell = 1.0e-6
    bcs = ["thermal", "thermal"] + ["periodic"] * (2 * nd - 2)
    bcf = ["clamp", "clamp"] + ["periodic"] * (2 * nd - 2)
    sp = SpeciesSpec("electron", -Q0, M0, bcs=bcs, density=1e24, temperature=(1e9,) * 3)
    sim = Simulation(nd, n, [0.0] * nd, [cell * x for x in n], shape="triangle", species=[sp], abi=product_abi, bc_field=bcf)
    for f in (1, 2):
        shape = (3,) + tuple(n[a] + 2 * NG for a in range(nd - 1, 0, -1))
        sim.ctx.set_thermal_boundary(0, f, np.full(shape, 1.0e7), seed=99)
    rng = np.random.default_rng(3)
    m = 20 * int(np.prod(n))
    P = np.empty((m, nd + 4))
    for a in range(nd):
        P[:, a] = rng.uniform(sim.min_local[a], sim.max_local[a], m)
    P[:, nd:nd + 3] = rng.normal(0.0, 1.0, (m, 3)) * math.sqrt(KB * 1e9 * M0)
    P[:, nd + 3] = 1e24 * cell ** nd / 20
    sim.ctx.upload_particles(0, P)
    for _ in range(12):
        sim.step_once()
    got = sim.ctx.download_particles(0)
    assert len(got) == m
    assert got[:, 0].min() >= sim.min_local[0] - 2.0 * cell and got[:, 0].max() <= sim.max_local[0] + 2.0 * cell
    assert int(sim.ctx.timer("cell_mismatch")[1]) == 0
    # (v_th = 0.41 c: in 12 steps only the particles that start within ~1.3 cells of a wall get 2 cells beyond it)
    assert int(sim.ctx.timer("thermal_reemitted")[1]) > m // 200
