"""Derived grid variables of io/calc_df.F90: oracle known answers and an
independent numpy restatement of the particle-to-grid weighting; the CUDA parity test is in
test_gpu_parity.py (it imports the cases from here)."""
import numpy as np
import pytest

from helpers import make_plasma, SHAPE_NAMES
from epoch_b200.driver import C, Q0, M0

KB = 1.3806488e-23
MU0 = 4e-7 * np.pi

# (kind, species, direction) combinations that cover every routine and every argument form
CASES = [("mass_density", -1, 0), ("mass_density", 1, 0), ("ekbar", 0, 0), ("ekbar", -1, 0),
         ("ekflux", 0, 1), ("ekflux", 0, -2), ("ekflux", -1, 3), ("poynt_flux", -1, 1), ("poynt_flux", -1, 2),
         ("poynt_flux", -1, 3), ("charge_density", -1, 0), ("charge_density", 0, 0), ("number_density", 0, 0),
         ("number_density", -1, 0), ("ppc", 0, 0), ("ppc", -1, 0), ("average_weight", 1, 0),
         ("temperature", 0, 0), ("temperature", 0, 1), ("temperature", -1, 3), ("species_current", 0, 1),
         ("species_current", -1, 2), ("species_current", 1, 3), ("average_momentum", 0, 2),
         ("average_momentum", -1, 1)]


def gx_weights(shape, frac):
    """include/<shape>/gxfac.inc for an array of cell_frac values -> [n, nw]"""
    f = frac
    if shape == "tophat":
        return np.stack([0.5 + f, 0.5 - f], 1)
    if shape == "triangle":
        return np.stack([0.5 * (0.25 + f * f + f), 0.75 - f * f, 0.5 * (0.25 + f * f - f)], 1)
    c2 = f * f
    return np.stack([(0.5 + f) ** 4 / 24, (1.1875 + 2.75 * f + c2 * (1.5 - f - c2)) / 6,
                     0.25 * (115.0 / 48.0 + c2 * (c2 - 2.5)), (1.1875 - 2.75 * f + c2 * (1.5 + f - c2)) / 6,
                     (0.5 - f) ** 4 / 24], 1)


def numpy_deposit(sim, P, wdata):
    """sum of gx*gy*gz*wdata onto a periodic grid (interior only), written from
    particle_to_grid.inc with numpy: the independent check of the oracle's weighting"""
    nd, shape = sim.ndims, sim.shape_name
    sf_min = {"tophat": 0, "triangle": -1, "bspline3": -2}[shape]
    n = sim.n_local
    out = np.zeros(n[::-1])
    cells, ws = [], []
    for d in range(nd):
        cr = (P[:, d] - sim.grid_min_local[d]) / sim.d[d] - (0.5 if shape == "tophat" else 0.0)
        cell = np.floor(cr + 0.5)
        ws.append(gx_weights(shape, cell - cr))
        cells.append(cell.astype(int) + 1)
    nw = ws[0].shape[1]
    for off in np.ndindex(*([nw] * nd)):
        g = np.ones(len(P))
        idx = []
        for d in range(nd):
            g = g * ws[d][:, off[d]]
            idx.append((cells[d] + sf_min + off[d] - 1) % n[d])       # periodic fold of the ghosts
        np.add.at(out, tuple(idx[::-1]), g * wdata)
    return out


@pytest.mark.parametrize("ndims", [1, 2, 3])
@pytest.mark.parametrize("shape", SHAPE_NAMES)
def test_densities_against_numpy(oracle_abi, ndims, shape):
    sim = make_plasma(oracle_abi, ndims, shape, [12, 8, 6], seed=3, ppc=3, u_th=0.5)
    sim.shape_name = shape
    vol = float(np.prod(sim.d))
    P0, P1 = sim.loads
    want_n = numpy_deposit(sim, P0, P0[:, -1]) / vol
    got_n = sim.ctx.interior(sim.ctx.calc_derived("number_density", 0))
    assert np.max(np.abs(got_n - want_n)) < 1e-12 * np.max(want_n)
    want_q = (numpy_deposit(sim, P0, -Q0 * P0[:, -1]) + numpy_deposit(sim, P1, Q0 * P1[:, -1])) / vol
    got_q = sim.ctx.interior(sim.ctx.calc_derived("charge_density", -1))
    assert np.max(np.abs(got_q - want_q)) < 1e-12 * np.max(np.abs(want_n)) * Q0
    # total number is conserved by the weighting
    assert np.isclose(got_n.sum() * vol, P0[:, -1].sum(), rtol=1e-12)
    pmc = C * M0
    root = 1.0 / np.sqrt(pmc ** 2 + (P0[:, ndims:ndims + 3] ** 2).sum(1))
    want_j = numpy_deposit(sim, P0, -Q0 * P0[:, -1] * P0[:, ndims + 1] * root) * C / vol
    got_j = sim.ctx.interior(sim.ctx.calc_derived("species_current", 0, 2))
    assert np.max(np.abs(got_j - want_j)) < 1e-12 * np.max(np.abs(want_j))


def test_known_answers(oracle_abi):
    """a cold drifting beam: ekbar = (gamma-1) m c^2, temperature = 0, average momentum = p,
    ppc counts every particle once; uniform fields: Poynting flux = E x B / mu0"""
    sim = make_plasma(oracle_abi, 2, "triangle", [10, 8], seed=4, ppc=4, two_species=False)
    P = sim.loads[0].copy()
    p0 = np.array([0.7, -0.2, 0.4]) * M0 * C
    P[:, 2:5] = p0
    sim.ctx.upload_particles(0, P)
    occupied = sim.ctx.interior(sim.ctx.calc_derived("number_density", 0)) > 0
    gamma = np.sqrt(1 + (p0 ** 2).sum() / (M0 * C) ** 2)
    ek = sim.ctx.interior(sim.ctx.calc_derived("ekbar", 0))
    assert np.allclose(ek[occupied], (gamma - 1) * M0 * C * C, rtol=1e-12)
    T = sim.ctx.interior(sim.ctx.calc_derived("temperature", 0))
    assert np.max(np.abs(T)) < 1e-12 * (p0 ** 2).sum() / M0 / KB
    for d in (1, 2, 3):
        am = sim.ctx.interior(sim.ctx.calc_derived("average_momentum", 0, d))
        assert np.allclose(am[occupied], p0[d - 1], rtol=1e-12)
    assert sim.ctx.interior(sim.ctx.calc_derived("ppc", 0)).sum() == len(P)
    aw = sim.ctx.interior(sim.ctx.calc_derived("average_weight", 0))
    cnt = sim.ctx.interior(sim.ctx.calc_derived("ppc", 0))
    assert np.isclose((aw * cnt).sum(), P[:, -1].sum(), rtol=1e-12)
    # forward flux along x only (p_x > 0), backward along y only
    assert sim.ctx.calc_derived("ekflux", 0, -1).max() == 0 and sim.ctx.calc_derived("ekflux", 0, 1).max() > 0
    assert sim.ctx.calc_derived("ekflux", 0, 2).max() == 0 and sim.ctx.calc_derived("ekflux", 0, -2).max() > 0
    shp = sim.ctx._np_shape()
    f = {k: np.zeros(shp) for k in ("ex", "ey", "ez", "bx", "by", "bz")}
    f["ey"][:] = 2.0e9
    f["bz"][:] = 3.0
    f["ez"][:] = -1.0e9
    f["bx"][:] = 0.5
    sim.ctx.upload_fields(**f)
    assert np.allclose(sim.ctx.interior(sim.ctx.calc_derived("poynt_flux", -1, 1)), 2.0e9 * 3.0 / MU0, rtol=1e-14)
    assert np.allclose(sim.ctx.interior(sim.ctx.calc_derived("poynt_flux", -1, 2)), -1.0e9 * 0.5 / MU0, rtol=1e-14)
    assert np.allclose(sim.ctx.interior(sim.ctx.calc_derived("poynt_flux", -1, 3)), -2.0e9 * 0.5 / MU0, rtol=1e-14)


def test_temperature_of_a_maxwellian(oracle_abi):
    """<(p - <p>)^2> / (m kB) per degree of freedom recovers the loaded temperature (statistical)"""
    sim = make_plasma(oracle_abi, 1, "triangle", [16], seed=6, ppc=4000, two_species=False)
    P = sim.loads[0].copy()
    T0 = 3.0e6
    rng = np.random.default_rng(1)
    P[:, 1:4] = rng.normal(0.0, np.sqrt(M0 * KB * T0), (len(P), 3)) + np.array([0.3, 0, 0]) * M0 * C
    sim.ctx.upload_particles(0, P)
    for d, tol in ((0, 0.03), (1, 0.05), (2, 0.05), (3, 0.05)):
        T = sim.ctx.interior(sim.ctx.calc_derived("temperature", 0, d))
        assert np.all(np.abs(T / T0 - 1) < tol), (d, T / T0)


def test_argument_errors(oracle_abi):
    sim = make_plasma(oracle_abi, 1, "triangle", [8], ppc=1)
    for bad in (("ekflux", 0, 0), ("species_current", 0, 0), ("mass_density", 0, -1), ("mass_density", 7, 0),
                ("poynt_flux", -1, 4)):
        with pytest.raises(Exception):
            sim.ctx.calc_derived(*bad)
