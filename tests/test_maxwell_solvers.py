"""Extended-stencil Maxwell solvers (set_maxwell_solver, fields.f90:53-179 and the
non-Yee branches of update_b_field).

The reference's own tests for this part (epoch{1,2,3}d/tests/test_maxwell_solvers.py)
inject a short laser pulse, follow the centroid of Ey^2 through the dumps and assert
that its velocity is the analytic group velocity of the scheme (rtol 0.022 / 0.012 /
0.007 in 1D / 2D / 3D).  The same decks are run here -- with simple_laser / open
boundaries instead of cpml_laser / cpml_outflow, which are outside this path -- and
the same assertion is made: on the CPU oracle in 1D and 2D (fast), on the CUDA library
in 3D (test_gpu_parity.py imports the 3D case from here)."""
import numpy as np
import pytest

from epoch_b200.driver import Simulation, LaserSpec, gauss, C

femto, micron = 1e-15, 1e-6
LAMBDA = 0.5 * micron
K_L = 2 * np.pi / LAMBDA
NX = 240
DX = 24 * micron / NX


def solver_deck(abi, ndims, solver, **kw):
    """epochNd/tests/maxwell_solvers/<solver>/input.deck"""
    n = [NX, NX // 3, NX // 3][:ndims]
    prof = {1: None, 2: lambda y: gauss(y, 0, 4 * micron),
            3: lambda y, z: gauss(np.sqrt(y ** 2 + z ** 2), 0, 4 * micron)}[ndims]
    L = LaserSpec(1, intensity_w_cm2=1e15, lambda_=LAMBDA, profile=prof,
                  t_profile=lambda t: gauss(t, 8 * femto, 1.8 * femto),
                  t_end=14 * femto if ndims == 1 else None)
    # The reference decks use CPML boundaries, for which set_dt does not apply the any_open
    # limit (setup.F90:729-732): dt = 0.95 * MIN(dx, dy, dz) / c for the extended solvers.  With
    # the simple_laser / open boundaries used here that time step is passed explicitly so that
    # the scheme under test is the one of the reference's run.
    if solver != "yee" and "dt" not in kw:
        kw["dt"] = 0.95 * DX / C
    return Simulation(ndims, n, [-12 * micron] * ndims, [12 * micron] * ndims,
                      bc_field=["simple_laser", "open"] + ["periodic"] * (2 * ndims - 2), lasers=[L],
                      t_end=75 * femto, maxwell_solver=solver, abi=abi, **kw)


def group_velocity_expected(sim, solver):
    """the formulas of the reference tests (test_maxwell_solvers.py:53-55; 3D :62-72)"""
    s = C * sim.dt / DX
    if solver.startswith("lehe"):
        return C * (1.0 + 2.0 * (1.0 - s) * (K_L * DX / 2.0) ** 2)
    return C * np.cos(K_L * DX / 2.0) / np.sqrt(1 - (s * np.sin(K_L * DX / 2.0)) ** 2)


def exact_group_velocity(sim):
    """d omega / d k at the laser wavenumber for a wave along x (k_y = k_z = 0: the beta and
    gamma terms fold into alpha, which leaves 1 - 3 delta_x)"""
    s, dl = C * sim.dt / DX, sim.coeffs[12]

    def w(k):
        x = k * DX / 2
        return 2 / sim.dt * np.arcsin(np.sqrt(s * s * np.sin(x) * ((1 - 3 * dl) * np.sin(x) + dl * np.sin(3 * x))))
    h = K_L * 1e-5
    return (w(K_L + h) - w(K_L - h)) / (2 * h)


def centroid_velocity(sim, dt_snapshot):
    """xt() + polyfit of the reference test: centroid of Ey^2 at every dump after the first"""
    pts = []
    prev = [0.0]
    x = sim.grid_min[0] + np.arange(sim.n_global[0]) * sim.d[0]

    def grab(s):
        e = s.ctx.interior(s.ctx.download_fields(["ey"])["ey"])       # [.., y, x]
        w = (e ** 2).reshape(-1, e.shape[-1]).sum(axis=0)
        if w.sum() > 0 and s.time > 0:
            pts.append((s.time, float((x * w).sum() / w.sum())))

    def out(s):
        fire = False
        if s.time >= prev[0] + dt_snapshot:
            while prev[0] + dt_snapshot <= s.time:
                prev[0] += dt_snapshot
            fire = True
        if fire or s.time >= s.t_end:
            grab(s)
    sim.start()
    while not sim.finished():
        sim.step_once(out)
    grab(sim)            # output_routines after the loop (epoch3d.F90:261)
    pts = np.array(pts)
    return np.polyfit(pts[:, 0], pts[:, 1], 1)[0], pts


def test_dt_and_coefficients(oracle_abi):
    """set_dt / set_maxwell_solver scalars against hand evaluation of the reference formulas"""
    s = solver_deck(oracle_abi, 1, "lehe_x", dt=None)
    assert np.isclose(s.dt, 0.95 * DX / C, rtol=1e-14)
    dx_cdt = 1 / 0.95
    dl = 0.25 * (1 - dx_cdt ** 2 * np.sin(0.5 * np.pi / dx_cdt) ** 2)
    assert np.isclose(s.coeffs[12], dl, rtol=1e-12) and np.isclose(s.coeffs[0], 1 - 3 * dl, rtol=1e-12)
    s = solver_deck(oracle_abi, 3, "cowan", dt=None)
    dy = 3 * DX
    assert np.isclose(s.dt, 0.95 * DX / C, rtol=1e-14)
    c1, c2 = 1.0, (DX / dy) ** 2
    cx1 = 1 / (c1 * c2 + c2 * c2 + c1 * c2)
    cx2 = 1 - c1 * c2 * c2 * cx1
    assert np.isclose(s.coeffs[5], 0.125 * c1 * cx2, rtol=1e-12)            # betayx
    assert np.isclose(s.coeffs[9], c2 * c2 * (0.0625 - 0.125 * c2 * c2 * cx1), rtol=1e-12)   # gammax
    assert np.isclose(s.coeffs[0], 1 - 2 * s.coeffs[3] - 2 * s.coeffs[4] - 4 * s.coeffs[9], rtol=1e-14)
    s = solver_deck(oracle_abi, 3, "lehe_x", dt=None)
    # any_open: MIN(lehe dt, 3D Yee CFL) (setup.F90:713-715,729-732)
    yee = DX * dy * dy / np.sqrt(2 * (DX * dy) ** 2 + dy ** 4) / C
    assert np.isclose(s.dt, 0.95 * min(DX / C, yee), rtol=1e-14)
    assert np.isclose(s.coeffs[3], 0.125 * (DX / dy) ** 2, rtol=1e-12) and s.coeffs[5] == 0.125 and s.coeffs[6] == 0.0


@pytest.mark.parametrize("ndims", [1, 2, 3])
def test_custom_stencil_without_terms_is_yee(oracle_abi, ndims):
    """alpha = 1, beta = gamma = delta = 0 must reproduce the Yee update exactly"""
    n = [20, 12, 10][:ndims]
    out = []
    for solver, kw in (("yee", {}), ("custom", {"stencil": {"dt": None}})):
        sim = Simulation(ndims, n, [0.0] * ndims, [1e-6 * k for k in n], abi=oracle_abi, dt=1e-15 / 3,
                         maxwell_solver=solver, **kw)
        rng = np.random.default_rng(4)
        shp = sim.ctx._np_shape()
        sim.ctx.upload_fields(**{k: rng.normal(0, 1, shp) for k in ("ex", "ey", "ez", "bx", "by", "bz")})
        sim.ctx.set_time(sim.dt, 0.5 * sim.dt, 0)
        sim.ctx.update_eb_fields_half()
        sim.ctx.update_eb_fields_final()
        out.append(sim.ctx.download_fields(["bx", "by", "bz", "ex", "ey", "ez"]))
    for k in out[0]:
        assert np.array_equal(out[0][k], out[1][k]), k


@pytest.mark.parametrize("ndims,solver,rtol,printed", [
    (1, "lehe_x", 0.022, 313164075.045), (1, "yee", 0.022, None),
    (2, "lehe_x", 0.012, 312016227.758), (2, "pukhov", 0.012, 292013249.255)])
def test_group_velocity_oracle(oracle_abi, ndims, solver, rtol, printed):
    sim = solver_deck(oracle_abi, ndims, solver)
    vg, pts = centroid_velocity(sim, 12 * femto if ndims == 1 else 25 * femto)
    assert len(pts) >= 3
    # the reference's assertion ...
    assert np.isclose(vg, group_velocity_expected(sim, solver), rtol=rtol), (vg, group_velocity_expected(sim, solver))
    # ... the value its authors recorded in the test source, to the same tolerance (CPML boundaries
    # there, open boundaries here; yee is commented out of the reference's solver list, so no
    # recorded value is used for it) ...
    if printed is not None:
        assert abs(vg / printed - 1) < rtol, (vg, printed)
    # ... and, much tighter, the group velocity of the scheme's exact dispersion relation on the
    # x axis: sin^2(w dt/2) = s^2 sin(kdx/2) [(1 - 3 delta) sin(kdx/2) + delta sin(3 kdx/2)]
    # (the formula of the reference test is its small-k expansion)
    assert abs(vg / exact_group_velocity(sim) - 1) < 3e-3, (vg, exact_group_velocity(sim))
