"""-m gpu: parity of the CUDA path (through the C ABI) against the CPU oracle on
identical seeded decks, against the reference's golden sums, and through
size-independent invariants.  Tolerances: fp64, 1e-12 relative (positions relative
to the domain length, grid quantities norm-wise), as BASELINE.json's north_star
states; golden sums at the reference's own np.isclose tolerance."""
import numpy as np
import pytest

from helpers import make_plasma, SHAPE_NAMES, ALL9, FIELDS6, match_particles, relerr
from test_oracle import (laser_1d, laser_2d, laser_3d, run_laser_deck, check_golden, GOLD_1D,
                         GOLD_2D, GOLD_3D, femto)
from epoch_b200.driver import Simulation, SpeciesSpec, Q0, M0, C

pytestmark = pytest.mark.gpu
TOL = 1e-12


def compare_fields(a, b, names, tol=TOL):
    fa, fb = a.ctx.download_fields(names), b.ctx.download_fields(names)
    for k in names:
        assert relerr(fa[k], fb[k]) < tol, (k, relerr(fa[k], fb[k]))


def compare_particles(gpu, orc, tol=TOL):
    for isp in range(len(gpu.species)):
        A, B = gpu.ctx.download_particles(isp), orc.ctx.download_particles(isp)
        assert A.shape == B.shape, (A.shape, B.shape)
        if len(A) == 0:
            continue
        A, B = match_particles(A, B)
        nd = gpu.ndims
        assert np.array_equal(A[:, -1], B[:, -1])
        for d in range(nd):
            assert np.max(np.abs(A[:, d] - B[:, d])) / gpu.length[d] < tol, ("pos", d)
        pscale = np.max(np.abs(B[:, nd:nd + 3]))
        assert np.max(np.abs(A[:, nd:nd + 3] - B[:, nd:nd + 3])) / pscale < tol, "momentum"


@pytest.mark.parametrize("ndims", [1, 2, 3])
@pytest.mark.parametrize("order", [2, 4, 6])
def test_field_half_steps(product_abi, oracle_abi, ndims, order):
    n = [40, 24, 20]
    sims = [make_plasma(a, ndims, "triangle", n, seed=5, ppc=1, field_order=order, two_species=False)
            for a in (product_abi, oracle_abi)]
    for s in sims:
        rng = np.random.default_rng(9)
        shp = s.ctx._np_shape()
        s.ctx.upload_fields(**{k: rng.normal(0, 1e6, shp) for k in ("jx", "jy", "jz")})
        s.ctx.set_time(s.dt, 0.5 * s.dt, 0)
        s.ctx.update_eb_fields_half()
        s.ctx.update_eb_fields_final()
    compare_fields(sims[0], sims[1], FIELDS6)


@pytest.mark.parametrize("ndims", [1, 2, 3])
@pytest.mark.parametrize("shape", SHAPE_NAMES)
def test_push_deposit_one_step(product_abi, oracle_abi, ndims, shape):
    """push_particles + current_finish on random relativistic particles in random
    fields (every particle crosses cells with probability ~1/2 per axis)."""
    n = [24, 12, 10]
    sims = [make_plasma(a, ndims, shape, n, seed=11, ppc=5, u_th=1.0) for a in (product_abi, oracle_abi)]
    for s in sims:
        s.ctx.set_time(s.dt, 0.5 * s.dt, 0)
        s.ctx.push_particles()
        s.ctx.current_finish()
    compare_particles(*sims)
    compare_fields(sims[0], sims[1], ("jx", "jy", "jz"))


@pytest.mark.parametrize("ndims,shape", [(1, "triangle"), (2, "triangle"), (3, "triangle"), (3, "tophat"), (2, "bspline3")])
def test_full_steps_energy_and_charge(product_abi, oracle_abi, ndims, shape):
    """10 full PIC steps: fields, particles, energy history and the Gauss-law
    invariant agree with the oracle; charge is conserved to rounding."""
    n = [32, 16, 12]
    sims = [make_plasma(a, ndims, shape, n, seed=21, ppc=4, u_th=0.2, e_amp=1e9, b_amp=3.0)
            for a in (product_abi, oracle_abi)]
    hist = [[], []]
    for k, s in enumerate(sims):
        def hook(sim, k=k):
            if sim.step == 0:
                sim.ctx.set_option("capture_charge_reference", "1")
            hist[k].append(sim.ctx.energy() + sim.ctx.charge_residual())
        for _ in range(10):
            s.step_once(after_half=hook)
    compare_fields(sims[0], sims[1], ALL9, tol=1e-10)
    compare_particles(sims[0], sims[1], tol=1e-10)
    for (fe_g, pe_g, res_g, rho_g), (fe_o, pe_o, res_o, rho_o) in zip(*hist):
        assert abs(fe_g - fe_o) <= 1e-11 * abs(fe_o)
        assert np.allclose(pe_g, pe_o, rtol=1e-11, atol=0)
        assert res_g <= 5e-12 * rho_g      # charge conservation at machine precision


@pytest.mark.parametrize("deck,comp,snap,gold", [(laser_1d, "ey", 8 * femto, GOLD_1D),
                                               (laser_2d, "ey", 25 * femto, GOLD_2D),
                                               (laser_3d, "ex", 25 * femto, GOLD_3D)])
def test_laser_goldens_on_gpu(product_abi, deck, comp, snap, gold):
    """The reference's own test_laser assertions, run through the CUDA path."""
    check_golden(run_laser_deck(deck(product_abi), comp, snap), gold)


@pytest.mark.parametrize("ndims", [1, 2, 3])
@pytest.mark.parametrize("pbc,fbc", [("reflect", "conduct"), ("open", "clamp"), ("open", "zero_gradient"), ("reflect", "open")])
def test_boundaries(product_abi, oracle_abi, ndims, pbc, fbc):
    """Non-periodic x (and, in 3D, z) boundaries: reflecting and open particles,
    conducting / clamped / zero-gradient / outflow fields; other axes periodic."""
    n = [20, 12, 10]
    bc_field = [fbc, fbc] + ["periodic"] * (2 * (ndims - 1))
    bcs = [pbc, pbc] + ["periodic"] * (2 * (ndims - 1))
    if ndims == 3:
        bc_field[4:6] = [fbc, fbc]
        bcs[4:6] = [pbc, pbc]
    sims = [make_plasma(a, ndims, "triangle", n, seed=31, ppc=4, u_th=1.0, bc_field=bc_field, bcs=bcs)
            for a in (product_abi, oracle_abi)]
    for s in sims:
        for _ in range(6):
            s.step_once()
    assert sims[0].ctx.particle_count(0) == sims[1].ctx.particle_count(0)
    if pbc == "open":
        assert sims[0].ctx.particle_count(0) < len(sims[0].loads[0])   # some particles really left
    compare_particles(sims[0], sims[1], tol=1e-10)
    compare_fields(sims[0], sims[1], ALL9, tol=1e-10)


@pytest.mark.parametrize("shape", SHAPE_NAMES)
def test_cell_edge_particles(product_abi, oracle_abi, shape):
    """Particles exactly on k*dx and (k+1/2)*dx with zero momentum: FLOOR(x+0.5)
    / FLOOR(x) conventions."""
    sims = []
    for a in (product_abi, oracle_abi):
        sp = SpeciesSpec("electron", -Q0, M0)
        s = Simulation(2, [16, 16], [0.0, 0.0], [16e-6, 16e-6], shape=shape, species=[sp], abi=a)
        k = np.arange(32)
        P = np.zeros((64, 6))
        P[:32, 0] = 0.5e-6 * k
        P[:32, 1] = 0.5e-6 * ((k * 7) % 32)
        P[32:, 0] = 0.5e-6 * ((k * 5) % 32)
        P[32:, 1] = 0.5e-6 * k
        P[:, 0:2] = np.minimum(P[:, 0:2], 16e-6 * (1 - 1e-16))
        P[32:, 2] = 1e-3 * M0 * C
        P[:, 5] = 1.0 + np.arange(64)
        rng = np.random.default_rng(2)
        shp = s.ctx._np_shape()
        s.ctx.upload_fields(**{f: rng.normal(0, 1e8 if f[0] == "e" else 0.3, shp) for f in FIELDS6})
        s.ctx.upload_particles(0, P)
        s.ctx.set_time(s.dt, 0.5 * s.dt, 0)
        s.ctx.push_particles()
        s.ctx.current_finish()
        sims.append(s)
    compare_particles(*sims)
    # J of these slow particles (displacement ~7e-4 cell) is the difference of nearly equal
    # shape weights: rounding of the weights (FMA vs. unfused) is amplified by ~1/7e-4, so
    # 1e-11 on J here corresponds to <1e-14 on the weights themselves.
    compare_fields(sims[0], sims[1], ("jx", "jy", "jz"), tol=1e-11)


def test_empty_and_ragged(product_abi, oracle_abi):
    """No particles at all; one empty and one populated species; zero-current and
    immobile species."""
    for a in (product_abi,):
        sp = [SpeciesSpec("e", -Q0, M0), SpeciesSpec("p", Q0, 1836 * M0), SpeciesSpec("tracer", -Q0, M0, zero_current=True),
              SpeciesSpec("fixed", Q0, 1836 * M0, immobile=True)]
        s = Simulation(3, [8, 8, 8], [0.0] * 3, [8e-6] * 3, species=sp, abi=a)
        for _ in range(2):
            s.step_once()
        assert s.ctx.particle_count(0) == 0
        j = s.ctx.download_fields(("jx",))["jx"]
        assert not j.any()
    sims = []
    for a in (product_abi, oracle_abi):
        s = Simulation(3, [8, 8, 8], [0.0] * 3, [8e-6] * 3, species=sp, abi=a)
        rng = np.random.default_rng(4)
        P = np.empty((37, 7))
        P[:, 0:3] = rng.uniform(0, 8e-6, (37, 3))
        P[:, 3:6] = rng.normal(0, 0.5, (37, 3)) * M0 * C
        P[:, 6] = 1e10 * (1 + np.arange(37))
        s.ctx.upload_particles(1, P * np.array([1, 1, 1, 1836, 1836, 1836, 1]))
        s.ctx.upload_particles(2, P)
        s.ctx.upload_particles(3, P)
        for _ in range(3):
            s.step_once()
        sims.append(s)
    compare_particles(*sims)
    compare_fields(sims[0], sims[1], ALL9, tol=1e-10)
    # immobile species untouched, tracer moved but deposited nothing on its own
    fixed = sims[0].ctx.download_particles(3)
    assert np.array_equal(np.sort(fixed[:, 0]), np.sort(P[:, 0]))


def test_upload_download_roundtrip(product_abi):
    s = make_plasma(product_abi, 3, "triangle", [6, 5, 4], seed=1, ppc=7)
    for isp in range(2):
        Q = s.ctx.download_particles(isp)
        A, B = match_particles(Q, s.loads[isp])
        assert np.array_equal(A, B)     # bit exact
    f = s.ctx.download_fields(FIELDS6)
    for k in FIELDS6:
        assert np.array_equal(f[k], s.init_fields[k])
    # chunked append
    P = s.loads[0]
    s.ctx.upload_particles(0, P[:10])
    s.ctx.upload_particles(0, P[10:], append=True)
    A, B = match_particles(s.ctx.download_particles(0), P)
    assert np.array_equal(A, B)
    # page-locked host buffers (eb200_host_alloc) take the direct path past the staging ring; several chunks
    s = make_plasma(product_abi, 3, "triangle", [6, 5, 4], seed=1, ppc=7, options={"pin_bytes": 4096})
    big = np.random.default_rng(3).normal(size=(5000, 7))
    hp = s.ctx.host_empty(big.shape)
    assert getattr(s.ctx, "pinned_host", False)
    hp[...] = big
    s.ctx.upload_particles(1, hp)
    out = s.ctx.host_empty(big.shape)
    got = s.ctx.download_particles(1, out=out)
    A, B = match_particles(got, big)
    assert np.array_equal(A, B)
    fo = s.ctx.download_fields(FIELDS6, out={k: s.ctx.host_empty(s.ctx._np_shape()) for k in FIELDS6})
    for k in FIELDS6:
        assert np.array_equal(fo[k], s.init_fields[k])


def test_error_behaviour(product_abi):
    from epoch_b200 import PicContext, EpochError
    c = PicContext(2, "triangle", abi=product_abi)
    with pytest.raises(EpochError):
        c.update_eb_fields_half()           # grid not set
    c.set_grid([8, 8], [1e-6, 1e-6], [0.5e-6] * 2, [0.0] * 2, [8e-6] * 2, [0.0] * 2, [8e-6] * 2)
    with pytest.raises(EpochError):
        c.push_particles()                  # dt not set
    with pytest.raises(EpochError):
        c.set_field_bcs(["cpml_laser", "periodic", "periodic", "periodic"])
    with pytest.raises(EpochError):
        c.set_species(0, -Q0, M0, bcs=["dump", "dump", "periodic", "periodic"])      # not a particle boundary
    with pytest.raises(EpochError):
        c.set_species(0, -Q0, M0, bcs=["periodic", "reflect", "periodic", "periodic"])   # periodic on one side only
    with pytest.raises(EpochError):
        c.upload_particles(1, np.zeros((1, 6)))   # species not defined
    # a thermal wall without its temperature table is refused when the particles reach it
    c.set_species(0, -Q0, M0, bcs=["thermal", "thermal", "periodic", "periodic"])
    c.set_field_bcs(["clamp", "clamp", "periodic", "periodic"])
    c.set_time(1e-16, 0.0, 0)
    c.upload_particles(0, np.array([[4e-6, 4e-6, 0.0, 0.0, 0.0, 1.0]]))
    with pytest.raises(EpochError, match="thermal"):
        c.particle_bcs()


# ---- extended-stencil Maxwell solvers (fields.f90:53-179, :655-732) -----------------------------
CUSTOM = {"betaxy": 0.03, "betaxz": 0.02, "betayx": 0.05, "betayz": 0.01, "betazx": 0.04, "betazy": 0.015,
          "gammax": 0.006, "gammay": 0.004, "gammaz": 0.005, "deltax": 0.02, "deltay": -0.01, "deltaz": 0.015,
          "dt": 1.0e-15}


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,solver", [(1, "lehe_x"), (1, "custom"), (2, "lehe_x"), (2, "lehe_y"), (2, "pukhov"),
                                          (2, "custom"), (3, "lehe_x"), (3, "lehe_y"), (3, "lehe_z"), (3, "cowan"),
                                          (3, "pukhov"), (3, "custom")])
def test_maxwell_solver_half_steps(product_abi, oracle_abi, ndims, solver):
    n = [40, 24, 20][:ndims]
    cell = [1.0e-6, 1.3e-6, 0.8e-6][:ndims]
    sims = []
    for a in (product_abi, oracle_abi):
        s = Simulation(ndims, n, [0.0] * ndims, [c * k for c, k in zip(cell, n)], abi=a, maxwell_solver=solver,
                       stencil=CUSTOM if solver == "custom" else None)
        rng = np.random.default_rng(11)
        shp = s.ctx._np_shape()
        s.ctx.upload_fields(**{k: rng.normal(0, 1e6 if k[0] == "j" else 1.0, shp) for k in ALL9})
        for it in range(3):
            s.ctx.set_time(s.dt, (it + 0.5) * s.dt, it)
            s.ctx.update_eb_fields_half()
            s.ctx.update_eb_fields_final()
        sims.append(s)
    assert sims[0].coeffs == sims[1].coeffs and any(c != 0 for c in sims[0].coeffs[3:])
    compare_fields(sims[0], sims[1], FIELDS6)


@pytest.mark.gpu
@pytest.mark.parametrize("solver,printed", [("lehe_x", 311952693.446), ("cowan", 291962719.038), ("pukhov", 291961761.344)])
def test_maxwell_solver_group_velocity_3d(product_abi, solver, printed):
    """epoch3d/tests/test_maxwell_solvers.py::test_group_velocity on the CUDA library: same deck
    (open instead of CPML boundaries), same assertion (rtol 0.007), plus the value recorded there."""
    from test_maxwell_solvers import solver_deck, centroid_velocity, group_velocity_expected, exact_group_velocity
    sim = solver_deck(product_abi, 3, solver)
    vg, pts = centroid_velocity(sim, 25e-15)
    assert len(pts) >= 3
    assert np.isclose(vg, group_velocity_expected(sim, solver), rtol=0.007), (vg, group_velocity_expected(sim, solver))
    # the recorded values sit up to 0.65 % above the scheme's exact dispersion value (CPML run)
    assert abs(vg / printed - 1) < 0.01, (vg, printed)
    assert abs(vg / exact_group_velocity(sim) - 1) < 3e-3


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["1d_optimized", "2d_optimized", "2d_optimized_symm", "2d_optimized_xaxis",
                                  "3d_optimized", "3d_optimized_xaxis", "3d_optimized_xaxis_soft"])
def test_custom_stencil_goldens_on_gpu(product_abi, name):
    """epoch{1,2,3}d/tests/test_custom_stencils.py on the CUDA library (these decks use the open
    boundaries of this path): the reference's assertion and its recorded velocities."""
    from test_maxwell_solvers import custom_deck, check_custom
    check_custom(custom_deck(product_abi, name), name)


@pytest.mark.gpu
@pytest.mark.parametrize("ndims", [1, 2, 3])
@pytest.mark.parametrize("fbc", ["periodic", "open"])
def test_current_smoothing_and_correction(product_abi, oracle_abi, ndims, fbc):
    """current_finish with smooth_current (2 passes + compensation pass, strides 1, 2, 3) and
    use_current_correction, after a real push: CUDA == oracle, ghosts included."""
    bc_field = [fbc, fbc] + ["periodic"] * (2 * ndims - 2)
    pbc = ["reflect" if fbc == "open" else "periodic"] * 2 + ["periodic"] * (2 * ndims - 2)
    sims = [make_plasma(a, ndims, "triangle", [16, 12, 10], seed=8, ppc=3, bc_field=bc_field, bcs=pbc)
            for a in (product_abi, oracle_abi)]
    for s in sims:
        s.ctx.set_current_options(smooth=True, smooth_its=2, smooth_comp_its=1, strides=[1, 2, 3],
                                  initial_j=[1.0e9, -2.0e9, 0.5e9])
        s.step_once()
    compare_fields(sims[0], sims[1], ALL9, tol=1e-11)


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,shape,fbc", [(1, "triangle", "open"), (2, "triangle", "open"), (2, "tophat", "open"),
                                             (3, "triangle", "open"), (3, "bspline3", "open"),
                                             (1, "triangle", "cpml_outflow"), (2, "triangle", "cpml_outflow"),
                                             (3, "triangle", "cpml_outflow")])
def test_moving_window(product_abi, oracle_abi, ndims, shape, fbc):
    """housekeeping/window.F90 through the ABI: a window at 0.9 c with particle insertion on the
    right, removal on the left, field shift + incoming-cell fix, 12 steps (4-7 shifts): CUDA ==
    oracle for fields and every particle.  With CPML layers on x the psi arrays shift as well."""
    from test_window import window_deck
    sims = [window_deck(a, ndims, shape, fbc=fbc) for a in (product_abi, oracle_abi)]
    for s in sims:
        for _ in range(12):
            s.step_once()
    assert sims[0].window_offset == sims[1].window_offset > 3 * sims[0].d[0]
    compare_fields(sims[0], sims[1], ALL9, tol=1e-10)
    compare_particles(sims[0], sims[1], tol=1e-10)


# ---- CPML absorbing boundaries (boundary.F90: set_cpml_helpers, cpml_advance_*_currents) --------
@pytest.mark.gpu
@pytest.mark.parametrize("ndims,order,solver,axes", [
    (1, 2, "yee", "x"), (1, 4, "yee", "x"), (1, 6, "yee", "x"), (1, 2, "lehe_x", "x"),
    (2, 2, "yee", "xy"), (2, 4, "yee", "x"), (2, 2, "pukhov", "xy"), (2, 6, "yee", "y"),
    (3, 2, "yee", "xyz"), (3, 4, "yee", "xz"), (3, 6, "yee", "y"), (3, 2, "cowan", "xyz"), (3, 2, "lehe_z", "z")])
def test_cpml_plasma_steps(product_abi, oracle_abi, ndims, order, solver, axes):
    """Six full PIC steps in a box whose listed axes end in CPML layers (cpml_laser with an
    attached laser on x_min, cpml_outflow elsewhere), open particle boundaries on those sides:
    fields (kappa-scaled updates + psi currents), the particles that survive and J agree."""
    from epoch_b200.driver import LaserSpec
    bc_field, bcs = ["periodic"] * (2 * ndims), ["periodic"] * (2 * ndims)
    for d, ax in enumerate("xyz"[:ndims]):
        if ax in axes:
            bc_field[2 * d:2 * d + 2] = ["cpml_laser" if ax == "x" else "cpml_outflow", "cpml_outflow"]
            bcs[2 * d:2 * d + 2] = ["open", "open"]
    lasers = []
    if "x" in axes:
        prof = {1: None, 2: lambda y: np.exp(-((y - 6e-6) / 3e-6) ** 2),
                3: lambda y, z: np.exp(-((y - 6e-6) ** 2 + (z - 5e-6) ** 2) / (3e-6) ** 2)}[ndims]
        lasers = [LaserSpec(1, intensity_w_cm2=1e15, lambda_=4e-6, profile=prof, pol_angle=0.3)]
    sims = [make_plasma(a, ndims, "triangle", [20, 12, 10], seed=41, ppc=3, u_th=1.0, bc_field=bc_field, bcs=bcs,
                        field_order=order, maxwell_solver=solver, lasers=lasers, e_amp=1e9, b_amp=3.0)
            for a in (product_abi, oracle_abi)]
    for s in sims:
        for _ in range(6):
            s.step_once()
    assert sims[0].ctx.particle_count(0) == sims[1].ctx.particle_count(0)
    compare_particles(sims[0], sims[1], tol=1e-10)
    compare_fields(sims[0], sims[1], ALL9, tol=1e-10)


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,solver,rtol,printed,ptol", [
    (1, "yee", 0.022, 298540658.530, 1e-10), (2, "yee", 0.012, 289199289.192, 1e-10),
    (2, "pukhov", 0.012, 292013249.255, 1e-10), (3, "yee", 0.007, 279231545.324, 1e-10),
    (3, "cowan", 0.007, 291962719.038, 1e-10), (3, "pukhov", 0.007, 291961761.344, 1e-10),
    (3, "lehe_x", 0.007, 311952693.446, 1e-3)])
def test_group_velocity_cpml_gpu(product_abi, ndims, solver, rtol, printed, ptol):
    """epoch{1,2,3}d/tests/test_maxwell_solvers.py with the decks' own cpml_laser / cpml_outflow
    boundaries on the CUDA library: the reference's assertion, and the velocities its comments
    record (Yee, Cowan, Pukhov to 10 digits; the Lehe recordings predate its present coefficients)."""
    from test_maxwell_solvers import cpml_deck, check_cpml
    check_cpml(cpml_deck(product_abi, ndims, solver), ndims, solver, rtol, printed, ptol)


# ---- derived grid variables (io/calc_df.F90) -------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("ndims,shape,pbc,fbc", [(1, "triangle", "periodic", "periodic"), (1, "tophat", "reflect", "open"),
                                                 (2, "triangle", "periodic", "periodic"), (2, "bspline3", "reflect", "conduct"),
                                                 (3, "triangle", "reflect", "open"), (3, "tophat", "periodic", "periodic"),
                                                 (3, "bspline3", "periodic", "periodic")])
def test_derived_grid_variables(product_abi, oracle_abi, ndims, shape, pbc, fbc):
    """every calc_df.F90 routine (mass/charge/number density, ekbar, ekflux, Poynting flux, ppc,
    average weight, temperature, per-species current, average momentum) after two PIC steps:
    CUDA == oracle on the whole array, ghosts included"""
    from test_derived import CASES
    bc_field = [fbc, fbc] + ["periodic"] * (2 * ndims - 2)
    bcs = [pbc, pbc] + ["periodic"] * (2 * ndims - 2)
    sims = [make_plasma(a, ndims, shape, [14, 10, 8], seed=17, ppc=3, u_th=0.8, bc_field=bc_field, bcs=bcs)
            for a in (product_abi, oracle_abi)]
    for s in sims:
        for _ in range(2):
            s.step_once()
    for kind, isp, direction in CASES:
        a, b = (s.ctx.calc_derived(kind, isp, direction) for s in sims)
        scale = np.max(np.abs(b))
        assert scale > 0, (kind, isp, direction)
        tol = 1e-9 if kind == "temperature" else 1e-11      # the variance subtracts the mean momentum
        assert np.max(np.abs(a - b)) <= tol * scale, (kind, isp, direction, np.max(np.abs(a - b)) / scale)


# ---- load-balancing hooks (housekeeping/balance.F90) ---------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("ndims,shape", [(1, "triangle"), (2, "tophat"), (3, "triangle"), (3, "bspline3")])
def test_get_load(product_abi, oracle_abi, ndims, shape):
    """part_load_func histograms after three steps: CUDA == oracle exactly (integers)"""
    sims = [make_plasma(a, ndims, shape, [14, 10, 8], seed=23, ppc=4, u_th=0.8) for a in (product_abi, oracle_abi)]
    for s in sims:
        for _ in range(3):
            s.step_once()
    la, lb = (s.ctx.get_load(s.grid_min, s.n_global) for s in sims)
    for a in range(ndims):
        assert np.array_equal(la[a], lb[a]) and la[a].sum() == sum(len(p) for p in sims[0].loads)


@pytest.mark.gpu
def test_redistribute_single_rank_identity(product_abi, oracle_abi):
    """re-slicing onto the same (whole) domain goes through the full device path (pack, copy,
    unpack, halo refresh, re-sort) and must change nothing"""
    sims = [make_plasma(a, 3, "triangle", [12, 10, 8], seed=29, ppc=3, bc_field=["open"] * 2 + ["periodic"] * 4,
                        bcs=["reflect"] * 2 + ["periodic"] * 4) for a in (product_abi, oracle_abi)]
    for s in sims:
        s.step_once()
    g = sims[0]
    g.ctx.redistribute(g.cell_mins, g.cell_maxs, g.cell_mins, g.cell_maxs, g.grid_min, g.grid_min_local,
                       g.min_local, g.max_local)
    g._setup_lasers()
    for s in sims:
        s.step_once()
    compare_fields(sims[0], sims[1], ALL9, tol=1e-10)
    compare_particles(sims[0], sims[1], tol=1e-10)


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,shape,pbc,n,window,expect", [
    (3, "triangle", "periodic", [16, 8, 48], 8000, True), (3, "tophat", "periodic", [16, 8, 48], 9000, True),
    (2, "triangle", "periodic", [32, 64], 6000, True), (2, "tophat", "reflect", [32, 64], 7000, True),
    (3, "triangle", "open", [16, 8, 48], 8000, True), (3, "triangle", "periodic", [16, 8, 48], 600, False),
    (3, "triangle", "periodic", [16, 8, 48], "auto", True), (3, "triangle", "periodic", [16, 8, 48], 0, False),
    (3, "bspline3", "periodic", [16, 8, 48], 9000, True), (2, "bspline3", "open", [32, 64], 7000, True),
    (1, "triangle", "periodic", [1500], 5000, True), (1, "tophat", "reflect", [1500], 5000, True),
    (1, "bspline3", "open", [1500], 5000, True)])
def test_deferred_permute(product_abi, oracle_abi, ndims, shape, pbc, n, window, expect):
    """The sort leaves its permute to the next push (gather through src, write into the other state of the
    arrays).  A window far smaller than the species forces many launches per step in both directions and,
    where the displacements do not fit (window 600 < one tile), the fall-back permute; downloads,
    energy sums and diagnostics in mid-run settle a deferred permute.  Everything must equal the oracle
    (as does window = 0, the plain sort)."""
    bcs = [pbc] * (2 * ndims)
    fbc = ["periodic" if pbc == "periodic" else "clamp"] * (2 * ndims)
    opts = {"defer_window": window, "deposit_scratch": 2048}
    sims = [make_plasma(a, ndims, shape, n, seed=5, ppc=6, u_th=0.4, e_amp=1e9, b_amp=3.0, bcs=bcs, bc_field=fbc,
                        options=opts) for a in (product_abi, oracle_abi)]
    for it in range(9):
        for s in sims:
            s.step_once()
        if it == 3:
            e = [s.ctx.energy() for s in sims]
            assert np.allclose(e[0][1], e[1][1], rtol=1e-11, atol=0)
        if it in (4, 8):
            compare_particles(sims[0], sims[1], tol=1e-10)
            compare_fields(sims[0], sims[1], ALL9, tol=1e-10)
    deferred, permuted = sims[0].ctx.timer("deferred_sorts")[1], sims[0].ctx.timer("permuted_sorts")[1]
    assert sims[0].ctx.timer("cell_mismatch")[1] == 0
    assert (deferred > 0) == expect, (deferred, permuted)


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,shape,pbc,window", [(3, "triangle", "periodic", "auto"), (3, "tophat", "reflect", 0),
                                                    (2, "triangle", "periodic", "auto"), (2, "tophat", "open", 4000),
                                                    (3, "bspline3", "periodic", "auto"), (2, "bspline3", "reflect", "auto"),
                                                    (1, "triangle", "periodic", "auto"), (1, "tophat", "open", 3000),
                                                    (1, "bspline3", "reflect", "auto")])
def test_deterministic_deposit(product_abi, oracle_abi, ndims, shape, pbc, window):
    """option deterministic_deposit: two runs of the same deck give bitwise identical currents, fields and
    particle arrays (canonical order inside a cell, fixed-point overflow sums, ordered sum of the tile
    halos), and they still equal the oracle within the usual tolerance."""
    n = [24, 16, 12] if ndims == 3 else ([40, 48] if ndims == 2 else [1200])
    bcs = [pbc] * (2 * ndims)
    fbc = ["periodic" if pbc == "periodic" else "clamp"] * (2 * ndims)
    opts = {"deterministic_deposit": 1, "defer_window": window, "deposit_scratch": 4096}

    def run(abi):
        s = make_plasma(abi, ndims, shape, n, seed=9, ppc=7, u_th=0.5, e_amp=1e9, b_amp=3.0, bcs=bcs, bc_field=fbc,
                        options=opts)
        for _ in range(6):
            s.step_once()
        return s
    a, b, o = run(product_abi), run(product_abi), run(oracle_abi)
    fa, fb = a.ctx.download_fields(ALL9), b.ctx.download_fields(ALL9)
    for k in ALL9:
        assert np.array_equal(fa[k], fb[k]), k
    for isp in range(len(a.species)):
        assert np.array_equal(a.ctx.download_particles(isp), b.ctx.download_particles(isp))
    compare_fields(a, o, ALL9, tol=1e-10)
    compare_particles(a, o, tol=1e-10)
    # not on the general kernel
    if ndims == 3:
        s = make_plasma(product_abi, 1, "triangle", [32], seed=1, ppc=2, options={"deterministic_deposit": 1, "push_kernel": "baseline"})
        with pytest.raises(Exception):
            s.step_once()


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,shape,n,ppc", [(2, "triangle", [32, 32], 70), (3, "triangle", [8, 8, 8], 70),
                                               (2, "bspline3", [16, 32], 80), (1, "triangle", [600], 90)])
def test_dense_tiles(product_abi, oracle_abi, ndims, shape, n, ppc):
    """tiles with more than PUSH_CHUNK (16384) particles are pushed by several CTAs; dense cells overflow the
    deposit's work-item list (inline path): same answers as the oracle"""
    sims = [make_plasma(a, ndims, shape, n, seed=13, ppc=ppc, u_th=0.5, e_amp=1e9, b_amp=3.0, two_species=False)
            for a in (product_abi, oracle_abi)]
    for _ in range(3):
        for s in sims:
            s.step_once()
    compare_particles(sims[0], sims[1], tol=1e-10)
    compare_fields(sims[0], sims[1], ALL9, tol=1e-10)
    assert sims[0].ctx.timer("cell_mismatch")[1] == 0
