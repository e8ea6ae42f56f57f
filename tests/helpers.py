"""Shared builders for the parity tests: the same seeded deck is handed to the
oracle and to the CUDA library through the same `Simulation` driver."""
import numpy as np

from epoch_b200.driver import Simulation, SpeciesSpec, LaserSpec, Q0, M0, C, gauss, PI  # noqa: F401

SHAPE_NAMES = ("tophat", "triangle", "bspline3")
FIELDS6 = ("ex", "ey", "ez", "bx", "by", "bz")
ALL9 = FIELDS6 + ("jx", "jy", "jz")


def make_plasma(abi, ndims, shape, n, seed=0, ppc=4, u_th=0.3, e_amp=1e11, b_amp=3e2,
                bc_field=None, bcs=None, two_species=True, cell=1.0e-6, **kw):
    """Random plasma in random fields: exercises gather, rotation, cell crossings."""
    rng = np.random.default_rng(seed)
    n = list(n)[:ndims]
    L = [cell * x for x in n]
    species = [SpeciesSpec("electron", -Q0, M0, bcs=bcs, density=1e25, temperature=(1e6,) * 3)]
    if two_species:
        species.append(SpeciesSpec("proton", Q0, 1836.2 * M0, bcs=bcs, density=1e25, temperature=(1e6,) * 3))
    sim = Simulation(ndims, n, [0.0] * ndims, L, shape=shape, species=species, abi=abi,
                     bc_field=bc_field, **kw)
    ncell = int(np.prod(sim.n_local))
    npart = ppc * ncell
    loads = []
    for isp, sp in enumerate(species):
        P = np.empty((npart, ndims + 4))
        for d in range(ndims):
            P[:, d] = rng.uniform(sim.min_local[d], sim.max_local[d], npart)
        P[:, ndims:ndims + 3] = rng.normal(0.0, u_th, (npart, 3)) * sp.mass * C
        P[:, ndims + 3] = rng.uniform(0.5, 1.5, npart) * 1e25 * cell ** ndims / ppc
        sim.ctx.upload_particles(isp, P)
        loads.append(P)
    shp = sim.ctx._np_shape()
    f = {k: rng.normal(0.0, 1.0, shp) * (e_amp if k[0] == "e" else b_amp) for k in FIELDS6}
    sim.ctx.upload_fields(**f)
    sim.loads = loads
    sim.init_fields = f
    return sim


def sort_particles(P):
    """canonical order for comparing particle sets (the device keeps them cell sorted)"""
    idx = np.lexsort(tuple(P[:, k] for k in range(P.shape[1] - 1, -1, -1)))
    return P[idx]


def match_particles(A, B):
    """Pair particles of two sets by their weight (unique per particle in the tests)."""
    ia = np.argsort(A[:, -1], kind="stable")
    ib = np.argsort(B[:, -1], kind="stable")
    return A[ia], B[ib]


def relerr(a, b):
    s = max(np.max(np.abs(b)), 1e-300)
    return float(np.max(np.abs(a - b)) / s)


# ---- the reference's own particle load (oracle/orc_loader.cpp: KISS + Box-Muller + auto_load) ----
def reference_load(oracle_abi, sim_oracle, loads, seed=7842432):
    """auto_load (helper.F90:89-163) on an ORACLE-backed Simulation: random_init(seed + rank) as
    set_initial_values does (setup.F90:505-510), then, species by species, positions, weights
    and momenta from the global generator.  `loads` = [(ppc, density, (Tx, Ty, Tz), (dpx, dpy, dpz)), ...]
    per species.  Returns the packed particle arrays (wire format) per species."""
    import ctypes as C
    lib = oracle_abi.lib
    lib.orc_random_init.argtypes = [C.c_int32]
    lib.orc_load_uniform.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.POINTER(C.c_double),
                                     C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    lib.orc_load_uniform.restype = C.c_int32
    assert sim_oracle.ctx.abi is oracle_abi
    lib.orc_random_init(int(seed + sim_oracle.rank))
    out = []
    for isp, (ppc, density, temp, drift) in enumerate(loads):
        n = C.c_int64(0)
        t3 = (C.c_double * 3)(*temp)
        d3 = (C.c_double * 3)(*drift)
        rc = lib.orc_load_uniform(sim_oracle.ctx._ctx, isp, int(ppc), float(density), t3, d3, C.byref(n))
        assert rc == 0, rc
        out.append(sim_oracle.ctx.download_particles(isp))
    return out
