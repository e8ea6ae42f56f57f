"""Mixed per-species particle boundaries (row a10: current_bcs(species), particles.F90:609,
boundary.F90:1123-1144,1976-1985): each species' current is folded with ITS boundary codes and the
ghosts cleared.  Oracle check (not gpu): the current is linear in the species, so the mixed run equals the
sum of two common-BC runs in which the other species carries no current; CUDA against the oracle (gpu)."""
import numpy as np
import pytest

from helpers import make_plasma, match_particles, relerr
from epoch_b200.driver import Q0, M0

JS = ("jx", "jy", "jz")


def deck(abi, nd, shape, bc0, bc1, zero=(False, False)):
    """electrons with x boundaries bc0, protons with bc1 (other axes periodic)"""
    sim = make_plasma(abi, nd, shape, [16, 10, 8], seed=23, ppc=5, u_th=0.8, e_amp=1e9, b_amp=3.0)
    per = ["periodic"] * (2 * nd - 2)
    sim.ctx.set_species(0, -Q0, M0, bcs=[bc0, bc0] + per, zero_current=zero[0])
    sim.ctx.set_species(1, Q0, 1836.2 * M0, bcs=[bc1, bc1] + per, zero_current=zero[1])
    return sim


def one_step(sim):
    sim.ctx.set_time(sim.dt, 0.5 * sim.dt, 0)
    sim.ctx.push_particles()
    sim.ctx.current_finish()
    return sim.ctx.download_fields(JS), [sim.ctx.download_particles(i) for i in range(2)]


@pytest.mark.parametrize("nd,shape", [(1, "triangle"), (2, "tophat"), (3, "triangle"), (2, "bspline3")])
def test_oracle_mixed_is_sum_of_common_runs(oracle_abi, nd, shape):
    jm, pm = one_step(deck(oracle_abi, nd, shape, "periodic", "reflect"))
    ja, pa = one_step(deck(oracle_abi, nd, shape, "periodic", "periodic", zero=(False, True)))
    jb, pb = one_step(deck(oracle_abi, nd, shape, "reflect", "reflect", zero=(True, False)))
    for k in JS:
        assert relerr(jm[k], ja[k] + jb[k]) < 1e-12, k
        assert np.max(np.abs(jb[k])) > 0 and np.max(np.abs(ja[k])) > 0
    A, B = match_particles(pm[0], pa[0])
    assert np.array_equal(A, B)
    A, B = match_particles(pm[1], pb[1])
    assert np.array_equal(A, B)


@pytest.mark.gpu
@pytest.mark.parametrize("nd,shape,bc0,bc1", [(1, "triangle", "periodic", "reflect"), (2, "tophat", "reflect", "open"),
                                              (3, "triangle", "periodic", "reflect"), (2, "bspline3", "open", "reflect")])
def test_cuda_mixed_bcs_against_oracle(product_abi, oracle_abi, nd, shape, bc0, bc1):
    fbc = ["periodic" if "periodic" in (bc0, bc1) else "clamp"] * 2 + ["periodic"] * (2 * nd - 2)
    sims = []
    for abi in (product_abi, oracle_abi):
        s = deck(abi, nd, shape, bc0, bc1)
        s.ctx.set_field_bcs(fbc)
        sims.append(s)
    for _ in range(3):
        for s in sims:
            s.step_once()
    fa, fb = sims[0].ctx.download_fields(), sims[1].ctx.download_fields()
    for k in fa:
        assert relerr(fa[k], fb[k]) < 1e-10, k
    for isp in range(2):
        A, B = match_particles(sims[0].ctx.download_particles(isp), sims[1].ctx.download_particles(isp))
        assert A.shape == B.shape
        assert np.max(np.abs(A[:, :nd] - B[:, :nd])) / sims[0].length[0] < 1e-10
