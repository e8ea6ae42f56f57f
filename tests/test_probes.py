"""Particle probes (particles.F90:567-606): the oracle against a numpy restatement of the crossing test
(not gpu), the CUDA library against the oracle (gpu)."""
import numpy as np
import pytest

from helpers import make_plasma, match_particles
from epoch_b200.driver import C


def setup(abi, nd, shape):
    sim = make_plasma(abi, nd, shape, [24, 16, 12], seed=19, ppc=6, u_th=0.6)
    L = sim.length
    mid = [0.5 * L[d] for d in range(nd)]
    nrm = [1.0, 0.5, -0.25][:nd]
    mc2 = sim.species[0].mass * C * C
    sim.ctx.set_probe(0, 0, mid, nrm, 0.0, 1e300)                    # every energy, oblique plane
    sim.ctx.set_probe(0, 3, mid, [-1.0] + [0.0] * (nd - 1), 0.05 * mc2, 0.4 * mc2)   # other direction, energy window
    sim.ctx.set_probe(1, 1, [0.3 * L[0]] + mid[1:], [1.0] + [0.0] * (nd - 1), 0.0, 1e300)
    return sim, mid, nrm, mc2


@pytest.mark.parametrize("nd,shape", [(1, "triangle"), (2, "tophat"), (3, "triangle")])
def test_oracle_probe_against_numpy(oracle_abi, nd, shape):
    sim, mid, nrm, mc2 = setup(oracle_abi, nd, shape)
    sim.ctx.set_time(sim.dt, 0.5 * sim.dt, 0)
    before = sim.ctx.download_particles(0)
    sim.ctx.push_particles()
    after = sim.ctx.download_particles(0)
    A, B = match_particles(before, after)
    B = B.copy()
    for d in range(nd):          # the probe sees the position before particle_bcs: undo the periodic wrap
        B[:, d] -= sim.length[d] * np.round((B[:, d] - A[:, d]) / sim.length[d])
    d_init = sum(nrm[d] * (mid[d] - A[:, d]) for d in range(nd))
    d_final = sum(nrm[d] * (mid[d] - B[:, d]) for d in range(nd))
    hit = (d_final < 0.0) & (d_init >= 0.0)
    got, n = sim.ctx.download_probe(0, 0)
    assert n == len(got) == int(hit.sum()) > 0
    G, H = match_particles(got, B[hit])
    assert np.array_equal(G[:, nd:], H[:, nd:])
    assert np.max(np.abs(G[:, :nd] - H[:, :nd])) < 1e-9 * sim.length[0]       # (the un-wrapping rounds)
    # the energy window of the second probe
    mcs = sim.species[0].mass * C
    u2 = np.sum((B[:, nd:nd + 3] / mcs) ** 2, axis=1)
    ek = u2 / (np.sqrt(u2 + 1.0) + 1.0) * mc2
    hit2 = ((mid[0] - B[:, 0]) * -1.0 < 0.0) & ((mid[0] - A[:, 0]) * -1.0 >= 0.0) & (ek > 0.05 * mc2) & (ek < 0.4 * mc2)
    got2, n2 = sim.ctx.download_probe(0, 3)
    assert n2 == int(hit2.sum()) > 0
    # cleared by the download
    assert sim.ctx.download_probe(0, 0)[1] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("nd,shape", [(1, "triangle"), (2, "tophat"), (2, "bspline3"), (3, "triangle")])
def test_cuda_probe_against_oracle(product_abi, oracle_abi, nd, shape):
    sims = [setup(a, nd, shape)[0] for a in (product_abi, oracle_abi)]
    for s in sims:
        for _ in range(3):
            s.step_once()
    for isp, ipr in ((0, 0), (0, 3), (1, 1)):
        (a, na), (b, nb) = sims[0].ctx.download_probe(isp, ipr), sims[1].ctx.download_probe(isp, ipr)
        assert na == nb > 0, (isp, ipr, na, nb)
        A, B = match_particles(a, b)
        assert np.array_equal(A[:, -1], B[:, -1])
        assert np.max(np.abs(A[:, :nd] - B[:, :nd])) / sims[0].length[0] < 1e-11
        assert np.max(np.abs(A[:, nd:nd + 3] - B[:, nd:nd + 3])) <= 1e-11 * np.max(np.abs(B[:, nd:nd + 3]))
