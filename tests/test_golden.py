"""Committed fixtures of tests/golden/ (made by tests/golden/make_fixtures.py):
* reference_goldens.json -- the numbers the reference's own tests assert (cited there); the oracle
  (not gpu) and the CUDA library (gpu) must reproduce them;
* push_deposit_*.npz -- ORACLE-generated inputs/outputs of two full PIC steps (the reference pins no
  particle-dependent number): the oracle must still reproduce them bit for bit (a guard against silent
  changes of the checker), the CUDA path within the 1e-12 tolerances of BASELINE.json's north_star."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_fixtures as mf                                   # noqa: E402
from helpers import ALL9, relerr                              # noqa: E402
import test_oracle as to                                      # noqa: E402

GOLD = json.load(open(os.path.join(HERE, "golden", "reference_goldens.json")))


def test_reference_goldens_file_matches_the_tests():
    """the tables used by test_oracle / test_gpu_parity are the committed reference numbers"""
    for dim, table in (("1d", to.GOLD_1D), ("2d", to.GOLD_2D), ("3d", to.GOLD_3D)):
        ref = {int(k): v for k, v in GOLD["laser_sum_of_squares"][dim]["dumps"].items()}
        assert ref == table


def _check_inputs(sim, fx):
    for i, L in enumerate(sim.loads):
        assert np.array_equal(L, fx["load%d" % i])
    for k, v in sim.init_fields.items():
        assert np.array_equal(v, fx["init_" + k])
    assert sim.dt == float(fx["dt"])


@pytest.mark.parametrize("ndims,shape,n", mf.CASES)
def test_oracle_reproduces_fixtures(oracle_abi, ndims, shape, n):
    fx = np.load(mf.name(ndims, shape))
    sim = mf.build(oracle_abi, ndims, shape, n)
    _check_inputs(sim, fx)
    out = mf.run(sim)
    for k, v in out.items():
        assert np.array_equal(v, fx[k]), k


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,shape,n", mf.CASES)
def test_cuda_matches_fixtures(product_abi, ndims, shape, n):
    """positions 1e-12 of the domain length, momenta 1e-12 relative, grid E/B/J 1e-12 norm-wise per step
    (two steps: 1e-11 allowed)"""
    fx = np.load(mf.name(ndims, shape))
    sim = mf.build(product_abi, ndims, shape, n)
    _check_inputs(sim, fx)
    out = mf.run(sim)
    for isp in range(len(sim.species)):
        A, B = out["part%d" % isp], fx["part%d" % isp]
        assert A.shape == B.shape and np.array_equal(A[:, -1], B[:, -1])
        for d in range(ndims):
            assert np.max(np.abs(A[:, d] - B[:, d])) / sim.length[d] < 1e-11
        assert relerr(A[:, ndims:ndims + 3], B[:, ndims:ndims + 3]) < 1e-11
    for k in ALL9:
        assert relerr(out[k], fx[k]) < 1e-11, k
