"""Boundary planes of setup_field_boundaries in outflow_bcs_* (housekeeping/setup.F90:426-510,
laser.f90:536,559-572 and the five cyclic copies).

Known answer: a uniform static field solves Maxwell's equations in vacuum, so with the planes
captured from the initial fields an open (simple_outflow) boundary must leave it unchanged --
bz(0) = sum (2 (Ey + c Bz) - 2 Ey + (lx - c) Bz) = Bz.  Without the planes the boundary treats
the field as an outgoing wave and radiates it away.  The normal E component is left at zero: the
reference clamps it on the boundary node (field_clamp_zero, boundary.F90:702-783)."""
import numpy as np
import pytest

from helpers import FIELDS6
from epoch_b200.driver import Simulation


def _deck(abi, ndims, axis):
    n = [24, 16, 12][:ndims]
    bc = ["periodic"] * (2 * ndims)
    bc[2 * axis:2 * axis + 2] = ["open", "open"]
    sim = Simulation(ndims, n, [0.0] * ndims, [1e-6 * x for x in n], bc_field=bc, abi=abi)
    shp = sim.ctx._np_shape()
    amp = {"ex": 3e9, "ey": -2e9, "ez": 1.5e9, "bx": 4.0, "by": -7.0, "bz": 5.0}
    amp["e" + "xyz"[axis]] = 0.0
    f = {k: np.full(shp, amp[k]) for k in FIELDS6}
    sim.ctx.upload_fields(**f)
    return sim, amp


def _interior(sim, a):
    ng = sim.ctx.ng
    return a[tuple(slice(ng, -ng) for _ in range(sim.ndims))]


def _run(abi, ndims, axis, planes=True, steps=25):
    sim, amp = _deck(abi, ndims, axis)
    if not planes:
        sim._bnd_fields_done = True          # skip setup_field_boundaries: planes stay zero
    sim.start()
    for _ in range(steps):
        sim.step_once()
    f = sim.ctx.download_fields()
    dev = max(np.max(np.abs(_interior(sim, f[k]) - amp[k])) / (3e9 if k[0] == "e" else 7.0) for k in FIELDS6)
    sim.ctx.close()
    return dev, f


CASES = [(1, 0), (2, 0), (2, 1), (3, 0), (3, 1), (3, 2)]


@pytest.mark.parametrize("ndims,axis", CASES)
def test_uniform_field_through_outflow_oracle(oracle_abi, ndims, axis):
    dev, _ = _run(oracle_abi, ndims, axis)
    assert dev < 1e-12
    dev0, _ = _run(oracle_abi, ndims, axis, planes=False, steps=5)
    assert dev0 > 1e-3          # the test has teeth: without the planes the field radiates away


def test_set_boundary_fields_upload(oracle_abi):
    """planes handed over by the host == planes captured on the device"""
    sim, amp = _deck(oracle_abi, 2, 0)
    shp_plane = (sim.n_local[1] + 2 * sim.ctx.ng,)
    for b in (1, 2):
        sim.ctx.set_boundary_fields(b, {k: np.full(shp_plane, amp[k]) for k in FIELDS6})
    sim._bnd_fields_done = True
    sim.start()
    for _ in range(10):
        sim.step_once()
    f = sim.ctx.download_fields()
    assert max(np.max(np.abs(_interior(sim, f[k]) - amp[k])) for k in ("by", "bz")) < 1e-11


@pytest.mark.gpu
@pytest.mark.parametrize("ndims,axis", CASES)
def test_uniform_field_through_outflow_gpu(product_abi, oracle_abi, ndims, axis):
    dev, fa = _run(product_abi, ndims, axis)
    assert dev < 1e-12
    _, fb = _run(oracle_abi, ndims, axis)
    for k in FIELDS6:
        assert np.max(np.abs(fa[k] - fb[k])) <= 1e-12 * (3e9 if k[0] == "e" else 7.0), k
