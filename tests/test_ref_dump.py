"""Parity against EPOCH's own output, for the day someone can run it: `--ref-dump DIR` (or
tests/golden/refdump/) holds ref{1,2,3}d_{0,1}.npz made by tools/ref_build/build_and_run.sh from
SDF dumps of the reference before and after ONE step (last bullet).  The first dump is
uploaded, one step is taken with the dump's own dt, and positions, momenta and ex..jz are compared
with the second dump at the north star's tolerance (1e-12).  The reference cannot be built in this
container, so without a dump directory these tests are skipped; the harness itself is exercised
by a dump written by the oracle in the same format."""
import os

import numpy as np
import pytest

from helpers import Simulation, SpeciesSpec, Q0, M0, ALL9, FIELDS6

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-12
MASS = {"electron": M0, "proton": 1836.2 * M0}
CHARGE = {"electron": -Q0, "proton": Q0}


def _dump_dir(request):
    d = request.config.getoption("--ref-dump")
    if not d and os.path.isdir(os.path.join(HERE, "golden", "refdump")):
        d = os.path.join(HERE, "golden", "refdump")
    return d


def _load(d, name):
    a, b = (np.load(os.path.join(d, f"{name}_{i}.npz")) for i in (0, 1))
    return a, b


def step_from_dump(abi, d0, d1, shape="triangle"):
    """one step from the state of dump d0, with dt = time(d1) - time(d0)"""
    cells = [int(x) for x in d0["cells"]]
    nd = len(cells)
    names = [str(s) for s in d0["species"]]
    species = [SpeciesSpec(n, CHARGE[n], MASS[n], density=1e25, temperature=(2e8,) * 3) for n in names]
    dt = float(d1["time"]) - float(d0["time"])
    sim = Simulation(nd, cells, list(d0["edges"][:, 0]), list(d0["edges"][:, 1]), shape=shape, species=species,
                     abi=abi, dt=dt)
    ng = sim.ctx.ng
    shp = sim.ctx._np_shape()
    inner = tuple(slice(ng, -ng) for _ in range(nd))
    f = {}
    for k in FIELDS6:
        a = np.zeros(shp)
        a[inner] = np.asarray(d0[k]).T if nd > 1 else np.asarray(d0[k])      # npz: Fortran (x first); device arrays: x last
        f[k] = a
    sim.ctx.upload_fields(**f)
    for isp, n in enumerate(names):
        cols = [d0[f"{n}_{a}"] for a in "xyz"[:nd]] + [d0[f"{n}_p{a}"] for a in "xyz"] + [d0[f"{n}_w"]]
        sim.ctx.upload_particles(isp, np.stack(cols, axis=1))
    sim.start()                # fills the ghost cells (efield_bcs / bfield_final_bcs, epoch3d.F90:141-158)
    sim.step_once()
    return sim, names, inner


def compare_with_dump(sim, names, inner, d1, tol=TOL):
    nd = sim.ndims
    f = sim.ctx.download_fields()
    for k in ALL9:
        ref = np.asarray(d1[k]).T if nd > 1 else np.asarray(d1[k])
        scale = max(np.max(np.abs(ref)), 1e-300)
        assert np.max(np.abs(f[k][inner] - ref)) <= tol * scale, k
    length = [sim.maxs[a] - sim.mins[a] for a in range(nd)]
    for isp, n in enumerate(names):
        P = sim.ctx.download_particles(isp)
        w = d1[f"{n}_w"]
        # the device keeps particles cell sorted: pair by weight + initial order is lost, so match on (w, x, px) sort
        ref = np.stack([d1[f"{n}_{a}"] for a in "xyz"[:nd]] + [d1[f"{n}_p{a}"] for a in "xyz"] + [w], axis=1)
        ia = np.lexsort((P[:, nd], P[:, 0]))
        ib = np.lexsort((ref[:, nd], ref[:, 0]))
        A, B = P[ia], ref[ib]
        for a in range(nd):
            dx = np.abs(A[:, a] - B[:, a])
            dx = np.minimum(dx, np.abs(dx - length[a]))          # periodic wrap
            assert np.max(dx) <= tol * length[a], (n, "pos", a)
        pm = max(np.max(np.abs(B[:, nd:nd + 3])), 1e-300)
        assert np.max(np.abs(A[:, nd:nd + 3] - B[:, nd:nd + 3])) <= tol * pm, (n, "momentum")


@pytest.mark.parametrize("name", ["ref1d", "ref2d", "ref3d"])
def test_oracle_against_reference_dump(request, oracle_abi, name):
    d = _dump_dir(request)
    if not d or not os.path.exists(os.path.join(d, f"{name}_0.npz")):
        pytest.skip("no reference dump (tools/ref_build/README.md): EPOCH cannot be built in this container")
    d0, d1 = _load(d, name)
    sim, names, inner = step_from_dump(oracle_abi, d0, d1)
    compare_with_dump(sim, names, inner, d1)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["ref1d", "ref2d", "ref3d"])
def test_cuda_against_reference_dump(request, product_abi, name):
    d = _dump_dir(request)
    if not d or not os.path.exists(os.path.join(d, f"{name}_0.npz")):
        pytest.skip("no reference dump (tools/ref_build/README.md): EPOCH cannot be built in this container")
    d0, d1 = _load(d, name)
    sim, names, inner = step_from_dump(product_abi, d0, d1)
    compare_with_dump(sim, names, inner, d1)


def _oracle_dump(oracle_abi, tmp, nd):
    """a dump pair in the reference-dump format, written by the ORACLE (exercises the harness only)"""
    from helpers import make_plasma
    n = [16, 12, 8][:nd]
    sim = make_plasma(oracle_abi, nd, "triangle", n, seed=5, ppc=4)
    sim.start()
    ng = sim.ctx.ng
    inner = tuple(slice(ng, -ng) for _ in range(nd))

    def dump(i):
        f = sim.ctx.download_fields()
        out = {"time": np.float64(sim.time), "step": np.int64(sim.step), "species": np.array(["electron", "proton"]),
               "cells": np.array(n), "edges": np.array([[sim.mins[a], sim.maxs[a]] for a in range(nd)])}
        for k in ALL9:
            out[k] = f[k][inner].T if nd > 1 else f[k][inner]
        for isp, name in enumerate(("electron", "proton")):
            P = sim.ctx.download_particles(isp)
            for a in range(nd):
                out[f"{name}_{'xyz'[a]}"] = P[:, a]
            for a in range(3):
                out[f"{name}_p{'xyz'[a]}"] = P[:, nd + a]
            out[f"{name}_w"] = P[:, nd + 3]
        np.savez(os.path.join(tmp, f"ref{nd}d_{i}.npz"), **out)
    dump(0)
    dt = sim.dt
    sim.step_once()
    dump(1)
    sim.ctx.close()
    return dt


@pytest.mark.parametrize("nd", [1, 2, 3])
def test_ref_dump_harness_selfcheck(oracle_abi, tmp_path, nd):
    _oracle_dump(oracle_abi, str(tmp_path), nd)
    d0, d1 = _load(str(tmp_path), f"ref{nd}d")
    sim, names, inner = step_from_dump(oracle_abi, d0, d1)
    compare_with_dump(sim, names, inner, d1, tol=1e-13)
    sim.ctx.close()
