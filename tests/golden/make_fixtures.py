"""Writes the committed fixtures of tests/golden/.

reference_goldens.json is typed in from the reference's own test sources (file:line in the file);
nothing here can run the reference (Fortran + MPI, not buildable in this image).

push_deposit_*.npz are produced by the CPU ORACLE (oracle/liborc.so), not by the reference: seeded
particle loads and fields in, particles and E/B/J after two full steps out.  The reference has no test
that pins a particle-dependent number, so these fixtures pin the ORACLE's behaviour (and, through
tests/test_golden.py, the CUDA path's) -- "parity unpinned by the reference" as DESIGN.md §5 says.

usage: python tests/golden/make_fixtures.py      (needs oracle/liborc.so: make -C oracle)"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

CASES = [(1, "triangle", [24]), (2, "tophat", [12, 10]), (2, "bspline3", [12, 10]),
         (3, "triangle", [10, 8, 6]), (3, "bspline3", [8, 8, 6])]
SEED, PPC, STEPS = 424242, 3, 2


def build(abi, ndims, shape, n):
    from helpers import make_plasma
    return make_plasma(abi, ndims, shape, n, seed=SEED, ppc=PPC, u_th=0.4, e_amp=1e10, b_amp=30.0)


def run(sim):
    from helpers import ALL9
    for _ in range(STEPS):
        sim.step_once()
    out = {}
    for isp in range(len(sim.species)):
        P = sim.ctx.download_particles(isp)
        out["part%d" % isp] = P[np.argsort(P[:, -1], kind="stable")]       # weights are unique: canonical order
    f = sim.ctx.download_fields(ALL9)
    for k in ALL9:
        out[k] = f[k]
    return out


def name(ndims, shape):
    return os.path.join(HERE, "push_deposit_%dd_%s.npz" % (ndims, shape))


if __name__ == "__main__":
    from epoch_b200 import _abi
    orc = _abi.Abi(os.path.join(HERE, "..", "..", "oracle", "liborc.so"), "orc_")
    for ndims, shape, n in CASES:
        sim = build(orc, ndims, shape, n)
        inp = {"load%d" % i: L for i, L in enumerate(sim.loads)}
        inp.update({"init_" + k: v for k, v in sim.init_fields.items()})
        out = run(sim)
        np.savez_compressed(name(ndims, shape), dt=sim.dt, **inp, **out)
        print(name(ndims, shape), os.path.getsize(name(ndims, shape)))
