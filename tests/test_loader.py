"""The reference's random generator and particle loader as restated in oracle/orc_loader.cpp
(housekeeping/random_generator.f90:45-173, user_interaction/helper.F90 load_particles /
setup_particle_density, particle_temperature.F90:30-85,380-390), cross-checked against an independent
pure-Python restatement written from the Fortran (32-bit two's-complement arithmetic made explicit).
The reference holds no vector of its generator, so this pins the C++ against a second reading, not
against EPOCH output ("parity unpinned")."""
import ctypes as C
import math

import numpy as np
import pytest

from helpers import Simulation, SpeciesSpec, Q0, M0, reference_load

M32 = 0xFFFFFFFF


def _i32(v):
    v &= M32
    return v - (1 << 32) if v & 0x80000000 else v


class Kiss:
    """random_generator.f90 with Fortran default INTEGER semantics: wrap-around products and sums,
    ISHFT(a, n < 0) = logical right shift of the 32-bit pattern"""

    def __init__(self, seed):
        self.x, self.y, self.z, self.w = [(v + seed) & M32 for v in (123456789, 362436069, 521288629, 916191069)]
        self.cached, self.cval = False, 0.0
        for _ in range(1000):
            self.random()

    def random(self):
        self.x = (69069 * self.x + 1327217885) & M32
        a2 = self.y ^ ((self.y << 13) & M32)
        a3 = a2 ^ (a2 >> 17)
        self.y = a3 ^ ((a3 << 5) & M32)
        self.z = (18000 * (self.z & 65535) + (self.z >> 16)) & M32
        self.w = (30903 * (self.w & 65535) + (self.w >> 16)) & M32
        kiss = (self.x + self.y + ((self.z << 16) & M32) + self.w) & M32
        return (float(_i32(kiss)) + 2147483648.0) / 4294967296.0

    def box_muller(self, stdev, mu=0.0):
        if self.cached:
            self.cached = False
            return self.cval * stdev + mu
        self.cached = True
        while True:
            r1 = 2.0 * self.random() - 1.0
            r2 = 2.0 * self.random() - 1.0
            w = r1 * r1 + r2 * r2
            if 2.2250738585072014e-308 < w < 1.0:
                break
        w = math.sqrt((-2.0 * math.log(w)) / w)
        self.cval = r2 * w
        return r1 * w * stdev + mu


def test_kiss_and_box_muller(oracle_abi):
    lib = oracle_abi.lib
    lib.orc_random.restype = C.c_double
    lib.orc_random_box_muller.restype = C.c_double
    lib.orc_random_box_muller.argtypes = [C.c_double, C.c_double]
    lib.orc_random_init.argtypes = [C.c_int32]
    for seed in (7842432, 7842433, 0, -5):
        k = Kiss(seed)
        lib.orc_random_init(seed)
        a = [lib.orc_random() for _ in range(3000)]
        b = [k.random() for _ in range(3000)]
        assert a == b
        assert 0.0 <= min(a) and max(a) < 1.0
        g = [lib.orc_random_box_muller(2.5, -1.0) for _ in range(2001)]      # odd count: leaves a cached value
        h = [k.box_muller(2.5, -1.0) for _ in range(2001)]
        assert g == h
        assert lib.orc_random() == k.random()
    u = np.array(a)
    assert abs(u.mean() - 0.5) < 0.02 and abs(u.var() - 1.0 / 12.0) < 0.01
    gg = np.array(g)
    assert abs(gg.mean() + 1.0) < 0.2 and abs(gg.std() / 2.5 - 1.0) < 0.06


@pytest.mark.parametrize("ndims,shape", [(1, "triangle"), (2, "tophat"), (2, "triangle"), (3, "triangle"), (3, "bspline3")])
def test_auto_load(oracle_abi, ndims, shape):
    """load order (cells in Fortran order, ppc each, x then y then z from the stream), weights = n dV / ppc to
    rounding, momenta drawn direction by direction over the whole list"""
    n = [12, 6, 5][:ndims]
    d = 2.0e-6
    ppc, dens, temp, drift = 3, 4.0e24, (1.0e5, 0.0, 3.0e5), (0.0, 1.0e-24, 0.0)
    sp = [SpeciesSpec("e", -Q0, M0, density=dens, temperature=temp)]
    sim = Simulation(ndims, n, [0.0] * ndims, [d * x for x in n], shape=shape, species=sp, abi=oracle_abi)
    P = reference_load(oracle_abi, sim, [(ppc, dens, temp, drift)])[0]
    ncell = int(np.prod(n))
    assert P.shape == (ncell * ppc, ndims + 4)
    k = Kiss(7842432)
    # positions: the first 3 * ndims numbers of the stream land in cell (1, 1, 1)
    for ip in range(ncell * ppc):
        cell = ip // ppc
        for a in range(ndims):
            ia = (cell // int(np.prod(n[:a]))) % n[a]
            x = (0.0 + d / 2.0) + ia * d + (k.random() - 0.5) * d
            assert P[ip, a] == pytest.approx(x, abs=1e-18, rel=1e-14)
    assert np.allclose(P[:, ndims + 3], dens * d ** ndims / ppc, rtol=1e-13)
    kb = 1.3806488e-23
    for a in range(3):
        for ip in range(ncell * ppc):
            v = k.box_muller(math.sqrt(temp[a] * kb * M0), drift[a])
            assert P[ip, ndims + a] == pytest.approx(v, rel=1e-12, abs=1e-40)
    sim.ctx.close()
