"""general_dist_fn (io/dist_fn.F90): the oracle's particle loop against an independent numpy restatement
(not gpu), the CUDA kernel against the oracle (gpu)."""
import numpy as np
import pytest

from helpers import make_plasma
from epoch_b200.driver import C

CASES = [
    (1, ["x", "px"], [(0.0, 0.0), (-3e-22, 3e-22)], [1, 40], None),
    (1, ["en"], [(0.0, 0.0)], [25], None),                                         # empty range -> -1 .. 1
    (2, ["x", "y"], [(3e-6, 9.5e-6), (0.0, 0.0)], [1, 1], {"px": (0.0, 1e-20)}),
    (2, ["xy_angle", "mod_p"], [(-np.pi, np.pi), (0.0, 4e-22)], [36, 20], None),
    (2, ["gamma_m1"], [(0.0, 0.5)], [50], {"y": (2e-6, 7e-6), "pz": (-1e-22, 1e-22)}),
    (3, ["px", "py", "pz"], [(-2e-22, 2e-22)] * 3, [12, 10, 8], None),
    (3, ["z", "yz_angle"], [(0.0, 0.0), (0.0, 2.5 * np.pi)], [1, 30], {"en": (0.0, 1e-13)}),
    (3, ["zx_angle"], [(-np.pi, np.pi)], [24], {"x": (1e-6, 1e-5), "gamma_m1": (0.01, 10.0)}),
]


def numpy_dist_fn(sim, P, mass, direction, ranges, resolution, dgrid, restrictions):
    """vectorised restatement of dist_fn.F90:543-628 for one species"""
    nd = sim.ndims
    px, py, pz, w = P[:, nd], P[:, nd + 1], P[:, nd + 2], P[:, nd + 3]
    mc = mass * C
    u2 = (px / mc) ** 2 + (py / mc) ** 2 + (pz / mc) ** 2
    gm1 = u2 / (np.sqrt(u2 + 1.0) + 1.0)
    data = {"x": P[:, 0], "px": px, "py": py, "pz": pz, "en": gm1 * mc * C, "gamma_m1": gm1,
            "mod_p": np.sqrt(px ** 2 + py ** 2 + pz ** 2)}
    if nd > 1:
        data["y"] = P[:, 1]
    if nd > 2:
        data["z"] = P[:, 2]
    keep = np.ones(len(P), dtype=bool)
    for name, (a0, a1) in (("xy_angle", (px, py)), ("yz_angle", (py, pz)), ("zx_angle", (pz, px))):
        if name in direction:
            amax = ranges[direction.index(name)][1]
            p = np.sqrt(a0 ** 2 + a1 ** 2)
            keep &= p >= 2.2250738585072014e-308
            th = np.arcsin(a1 / np.where(p > 0, p, 1.0))
            th = np.where(a0 < 0.0, np.copysign(1.0, a1) * np.pi - th, th)
            th = np.where(th + 2 * np.pi < amax, th + 2 * np.pi, th)
            data[name] = th
    for name, (lo, hi) in (restrictions or {}).items():
        v = data.get(name, np.zeros(len(P)))
        keep &= ~((v < lo) | (v > hi))
    idx = np.zeros(len(P), dtype=np.int64)
    mul = 1
    for name, (lo, _), res, dg in zip(direction, ranges, resolution, dgrid):
        cell = np.floor((data[name] - lo) / dg).astype(np.int64) + 1
        keep &= (cell >= 1) & (cell <= res)
        idx += mul * (np.clip(cell, 1, res) - 1)
        mul *= res
    out = np.zeros(mul)
    np.add.at(out, idx[keep], w[keep])
    return out.reshape(tuple(reversed(resolution)))


def resolved(sim, direction, ranges_in, resolution_in):
    """what general_dist_fn hands to the loop (ranges, resolution, dgrid) -- recomputed from its output grids"""
    arr, grids = sim.general_dist_fn(0, direction, ranges_in, resolution_in)
    res = [len(g) for g in grids]
    dg = [(g[1] - g[0]) if len(g) > 1 else None for g in grids]
    return arr, grids, res, dg


def build(abi, nd):
    n = [20, 12, 8]
    return make_plasma(abi, nd, "triangle", n, seed=3, ppc=6, u_th=0.3)


@pytest.mark.parametrize("nd,direction,ranges_in,resolution_in,restr", CASES)
def test_oracle_dist_fn_against_numpy(oracle_abi, nd, direction, ranges_in, resolution_in, restr):
    sim = build(oracle_abi, nd)
    arr, grids = sim.general_dist_fn(0, direction, ranges_in, resolution_in, restr)
    # the resolved ranges from the grids: grid(i) = lo + (i - 1/2) dgrid
    res = [len(g) for g in grids]
    dgrid, ranges = [], []
    for name, g, (r1, r2), r_in in zip(direction, grids, ranges_in, resolution_in):
        if name in ("x", "y", "z"):
            dg = sim.d["xyz".index(name)]
        else:
            lo, hi = (r1, r2) if abs(r1 - r2) > 0 else (-1.0, 1.0)
            dg = (hi - lo) / r_in
        dgrid.append(dg)
        ranges.append((g[0] - 0.5 * dg, g[0] - 0.5 * dg + len(g) * dg))
    # the angle wrap uses the range's upper end as handed over by the host
    for i, name in enumerate(direction):
        if name.endswith("_angle"):
            ranges[i] = (ranges_in[i][0], ranges_in[i][1])
    P = sim.ctx.download_particles(0)
    ref = numpy_dist_fn(sim, P, sim.species[0].mass, list(direction), ranges, res, dgrid, restr)
    assert arr.shape == ref.shape
    scale = max(ref.max(), 1e-300)
    assert np.max(np.abs(arr - ref)) <= 1e-12 * scale
    assert ref.sum() > 0.0
    if all(n in ("x", "y", "z") for n in direction) and not restr and all(abs(a - b) == 0 for a, b in ranges_in):
        assert abs(arr.sum() - P[:, -1].sum()) <= 1e-12 * P[:, -1].sum()      # every particle lands in a cell


@pytest.mark.gpu
@pytest.mark.parametrize("nd,direction,ranges_in,resolution_in,restr", CASES)
def test_cuda_dist_fn_against_oracle(product_abi, oracle_abi, nd, direction, ranges_in, resolution_in, restr):
    sims = [build(a, nd) for a in (product_abi, oracle_abi)]
    for s in sims:
        for _ in range(2):
            s.step_once()               # (a deferred permute is settled by the call)
    a, ga = sims[0].general_dist_fn(0, direction, ranges_in, resolution_in, restr)
    b, gb = sims[1].general_dist_fn(0, direction, ranges_in, resolution_in, restr)
    assert a.shape == b.shape and all(np.array_equal(x, y) for x, y in zip(ga, gb))
    # bins at the edge of a cell can differ by one particle when positions differ in the last bit: compare sums
    # exactly where the particle sets agree, else within one particle weight per bin
    wmax = sims[1].ctx.download_particles(0)[:, -1].max()
    assert np.max(np.abs(a - b)) <= 1e-12 * max(b.max(), 1e-300) or np.max(np.abs(a - b)) <= 2.0 * wmax
    assert abs(a.sum() - b.sum()) <= 2.0 * wmax + 1e-12 * b.sum()
