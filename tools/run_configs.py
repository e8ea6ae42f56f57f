"""BASELINE.json configs[1] and configs[3] at full size on ONE B200 (they are parity-test cases, not
bench lines): throughput of the full step plus the size-independent checks that the path offers --
no particle lost (periodic axes) / only removed at open boundaries, key == cell for every particle
(cell_mismatch = 0), finite energies, laser energy entering the box.
  cfg2: epoch2d laser-plasma, 4096 x 4096 cells, 32 electrons + 32 protons per cell (64 ppc), triangle
  cfg4: epoch3d laser wakefield, 2048 x 256 x 256 cells, 8 electrons per cell, bspline3, moving window
usage: python tools/run_configs.py [cfg2|cfg4|both] [scale]   (scale < 1 shrinks the grid for a smoke run)
One JSON line per config on stdout."""
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from epoch_b200 import _abi                                                    # noqa: E402
from epoch_b200.driver import Simulation, SpeciesSpec, LaserSpec, Q0, M0, C     # noqa: E402

KB = 1.3806488e-23


def load_uniform(sim, isp, sp, ppc, density, t_kelvin, chunk_cells, seed):
    """uniform plasma, `ppc` per cell, thermal momenta; uploaded in chunks of whole slabs of the slowest axis"""
    nd = sim.ndims
    n = sim.n_local
    slow = n[nd - 1]
    per_slab = int(np.prod(n[:nd - 1])) * ppc
    slabs = max(1, min(slow, chunk_cells // max(1, per_slab // ppc)))
    w = density * float(np.prod(sim.d)) / ppc
    sigma = math.sqrt(sp.mass * KB * t_kelvin)
    sim.ctx.set_option("reserve", "%d:%d" % (isp, int(slow * per_slab * 1.03) + 4096))     # one allocation (+ room for inserts)
    buf = sim.ctx.host_empty((slabs * per_slab, nd + 4))
    rng = np.random.default_rng(seed)
    first = True
    for s0 in range(0, slow, slabs):
        s1 = min(slow, s0 + slabs)
        m = (s1 - s0) * per_slab
        P = buf[:m]
        for d in range(nd - 1):
            P[:, d] = rng.random(m) * (sim.max_local[d] - sim.min_local[d]) + sim.min_local[d]
        lo = sim.min_local[nd - 1] + s0 * sim.d[nd - 1]
        P[:, nd - 1] = rng.random(m) * ((s1 - s0) * sim.d[nd - 1]) + lo
        P[:, nd:nd + 3] = rng.standard_normal((m, 3)) * sigma
        P[:, nd + 3] = w
        sim.ctx.upload_particles(isp, P, append=not first)
        first = False


PHASES = ("push_A", "push_B", "sort", "fields", "bcs", "halo", "window")


def timed_steps(sim, warm, steps):
    for _ in range(warm):
        sim.step_once()
    sim.ctx.sync()
    for k in PHASES:                      # the phase timers count from here (not from the context's creation)
        sim.ctx.timer(k, reset=True)
    sim.ctx.mark("cfg", True)
    t0 = time.perf_counter()
    for _ in range(steps):
        sim.step_once()
    sim.ctx.mark("cfg", False)
    sim.ctx.sync()
    wall = time.perf_counter() - t0
    return sim.ctx.timer("cfg")[0] * 1e-3, wall


def run_cfg2(abi, scale, steps=8, warm=3):
    n = max(64, int(4096 * scale))
    lam = 1.0e-6
    dx = lam / 20.0
    dens = 5.0e26                                   # ~0.45 n_c at 1 um
    species = [SpeciesSpec("electron", -Q0, M0, bcs=["open", "open", "periodic", "periodic"], density=dens, temperature=(1e6,) * 3),
               SpeciesSpec("proton", Q0, 1836.2 * M0, bcs=["open", "open", "periodic", "periodic"], density=dens, temperature=(1e6,) * 3)]
    las = LaserSpec(1, intensity_w_cm2=1.0e18, lambda_=lam, t_profile=lambda t: min(1.0, t / 10e-15))
    sim = Simulation(2, [n, n], [0.0, 0.0], [n * dx, n * dx], shape="triangle", species=species, lasers=[las],
                     bc_field=["simple_laser", "simple_outflow", "periodic", "periodic"], abi=abi,
                     options={"timers": 1})
    t0 = time.perf_counter()
    for isp, sp in enumerate(species):
        load_uniform(sim, isp, sp, 32, dens, 1e6, 1 << 20, 100 + isp)
    t_load = time.perf_counter() - t0
    n0 = [sim.ctx.particle_count(i) for i in range(2)]
    dev_s, wall = timed_steps(sim, warm, steps)
    n1 = [sim.ctx.particle_count(i) for i in range(2)]
    fe, ke = sim.ctx.energy()
    npart = sum(n0)
    return {"config": "cfg2 epoch2d laser-plasma %dx%d cells, 32 e- + 32 p+ per cell, triangle, laser 1e18 W/cm^2 at x_min, "
                      "open x / periodic y, one B200" % (n, n),
            "particles": npart, "steps": steps, "ms_per_step": 1e3 * dev_s / steps, "wall_ms_per_step": 1e3 * wall / steps,
            "particle_steps_per_s": npart * steps / dev_s, "load_seconds": t_load,
            "checks": {"cell_mismatch": sim.ctx.timer("cell_mismatch")[1], "particles_before": n0, "particles_after": n1,
                       "only_removed": all(b <= a for a, b in zip(n0, n1)), "field_energy_J_per_m": fe,
                       "kinetic_finite": bool(np.all(np.isfinite(ke))), "laser_energy_entered": bool(fe > 0.0)},
            "sorts": {"deferred": sim.ctx.timer("deferred_sorts")[1], "permuted": sim.ctx.timer("permuted_sorts")[1]},
            "phases_ms_per_step": {k: sim.ctx.timer(k)[0] / steps for k in PHASES}}


def run_cfg4(abi, scale, steps=10, warm=3):
    nx, ny = max(64, int(2048 * scale)), max(32, int(256 * scale))
    lam = 0.8e-6
    dx, dyz = lam / 16.0, lam / 2.0
    dens, ppc = 1.0e24, 8
    sp = SpeciesSpec("electron", -Q0, M0, bcs=["open", "open", "periodic", "periodic", "periodic", "periodic"], density=dens,
                     temperature=(1e4,) * 3)
    las = LaserSpec(1, intensity_w_cm2=2.0e18, lambda_=lam, t_profile=lambda t: math.exp(-((t - 30e-15) / 12e-15) ** 2))
    sim = Simulation(3, [nx, ny, ny], [0.0, 0.0, 0.0], [nx * dx, ny * dyz, ny * dyz], shape="bspline3", species=[sp], lasers=[las],
                     bc_field=["simple_laser", "simple_outflow", "periodic", "periodic", "periodic", "periodic"], abi=abi,
                     options={"timers": 1})
    t0 = time.perf_counter()
    load_uniform(sim, 0, sp, ppc, dens, 1e4, 1 << 23, 200)
    t_load = time.perf_counter() - t0
    w = dens * dx * dyz * dyz / ppc
    sigma = math.sqrt(M0 * KB * 1e4)
    rng = np.random.default_rng(7)
    col = sim.ctx.host_empty((ny * ny * ppc, 7))
    shifts = [0]

    def insert(s, x0):        # stands for insert_particles (window.F90:198-358): the incoming cell column
        m = col.shape[0]
        col[:, 0] = x0 + rng.random(m) * s.d[0]
        for d in (1, 2):
            col[:, d] = rng.random(m) * (s.max_local[d] - s.min_local[d]) + s.min_local[d]
        col[:, 3:6] = rng.standard_normal((m, 3)) * sigma
        col[:, 6] = w
        shifts[0] += 1
        return [col]
    sim.enable_window(v_x=C, insert=insert)
    n0 = sim.ctx.particle_count(0)
    dev_s, wall = timed_steps(sim, warm, steps)
    n1 = sim.ctx.particle_count(0)
    fe, ke = sim.ctx.energy()
    return {"config": "cfg4 epoch3d laser wakefield %dx%dx%d cells, 8 e- per cell, bspline3, moving window at c "
                      "(host inserts the incoming column), laser 2e18 W/cm^2, one B200" % (nx, ny, ny),
            "particles": n0, "steps": steps, "ms_per_step": 1e3 * dev_s / steps, "wall_ms_per_step": 1e3 * wall / steps,
            "particle_steps_per_s": n0 * steps / dev_s, "load_seconds": t_load, "window_shifts": shifts[0],
            "window_offset_cells": sim.window_offset / sim.d[0],
            "checks": {"cell_mismatch": sim.ctx.timer("cell_mismatch")[1], "particles_before": n0, "particles_after": n1,
                       "count_within_2_percent": bool(abs(n1 - n0) < 0.02 * n0), "field_energy_J": fe,
                       "kinetic_finite": bool(np.all(np.isfinite(ke))), "laser_energy_entered": bool(fe > 0.0)},
            "sorts": {"deferred": sim.ctx.timer("deferred_sorts")[1], "permuted": sim.ctx.timer("permuted_sorts")[1]},
            "phases_ms_per_step": {k: sim.ctx.timer(k)[0] / steps for k in PHASES}}


def run_cfg5(abi, scale, steps=8, warm=3):
    """configs[4]: 2D laser-solid foil, highly non-uniform particles per cell (a thin dense slab, vacuum elsewhere)"""
    nx, ny = max(256, int(4096 * scale)), max(128, int(2048 * scale))
    lam = 1.0e-6
    dx = lam / 32.0
    dens, ppc = 5.0e28, 256                      # ~45 n_c, 256 e- + 256 p+ per cell inside the foil
    thick = max(8, int(32 * scale))
    species = [SpeciesSpec("electron", -Q0, M0, bcs=["open", "open", "periodic", "periodic"], density=dens, temperature=(1e5,) * 3),
               SpeciesSpec("proton", Q0, 1836.2 * M0, bcs=["open", "open", "periodic", "periodic"], density=dens, temperature=(1e5,) * 3)]
    las = LaserSpec(1, intensity_w_cm2=1.0e20, lambda_=lam, t_profile=lambda t: min(1.0, t / 10e-15))
    sim = Simulation(2, [nx, ny], [0.0, 0.0], [nx * dx, ny * dx], shape="triangle", species=species, lasers=[las],
                     bc_field=["simple_laser", "simple_outflow", "periodic", "periodic"], abi=abi, options={"timers": 1})
    x0 = sim.min_local[0] + (nx // 2) * dx
    w = dens * dx * dx / ppc
    m = thick * ny * ppc
    rng = np.random.default_rng(11)
    t0 = time.perf_counter()
    for isp, sp in enumerate(species):
        P = sim.ctx.host_empty((m, 6))
        P[:, 0] = x0 + rng.random(m) * (thick * dx)
        P[:, 1] = rng.random(m) * (sim.max_local[1] - sim.min_local[1]) + sim.min_local[1]
        P[:, 2:5] = rng.standard_normal((m, 3)) * math.sqrt(sp.mass * KB * 1e5)
        P[:, 5] = w
        sim.ctx.upload_particles(isp, P)
    t_load = time.perf_counter() - t0
    n0 = [sim.ctx.particle_count(i) for i in range(2)]
    dev_s, wall = timed_steps(sim, warm, steps)
    n1 = [sim.ctx.particle_count(i) for i in range(2)]
    fe, ke = sim.ctx.energy()
    npart = sum(n0)
    return {"config": "cfg5 epoch2d laser-foil %dx%d cells, foil %d cells thick with 256 e- + 256 p+ per cell (vacuum elsewhere), "
                      "triangle, laser 1e20 W/cm^2, one B200" % (nx, ny, thick),
            "particles": npart, "steps": steps, "ms_per_step": 1e3 * dev_s / steps,
            "particle_steps_per_s": npart * steps / dev_s, "load_seconds": t_load,
            "checks": {"cell_mismatch": sim.ctx.timer("cell_mismatch")[1], "particles_before": n0, "particles_after": n1,
                       "none_lost": n0 == n1, "field_energy_J_per_m": fe, "kinetic_finite": bool(np.all(np.isfinite(ke)))},
            "sorts": {"deferred": sim.ctx.timer("deferred_sorts")[1], "permuted": sim.ctx.timer("permuted_sorts")[1]},
            "phases_ms_per_step": {k: sim.ctx.timer(k)[0] / steps for k in PHASES}}


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "both"
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    abi = _abi.product()
    if which in ("cfg2", "both"):
        print(json.dumps(run_cfg2(abi, scale)), flush=True)
    if which in ("cfg4", "both"):
        print(json.dumps(run_cfg4(abi, scale)), flush=True)
    if which == "cfg5":
        print(json.dumps(run_cfg5(abi, scale)), flush=True)
