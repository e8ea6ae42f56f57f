"""BASELINE.json configs[3] and configs[4] on N GPUs of one node with the decompositions SURVEY 8(e) names
(they are parity-test cases, not bench lines; launched with torchrun, one rank per GPU):
  cfg4: epoch3d laser wakefield, 2048 x 256 x 256 cells, 8 electrons per cell, bspline3, moving window at c --
        y x z pencils (nprocdir 1 x 2 x 4 on eight GPUs): the window moves along x, so no rank ever hands its
        plasma to another because of the shift (housekeeping/window.F90:388-436); every rank removes the particles
        behind x_min, shifts its own x lines and inserts the incoming column of its own y-z range
  cfg5: epoch2d laser-solid foil, 4096 x 2048 cells, a 32-cell slab with 256 e- + 256 p+ per cell, vacuum elsewhere --
        cut along y, the foil's transverse axis, and re-balanced every 4 steps the way balance_workload does
        (get_load -> calculate_breaks -> redistribute_domain; housekeeping/balance.F90:107-319,2526-2669)
One JSON line per config from rank 0: device time (max over ranks), per-phase ms (max over ranks), particles per
rank (max / mean), the size-independent checks of tools/run_configs.py summed over the ranks.
usage: python -m torch.distributed.run --nproc-per-node N tools/run_configs_mgpu.py cfg4|cfg5|both [scale]"""
import json
import math
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from epoch_b200 import _abi                                                    # noqa: E402
from epoch_b200.driver import Simulation, SpeciesSpec, LaserSpec, Q0, M0, C     # noqa: E402
from run_configs import load_uniform, PHASES, KB                                # noqa: E402

RANK = int(os.environ.get("RANK", "0"))
WORLD = int(os.environ.get("WORLD_SIZE", "1"))
LOCAL = int(os.environ.get("LOCAL_RANK", "0"))


def nccl_id(abi):
    if WORLD == 1:
        return None
    import ctypes
    buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if RANK == 0:
        raw = (ctypes.c_char * 128)()
        assert abi.nccl_unique_id(raw) == 0
        buf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
    dist.broadcast(buf, 0)
    return bytes(buf.cpu().numpy().tobytes())


def red(x, op="max"):
    if WORLD == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "sum": dist.ReduceOp.SUM}[op])
    return float(t.item())


def allreduce_np(a):
    if WORLD == 1:
        return a
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    dist.all_reduce(t)
    return t.cpu().numpy()


def summary(sim, name, npart0, steps, dev_ms, wall_s, extra):
    counts = [sim.ctx.particle_count(i) for i in range(len(sim.species))]
    n_rank = float(sum(counts))
    out = {"config": name, "n_gpus": WORLD, "nprocdir": list(sim.nprocdir), "cells_per_rank": list(sim.n_local),
           "particles": int(npart0), "steps": steps,
           "ms_per_step": red(dev_ms) / steps, "wall_ms_per_step": 1e3 * red(wall_s) / steps,
           "particles_per_rank_max": red(n_rank), "particles_per_rank_mean": red(n_rank, "sum") / WORLD,
           "particles_after": int(red(n_rank, "sum")),
           "phases_ms_per_step_max_over_ranks": {k: red(sim.ctx.timer(k)[0]) / steps for k in PHASES},
           "checks": {"cell_mismatch": int(red(sim.ctx.timer("cell_mismatch")[1], "sum"))},
           "sorts": {"deferred": sim.ctx.timer("deferred_sorts")[1], "permuted": sim.ctx.timer("permuted_sorts")[1]}}
    out["particle_steps_per_s"] = npart0 * steps / (out["ms_per_step"] * steps * 1e-3)
    fe, ke = sim.ctx.energy()
    out["checks"]["field_energy"] = red(fe, "sum")
    out["checks"]["kinetic_finite"] = bool(red(0.0 if np.all(np.isfinite(ke)) else 1.0) == 0.0)
    out.update(extra)
    return out


def run_cfg4(abi, scale, steps=10, warm=3):
    nx, ny = max(64, int(2048 * scale)), max(32, int(256 * scale))
    lam = 0.8e-6
    dx, dyz = lam / 16.0, lam / 2.0
    dens, ppc = 1.0e24, 8
    npd = {1: (1, 1, 1), 2: (1, 1, 2), 4: (1, 2, 2), 8: (1, 2, 4)}[WORLD]
    pb = ["open", "open", "periodic", "periodic", "periodic", "periodic"]
    sp = SpeciesSpec("electron", -Q0, M0, bcs=pb, density=dens, temperature=(1e4,) * 3)
    las = LaserSpec(1, intensity_w_cm2=2.0e18, lambda_=lam, t_profile=lambda t: math.exp(-((t - 30e-15) / 12e-15) ** 2))
    sim = Simulation(3, [nx, ny, ny], [0.0, 0.0, 0.0], [nx * dx, ny * dyz, ny * dyz], shape="bspline3", species=[sp], lasers=[las],
                     bc_field=["simple_laser", "simple_outflow", "periodic", "periodic", "periodic", "periodic"], abi=abi,
                     device=LOCAL, rank=RANK, nproc=WORLD, nprocdir=npd, nccl_id=nccl_id(abi), options={"timers": 1})
    t0 = time.perf_counter()
    load_uniform(sim, 0, sp, ppc, dens, 1e4, 1 << 23, 200 + RANK)
    t_load = time.perf_counter() - t0
    w = dens * dx * dyz * dyz / ppc
    sigma = math.sqrt(M0 * KB * 1e4)
    rng = np.random.default_rng([7, RANK])
    m = sim.n_local[1] * sim.n_local[2] * ppc
    col = sim.ctx.host_empty((m, 7))
    shifts = [0]

    def insert(s, x0):        # stands for insert_particles (window.F90:198-358): this rank's part of the incoming column
        col[:, 0] = x0 + rng.random(m) * s.d[0]
        for d in (1, 2):
            col[:, d] = rng.random(m) * (s.max_local[d] - s.min_local[d]) + s.min_local[d]
        col[:, 3:6] = rng.standard_normal((m, 3)) * sigma
        col[:, 6] = w
        shifts[0] += 1
        return [col]
    sim.enable_window(v_x=C, insert=insert)
    n0 = red(sim.ctx.particle_count(0), "sum")
    for _ in range(warm):
        sim.step_once()
    sim.ctx.sync()
    for k in PHASES:
        sim.ctx.timer(k, reset=True)
    if WORLD > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sim.ctx.mark("cfg", True)
    tw = time.perf_counter()
    for _ in range(steps):
        sim.step_once()
    sim.ctx.mark("cfg", False)
    sim.ctx.sync()
    wall = time.perf_counter() - tw
    dev_ms = sim.ctx.timer("cfg")[0]
    name = ("cfg4 epoch3d laser wakefield %dx%dx%d cells, 8 e- per cell, bspline3, moving window at c (every rank inserts its "
            "part of the incoming column), laser 2e18 W/cm^2; y x z pencils" % (nx, ny, ny))
    out = summary(sim, name, n0, steps, dev_ms, wall, {"load_seconds": t_load, "window_shifts": shifts[0],
                                                     "window_offset_cells": sim.window_offset / sim.d[0]})
    out["checks"]["count_within_2_percent"] = bool(abs(out["particles_after"] - n0) < 0.02 * n0)
    return out


def run_cfg5(abi, scale, steps=12, warm=3, every=4):
    nx, ny = max(256, int(4096 * scale)), max(128, int(2048 * scale))
    lam = 1.0e-6
    dx = lam / 32.0
    dens, ppc = 5.0e28, 256
    thick = max(8, int(32 * scale))
    pb = ["open", "open", "periodic", "periodic"]
    species = [SpeciesSpec("electron", -Q0, M0, bcs=pb, density=dens, temperature=(1e5,) * 3),
               SpeciesSpec("proton", Q0, 1836.2 * M0, bcs=pb, density=dens, temperature=(1e5,) * 3)]
    las = LaserSpec(1, intensity_w_cm2=1.0e20, lambda_=lam, t_profile=lambda t: min(1.0, t / 10e-15))
    sim = Simulation(2, [nx, ny], [0.0, 0.0], [nx * dx, ny * dx], shape="triangle", species=species, lasers=[las],
                     bc_field=["simple_laser", "simple_outflow", "periodic", "periodic"], abi=abi,
                     device=LOCAL, rank=RANK, nproc=WORLD, nprocdir=(1, WORLD), nccl_id=nccl_id(abi), options={"timers": 1})
    x0 = sim.mins[0] + (nx // 2) * dx
    w = dens * dx * dx / ppc
    # a foil whose areal density varies along y (a factor 3 between its thinnest and thickest part): equal cuts along y
    # are then unequal in particles and calculate_breaks has something to move
    rng = np.random.default_rng([11, RANK])
    t0 = time.perf_counter()
    for isp, sp in enumerate(species):
        cols = []
        for j in range(sim.n_local[1]):
            yg = (sim.cell_min[1] - 1 + j + 0.5) / ny
            nj = int(thick * ppc * (0.5 + 1.0 * yg))              # particles of this row of cells
            Q = np.empty((nj, 6))
            Q[:, 0] = x0 + rng.random(nj) * (thick * dx)
            Q[:, 1] = sim.min_local[1] + (j + rng.random(nj)) * dx
            cols.append(Q)
        P = np.concatenate(cols)
        P[:, 2:5] = rng.standard_normal((len(P), 3)) * math.sqrt(sp.mass * KB * 1e5)
        P[:, 5] = w
        sim.ctx.set_option("reserve", "%d:%d" % (isp, int(len(P) * 1.8) + 4096))
        sim.ctx.upload_particles(isp, P)
    t_load = time.perf_counter() - t0
    n0 = red(sum(sim.ctx.particle_count(i) for i in range(2)), "sum")
    n_rank0 = red(sum(sim.ctx.particle_count(i) for i in range(2)))
    for _ in range(warm):
        sim.step_once()
    sim.ctx.sync()
    for k in PHASES:
        sim.ctx.timer(k, reset=True)
    if WORLD > 1:
        dist.barrier()
    torch.cuda.synchronize()
    dev_ms, wall, t_bal, moved = 0.0, 0.0, 0.0, 0
    done = 0
    while done < steps:
        sim.ctx.mark("cfg", True)
        tw = time.perf_counter()
        for _ in range(min(every, steps - done)):
            sim.step_once()
            done += 1
        sim.ctx.mark("cfg", False)
        sim.ctx.sync()
        wall += time.perf_counter() - tw
        if WORLD > 1 and done < steps:
            tb = time.perf_counter()
            moved += 1 if sim.balance_workload(allreduce_np) else 0
            sim.ctx.sync()
            t_bal += time.perf_counter() - tb
    dev_ms = sim.ctx.timer("cfg")[0]
    name = ("cfg5 epoch2d laser-foil %dx%d cells, foil %d cells thick with ~256 e- + 256 p+ per cell (areal density x 0.5 .. 1.5 "
            "along y; vacuum elsewhere), triangle, laser 1e20 W/cm^2; cut along y, re-balanced every %d steps" % (nx, ny, thick, every))
    out = summary(sim, name, n0, steps, dev_ms, wall, {"load_seconds": t_load, "rebalances_that_moved_a_break": moved,
                                                     "rebalance_seconds_total": red(t_bal),
                                                     "particles_per_rank_max_before": n_rank0,
                                                     "cell_maxs_y": [int(v) for v in sim.cell_maxs[1]]})
    out["checks"]["none_lost"] = bool(out["particles_after"] == int(n0))
    return out


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "both"
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
    torch.cuda.set_device(LOCAL)
    if WORLD > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", LOCAL))
    abi = _abi.product()
    for name, fn in (("cfg5", run_cfg5), ("cfg4", run_cfg4)):
        if which in (name, "both"):
            res = fn(abi, scale)
            if RANK == 0:
                print(json.dumps(res), flush=True)
    if WORLD > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
