#!/usr/bin/env python
"""bench.py -- throughput of the EPOCH PIC hot path on B200 (BASELINE.json metric).

Workload (configs[2], the one the metric's target is quoted on): epoch3d uniform
thermal plasma, 512^3 cells x 16 ppc electrons (neutralising background implied,
deck_control_block.F90:519-526), periodic, triangle shape, fp64; n = 1e24 m^-3,
T = 1 keV, dx = lambda_D / 2, dt = 0.95 x CFL.  Synthetic data: the
particle load follows the reference loader's rule (16 particles per cell at
x(ix) + (U - 1/2) dx, helper.F90:599-601; Maxwellian momenta sigma = sqrt(m kB T),
particle_temperature.F90:397-409) with numpy's generator instead of KISS.

A "step" is one iteration of PROGRAM pic's main loop as far as the hot path goes
(epoch3d.F90:184-249): update_eb_fields_half, push_particles (push + deposit +
particle_bcs + re-sort), current_finish, update_eb_fields_final -- all of it inside
the timed region.  `value` = particles x steps / device time (CUDA events on the
library's stream, max over ranks) with the state resident in HBM; `e2e` = the same
through the C ABI from HOST buffers: upload of all particles and fields, K steps,
download of all particles and fields, wall clock (the context with its NCCL
communicator is created before the clock starts and timed apart).

N > 1 (torchrun): the line printed is STRONG scaling -- the same global 512^3 box cut into N slabs
of 512 x 512 x 512/N along z (`"scaling": "strong"`, e2e on that same deck) -- and a key `weak` holds
the device-timed run with one 512^3 slab per GPU stacked along z (--scaling weak prints that one as
the line instead).  NCCL send/recv of field halos, current ghost sums and migrating particles.
Every line carries `parity_check`: global particle count, total-energy drift and the Gauss-law
residual (max over ranks), taken outside the timed region; a failed check makes the run exit 3.

`--impl reference` times the CPU restatement of the reference (oracle/, built -O3)
on all host cores, one independent periodic box per process.
"""
import argparse
import json
import math
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

Q0, M0, C0, KB, EPS0 = 1.602176565e-19, 9.10938291e-31, 2.99792458e8, 1.3806488e-23, 8.854187817620389850536563031710750e-12
DENSITY = 1.0e24                   # m^-3
T_KELVIN = 1.0e3 * Q0 / KB         # 1 keV
PPC = 16
ORACLE_FAST = os.path.join(ROOT, "oracle", "liborc_fast.so")


def deck_geometry(cells):
    lam_d = math.sqrt(EPS0 * KB * T_KELVIN / (DENSITY * Q0 * Q0))
    dx = lam_d / 2.0
    return dx, [c * dx for c in cells]


SHAPE = "triangle"      # --shape (tuning / comparison runs; the bench line is quoted on triangle)


def make_sim(abi, cells_global, rank=0, nproc=1, nccl_id=None, device=0):
    from epoch_b200.driver import Simulation, SpeciesSpec
    dx, length = deck_geometry(cells_global)
    sp = SpeciesSpec("electron", -Q0, M0, density=DENSITY, temperature=(T_KELVIN,) * 3)
    return Simulation(3, list(cells_global), [0.0, 0.0, 0.0], length, shape=SHAPE, species=[sp],
                      abi=abi, device=device, rank=rank, nproc=nproc, nprocdir=(1, 1, nproc), nccl_id=nccl_id)


def fill_particles(out, sim, seed, nthreads, k_begin=0, k_end=None):
    """Reference loader rule for a uniform plasma on this rank's cells (z planes
    k_begin..k_end), written into `out` ((planes*nx*ny*PPC, 7) float64, partlist wire
    order x,y,z,px,py,pz,w)."""
    nx, ny, nz = sim.n_local
    k_end = nz if k_end is None else k_end
    dx = sim.d[0]
    weight = DENSITY * dx ** 3 / PPC
    sigma = math.sqrt(M0 * KB * T_KELVIN)
    per_plane = nx * ny * PPC
    cx = sim.grid_min_local[0] + np.arange(nx) * sim.d[0]
    cy = sim.grid_min_local[1] + np.arange(ny) * sim.d[1]

    def work(k0, k1):
        for k in range(k0, k1):
            rng = np.random.default_rng([seed, sim.cell_min[2] - 1 + k])     # by GLOBAL plane: the same load however the box is cut
            P = out[(k - k_begin) * per_plane:(k - k_begin + 1) * per_plane]
            u = rng.random((per_plane, 3))
            u -= 0.5
            cells = np.arange(nx * ny).repeat(PPC)
            P[:, 0] = cx[cells % nx] + u[:, 0] * dx
            P[:, 1] = cy[cells // nx] + u[:, 1] * dx
            P[:, 2] = (sim.grid_min_local[2] + k * sim.d[2]) + u[:, 2] * dx
            P[:, 3:6] = rng.standard_normal((per_plane, 3)) * sigma
            P[:, 6] = weight
    nthreads = max(1, min(nthreads, k_end - k_begin))
    bounds = np.linspace(k_begin, k_end, nthreads + 1).astype(int)
    ts = [threading.Thread(target=work, args=(bounds[i], bounds[i + 1])) for i in range(nthreads)]
    [t.start() for t in ts]
    [t.join() for t in ts]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device=0):
        self.proc = None
        self.device = device

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------
# CPU arm: the oracle port (CPU restatement of the reference), one independent
# periodic box per process.  Only bench.py's cpu legs and tests/ may load oracle/.
# ---------------------------------------------------------------------------
def _cpu_worker(task):
    cells, steps, warmup, seed = task
    from epoch_b200 import _abi
    abi = _abi.Abi(ORACLE_FAST, "orc_")
    sim = make_sim(abi, (cells, cells, cells))
    n = cells ** 3 * PPC
    P = np.empty((n, 7))
    fill_particles(P, sim, seed, 1)
    sim.ctx.upload_particles(0, P)
    for _ in range(warmup):
        sim.step_once()
    t0 = time.perf_counter()
    for _ in range(steps):
        sim.step_once()
    return n, time.perf_counter() - t0


def cpu_arm(cores, cells, steps, warmup):
    """particle-steps/s of the CPU port on `cores` processes (max-over-workers time)."""
    if not os.path.exists(ORACLE_FAST):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    ctx = mp.get_context("spawn")
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, [(cells, steps, warmup, 1000 + i) for i in range(cores)])
    total = sum(r[0] for r in res) * steps
    tmax = max(r[1] for r in res)
    return total / tmax, tmax / steps


def bind_near_gpu(dev):
    """What `mpirun --bind-to numa` / a job launcher does for the Fortran driver: run this rank's host threads -- and
    so place its page-locked buffers (first touch) -- on the NUMA node the GPU hangs off.  Eight unbound ranks on one
    host copied at 11.6 GB/s down and 22.6 GB/s up per rank (round 2, N = 8).  Returns (node, cpus) or None."""
    try:
        import torch
        p = torch.cuda.get_device_properties(dev)
        bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if len(allowed) < 2:
            return None
        os.sched_setaffinity(0, allowed)
        return node, len(allowed)
    except Exception:
        return None


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    cells = args.cpu_cells
    rate, s_per_step = cpu_arm(cores, cells, args.steps, args.warmup)
    ref_scaling = "strong" if (args.gpus > 1 and args.scaling != "weak") else "weak"
    sample = (f"{cores} independent periodic boxes of {cells}^3 cells x {PPC} ppc "
              f"({cores * cells ** 3 * PPC} particles), full step, {args.steps} steps after {args.warmup} warm-up")
    line = {
        "impl": "reference", "metric": "particle-steps/s (push+deposit, full PIC step)", "value": rate,
        "unit": "particle-steps/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": s_per_step * 1e3, "higher_is_better": True, "scaling": ref_scaling, "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, args.gpus, ref_scaling),
        "cpu_baseline": {"value": rate, "unit": "particle-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "particle-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "CPU restatement of the reference (oracle/, g++ -O3 -march=x86-64-v3, no FMA contraction); "
                "the Fortran reference cannot be built (no Fortran compiler, SDF submodule missing)",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, n_gpus, scaling="weak"):
    c = args.cells
    if scaling == "strong" and n_gpus > 1:
        glob, per = (c, c, c), (c, c, c // n_gpus)
        what = f"global {c}^3 cut into {n_gpus} slabs of {c}x{c}x{c // n_gpus} along z (strong scaling)"
    else:
        glob, per = (c, c, c * n_gpus), (c, c, c)
        what = f"{c}^3 per GPU, slabs along z" + (" (weak scaling)" if n_gpus > 1 else "")
    return {"workload": f"epoch3d uniform thermal plasma {glob[0]}x{glob[1]}x{glob[2]} cells x {PPC} ppc electrons, periodic, "
                        f"{SHAPE} shape, fp64 (BASELINE.json configs[2]; {what})",
            "cells_per_gpu": list(per), "ppc": PPC, "particles_per_gpu": per[0] * per[1] * per[2] * PPC,
            "density_m3": DENSITY, "temperature_keV": 1.0, "dx_over_debye": 0.5, "dt_over_cfl": 0.95,
            "l2_policy": "inputs larger than L2 (particle arrays %.1f GB per GPU)" % (per[0] * per[1] * per[2] * PPC * 56 / 1e9),
            "parallelism": f"domain decomposition 1x1x{n_gpus} (slab along z)"}


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except OSError:
        return {"hbm_gbs": 6650.0}, "fallback"


def run_ours(args):
    import torch
    import torch.distributed as dist
    from epoch_b200 import _abi
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product library has no CPU path")
    torch.cuda.set_device(local)
    numa = bind_near_gpu(local) if world > 1 else None
    abi = _abi.product()
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def fresh_nccl_id():
        """one ncclUniqueId per communicator (rank 0 makes it, torch.distributed carries it)"""
        if world == 1:
            return None
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            raw = (128 * __import__("ctypes").c_char)()
            assert abi.nccl_unique_id(raw) == 0
            buf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        return bytes(buf.cpu().numpy().tobytes())

    def barrier():
        if world > 1:
            dist.barrier()

    def reduce_ranks(x, op="max"):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op={"max": dist.ReduceOp.MAX, "sum": dist.ReduceOp.SUM, "min": dist.ReduceOp.MIN}[op])
        return float(t.item())

    import psutil
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    gen_threads = max(1, host_cores() if numa else host_cores() // max(1, local_world))
    peaks, how = measured_peaks()
    c = args.cells

    def sim_host_empty(pinned_host, shape):
        """page-locked float64 array (no context needed), ordinary memory if that fails"""
        import ctypes
        n = int(np.prod(shape))
        p = ctypes.c_void_p()
        if not pinned_host or abi.host_alloc(ctypes.byref(p), n * 8) != 0 or not p.value:
            return np.empty(shape)
        buf = (ctypes.c_double * n).from_address(p.value)
        import weakref
        weakref.finalize(buf, abi.host_free, p.value)
        return np.frombuffer(buf, dtype=np.float64).reshape(shape)

    def set_options(sim, n_part):
        # deferred permute: head room between the two states of the particle arrays (7 % of the species:
        # a quarter for the displacements -- one layer of tiles -- and three quarters for the particles of a launch)
        # (one 512^3 slab per GPU is memory-tight: a window of n / 14; smaller slabs take the library's own rule, which
        #  also keeps the window above 4.5 layers of tiles -- the thin z slabs of the strong-scaling cut need that)
        sim.ctx.set_option("defer_window", max(n_part // 14, min(n_part // 2, 1 << 24)) if n_part > 1_200_000_000 else "auto")
        for kv in args.opt:                       # tuning experiments: --opt push_kernel=tiled
            k, v = kv.split("=", 1)
            sim.ctx.set_option(k, v)
        sim.ctx.set_option("reserve", f"0:{n_part + max(n_part // 64, 1 << 18) if world > 1 else n_part}")

    def measure(cells_global, mode, want_e2e):
        """device-timed K steps (+ the e2e leg) of the deck on `cells_global`, slabs along z"""
        nz_local = cells_global[2] // world
        n_part = cells_global[0] * cells_global[1] * nz_local * PPC
        per_plane = cells_global[0] * cells_global[1] * PPC
        t_setup = time.perf_counter()
        sim = make_sim(abi, cells_global, rank, world, fresh_nccl_id(), device=local)
        set_options(sim, n_part)
        sim.ctx.set_option("timers", 1)
        # host copy of this rank's particle load (also the e2e leg's host buffer) -- if the host
        # has the memory for one per rank; else the load is generated and uploaded in slabs and
        # the e2e leg (which needs the whole state in host buffers) is skipped
        # (the host buffers are page-locked: keep a good part of the host's memory unpinned)
        host_budget = 0.70 * psutil.virtual_memory().available / max(1, local_world)
        if args.host_gb is not None:
            host_budget = args.host_gb * 1e9
        full_host = n_part * 56 * 1.04 < host_budget
        spare = max(n_part // 64, 1 << 18) if world > 1 else 0
        if full_host:
            host_all = sim.ctx.host_empty((n_part + spare, 7))   # room for migrated-in particles
            host = host_all[:n_part]
            fill_particles(host, sim, args.seed, gen_threads)
            t_gen = time.perf_counter() - t_setup
            sim.ctx.upload_particles(0, host)
        else:
            planes = max(1, min(nz_local, int(host_budget / (per_plane * 56 * (1 + 1 / 64)))))
            host_all = sim.ctx.host_empty((planes * per_plane + planes * per_plane // 64, 7))
            host = host_all[:planes * per_plane]
            t_gen = 0.0
            # (the extra weak-scaling run of an N > 1 invocation generates its first chunk of planes and uploads
            #  z-shifted copies of it for the others: the same uniform / Maxwellian statistics, minutes less numpy)
            replicate = mode == "weak" and not want_e2e
            for k0 in range(0, nz_local, planes):
                k1 = min(nz_local, k0 + planes)
                tg = time.perf_counter()
                if k0 == 0 or not replicate or k1 - k0 != planes:
                    fill_particles(host, sim, args.seed, gen_threads, k0, k1)
                else:
                    host[:, 2] += planes * sim.d[2]
                t_gen += time.perf_counter() - tg
                sim.ctx.upload_particles(0, host[:(k1 - k0) * per_plane], append=k0 > 0)
        sim.ctx.sync()
        t_up = time.perf_counter() - t_setup - t_gen

        # parity check, outside the timed region: the Gauss-law residual is captured at the first half step
        # (E at t + dt/2, particles at x(t): div E - rho(x + half drift)/eps0 is invariant under the Esirkepov
        # deposit), energy and particle count are taken before the warm-up
        def capture(s_):
            s_.ctx.set_option("capture_charge_reference", "1")
        check = {}
        i_cap = min(1, args.warmup - 1)      # (second warm-up step: the spare array of the first sort has been given back)
        for i in range(args.warmup):
            sim.step_once(after_half=capture if (i == i_cap and not args.no_check) else None)
            if i == i_cap and not args.no_check:
                fe0, pe0 = sim.ctx.energy()
                check["e0"] = (reduce_ranks(fe0, "sum"), reduce_ranks(sum(pe0), "sum"))
        sim.ctx.sync()
        for ph in ("push", "push_kernel", "push_A", "push_B", "sort", "fields", "halo", "bcs", "step"):
            sim.ctx.timer(ph, reset=True)
        l0 = sim.ctx.timer("launches")[1]
        clocks = ClockSampler(local)
        barrier(); torch.cuda.synchronize()
        if rank == 0:
            clocks.start()
        sim.ctx.mark("step", 1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sim.step_once()
        sim.ctx.mark("step", 0)
        sim.ctx.sync()
        wall = time.perf_counter() - t0
        barrier(); torch.cuda.synchronize()
        clk = clocks.stop() if rank == 0 else None
        dev_ms = sim.ctx.timer("step")[0]
        launches = sim.ctx.timer("launches")[1] - l0
        ms = reduce_ranks(dev_ms)
        phases = {ph: sim.ctx.timer(ph) for ph in ("push", "push_kernel", "push_A", "push_B", "sort", "fields", "halo", "bcs")}
        phases_max = {ph: reduce_ranks(v[0]) / args.steps for ph, v in phases.items()}
        sorts = {"deferred_to_next_push": sim.ctx.timer("deferred_sorts")[1], "permuted_in_sort": sim.ctx.timer("permuted_sorts")[1]}
        n_live = sim.ctx.particle_count(0)
        n_global = int(round(reduce_ranks(float(n_live), "sum")))
        n_expected = cells_global[0] * cells_global[1] * cells_global[2] * PPC
        value = n_expected * args.steps / (ms * 1e-3)
        parity = None
        if not args.no_check:
            res = {}

            def residual(s_):
                res["r"] = s_.ctx.charge_residual()
            sim.step_once(after_half=residual)       # one more (untimed) step: residual at its half step
            fe1, pe1 = sim.ctx.energy()
            e1 = (reduce_ranks(fe1, "sum"), reduce_ranks(sum(pe1), "sum"))
            r_max, rho_max = reduce_ranks(res["r"][0]), reduce_ranks(res["r"][1])
            e0 = check["e0"]
            drift = (e1[0] + e1[1]) / (e0[0] + e0[1]) - 1.0
            n_after = int(round(reduce_ranks(float(sim.ctx.particle_count(0)), "sum")))
            parity = {"particles": n_after, "particles_expected": n_expected,
                      "energy_field_J": e1[0], "energy_kinetic_J": e1[1],
                      "energy_total_drift_rel": drift, "steps_between_energy_samples": args.warmup - 1 - i_cap + args.steps + 1,
                      "charge_residual_max": r_max, "rho_over_eps0_max": rho_max,
                      "charge_residual_rel": r_max / max(rho_max, 1e-300),
                      "what": "global particle count (periodic box: conserved exactly), total energy drift over the run, "
                              "max |div E - rho/eps0| against its value at the first step (exact charge conservation "
                              "of the Esirkepov deposit); sums / maxima over all ranks"}
            parity["ok"] = bool(n_after == n_expected and abs(drift) < 1e-2 and parity["charge_residual_rel"] < 1e-9)

        # roofline of the dominant kernel pair (push + deposit): algorithmic bytes per launch
        # = 104 B per particle (read x,y,z,px,py,pz,w; write x,y,z,px,py,pz) + 72 B per
        # cell (read 6 field components, write 3 current components)   [SURVEY 8d]
        pk_ms, pk_launch = phases["push_kernel"]
        roof = None
        if pk_launch > 0 and pk_ms > 0:
            n_launch = args.steps      # one species => one push launch per step
            ncell_local = cells_global[0] * cells_global[1] * nz_local
            alg_bytes = 104.0 * n_live + 72.0 * ncell_local
            t_launch = pk_ms / n_launch * 1e-3
            ach = alg_bytes / t_launch / 1e9
            roof = {"bound": "hbm", "kernel": "k_push_split + k_deposit_split (one push+deposit pass over all slabs = one 'launch')", "achieved": ach, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "frac": ach / peaks["hbm_gbs"], "traffic": None, "peak_source": how,
                    "ms_per_launch": pk_ms / n_launch, "algorithmic_bytes_per_launch": alg_bytes,
                    "particle_steps_per_s_kernel_only": n_live / t_launch,
                    "share_of_step": pk_ms / dev_ms}
            tr = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tr):
                try:
                    tj = json.load(open(tr))
                    if tj.get("push_bytes_per_particle") is not None:       # ncu DRAM bytes per particle (profiles/) x particles of a pass
                        roof["traffic"] = tj["push_bytes_per_particle"] * n_live
                        roof["traffic_source"] = tj.get("source", "profiles/traffic.json (ncu --set full dram bytes per particle)")
                        roof["frac_of_hbm_with_measured_traffic"] = roof["traffic"] / t_launch / 1e9 / peaks["hbm_gbs"]
                    # the pair is not HBM bound: the fp64 pipe and the shared-memory data stage are busier.  Their fractions
                    # come from the ncu capture (instructions / wavefronts per particle) x the live particle rate.
                    if tj.get("fp64_inst_per_particle") is not None:
                        dfma_peak = tj.get("dfma_peak_per_s", 17.245e12)      # measured: 34.49 TFLOP/s DFMA (profiles/r01_microbench_b200.txt)
                        roof["fp64"] = {"achieved_inst_per_s": tj["fp64_inst_per_particle"] * n_live / t_launch,
                                        "peak_inst_per_s": dfma_peak,
                                        "frac": tj["fp64_inst_per_particle"] * n_live / t_launch / dfma_peak,
                                        "ncu_sm__pipe_fp64_cycles_active_pct": tj.get("ncu_fp64_pipe_pct")}
                    if tj.get("lsu_wavefronts_per_particle") is not None:
                        wf_peak = 148 * tj.get("sm_clock_hz", 1.93e9)         # one shared-memory / L1 data-stage wavefront per clock per SM
                        roof["l1_data_stage"] = {"frac": tj["lsu_wavefronts_per_particle"] * n_live / t_launch / wf_peak,
                                                 "ncu_l1tex__data_pipe_lsu_wavefronts_pct": tj.get("ncu_l1_wavefront_pct")}
                    fr = {"hbm": roof.get("frac_of_hbm_with_measured_traffic", roof["frac"]),
                          "fp64": roof.get("fp64", {}).get("frac", 0.0), "l1": roof.get("l1_data_stage", {}).get("frac", 0.0)}
                    roof["binding"] = max(fr, key=fr.get)
                    roof["bound"] = "hbm" if roof["binding"] == "hbm" else roof["binding"]
                except Exception:
                    pass
        fields_ms, fields_l = phases["fields"]
        field_line = None
        if fields_l > 0:
            cells = (cells_global[0] + 1) * (cells_global[1] + 1) * (nz_local + 1)
            fb = (2 * 96.0 + 2 * 72.0) * cells * args.steps      # 2 E + 2 B half-steps per step
            field_line = {"cell_updates_per_s": 4 * cells * args.steps / (fields_ms * 1e-3),
                          "achieved_gbs": fb / (fields_ms * 1e-3) / 1e9,
                          "frac_of_hbm": fb / (fields_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}
        sim_ng = sim.ctx.ng
        pinned_host = bool(getattr(sim.ctx, "pinned_host", False))
        sim.ctx.close()
        del sim

        # e2e: same metric through the C ABI from host buffers (upload, K steps, download)
        def e2e_leg(steps):
            if not full_host:
                return None
            shp = tuple(reversed([n + 2 * sim_ng for n in (cells_global[0], cells_global[1], nz_local)]))
            fzero = sim_host_empty(pinned_host, shp)
            fzero[...] = 0.0
            fout = {k: sim_host_empty(pinned_host, shp) for k in ("ex", "ey", "ez", "bx", "by", "bz", "jx", "jy", "jz")}
            for a in fout.values():
                a[...] = 0.0          # touch the pages before the clock starts
            nid = fresh_nccl_id()
            # the context (stream, NCCL communicator with its neighbour links, particle storage) is the one-time set-up
            # of a run -- MPI_INIT and the ALLOCATEs of the Fortran driver -- and stays outside the timed region
            tc = time.perf_counter()
            sim2 = make_sim(abi, cells_global, rank, world, nid, device=local)
            set_options(sim2, n_part)
            sim2.ctx.sync()
            t_create = time.perf_counter() - tc
            barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            sim2.ctx.upload_fields(ex=fzero, ey=fzero, ez=fzero, bx=fzero, by=fzero, bz=fzero)
            t1 = time.perf_counter()
            sim2.ctx.upload_particles(0, host)
            t2 = time.perf_counter()
            t2a = t2
            for i in range(steps):
                sim2.step_once()
                if i == 0:            # the first step carries the initial sort and the one-time allocations
                    sim2.ctx.sync()
                    t2a = time.perf_counter()
            sim2.ctx.sync()
            t3 = time.perf_counter()
            got = sim2.ctx.download_particles(0, out=host_all)
            fld = sim2.ctx.download_fields(out=fout)
            sim2.ctx.sync()
            t4 = time.perf_counter()
            t_e2e = reduce_ranks(t4 - t0)
            h2d = host.nbytes + 6 * fzero.nbytes
            d2h = got.nbytes + 9 * fzero.nbytes
            out = {"value": n_expected * steps / t_e2e, "unit": "particle-steps/s",
                   "h2d_bytes_per_step": h2d * world / steps, "d2h_bytes_per_step": d2h * world / steps,
                   "seconds": t_e2e, "steps_per_residency": steps,
                   "seconds_create_untimed": t_create,
                   "seconds_split": {"upload_fields": t1 - t0, "upload_particles": t2 - t1,
                                     "first_step": t2a - t2, "other_steps": t3 - t2a, "download": t4 - t3},
                   "h2d_gbs": host.nbytes / (t2 - t1) / 1e9, "d2h_gbs": (got.nbytes + 9 * fzero.nbytes) / (t4 - t3) / 1e9,
                   "host_buffers": "page-locked (eb200_host_alloc)" if pinned_host else "pageable (staged through the library's pinned ring)",
                   "what": "upload_fields + upload_particles from host buffers + K full steps + "
                           "download_particles + download_fields into host buffers + sync, wall clock, max over ranks, "
                           "on the SAME deck and size as `value`; the state lives on the device between steps, so the "
                           "copies are amortised over K steps; eb200_create (context, NCCL communicator, storage) is the "
                           "run's one-time set-up and is timed apart (seconds_create_untimed)"}
            del fld
            sim2.ctx.close()
            return out
        e2e = None
        if want_e2e and not args.no_e2e:
            e2e = e2e_leg(args.steps)
            if e2e is None:
                e2e = {"value": None, "unmeasured": "the host cannot hold this rank's particle state in page-locked buffers "
                                                    f"({n_part * 56 / 1e9:.0f} GB per rank, {local_world} ranks on this host)"}
            elif args.e2e_long > 0 and world == 1:
                lg = e2e_leg(args.e2e_long)     # the same with more steps per residency: the copies amortise
                e2e["longer_residency"] = {k: lg[k] for k in ("value", "seconds", "steps_per_residency", "seconds_split", "seconds_create_untimed")}
        return {
            "value": value, "ms_per_step": ms / args.steps, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
            "roofline": roof, "fields": field_line, "parity_check": parity,
            "phases_ms_per_step": {k: v[0] / args.steps for k, v in phases.items()},
            "phases_ms_per_step_max_over_ranks": phases_max,
            "sorts_whole_run": sorts, "wall_ms_per_step": wall * 1e3 / args.steps,
            "setup_seconds": {"generate_host": t_gen, "upload": t_up},
            "cells_global": list(cells_global), "cells_per_gpu": [cells_global[0], cells_global[1], nz_local],
            "particles_global": n_expected, "scaling": mode,
        }

    scaling = args.scaling
    if scaling == "auto":
        scaling = "strong" if world > 1 else "weak"
    if world > 1 and c % world != 0:
        raise SystemExit("--cells must be a multiple of the number of GPUs for the strong-scaling mode")
    main_res = measure((c, c, c) if scaling == "strong" else (c, c, c * world), scaling, True)
    other = None
    if world > 1 and args.scaling == "auto" and not args.one_mode:
        # the other mode as an extra key (device-timed only): weak = one 512^3 slab per GPU
        other = measure((c, c, c * world), "weak", False)
    if rank != 0:
        if main_res["parity_check"] is not None and not main_res["parity_check"]["ok"]:
            sys.exit(3)
        return
    line = {
        "metric": "particle-steps/s (push+deposit, full PIC step)", "value": main_res["value"], "unit": "particle-steps/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": main_res["ms_per_step"],
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, world, scaling), "clocks": main_res["clocks"], "e2e": main_res["e2e"],
        "gpu_launches": main_res["gpu_launches"], "roofline": main_res["roofline"], "fields": main_res["fields"],
        "parity_check": main_res["parity_check"],
        "phases_ms_per_step": main_res["phases_ms_per_step"],
        "phases_ms_per_step_max_over_ranks": main_res["phases_ms_per_step_max_over_ranks"],
        "sorts_whole_run": main_res["sorts_whole_run"], "wall_ms_per_step": main_res["wall_ms_per_step"],
        "setup_seconds": main_res["setup_seconds"],
        "host_binding": None if numa is None else {"numa_node_rank0": numa[0], "cpus": numa[1]},
    }
    if other is not None:
        line["weak"] = {k: other[k] for k in ("value", "ms_per_step", "cells_global", "cells_per_gpu", "particles_global",
                                              "phases_ms_per_step_max_over_ranks", "parity_check", "gpu_launches")}
        line["weak"]["what"] = (f"the same deck with one {c}^3 slab per GPU stacked along z (weak scaling), device-timed; "
                                "efficiency = weak.value / (n_gpus x the 1-GPU value)")
    if world == 1 and not args.no_cpu:
        cores = host_cores()
        rate, s_per = cpu_arm(cores, args.cpu_cells, 3, 1)
        line["cpu_baseline"] = {"value": rate, "unit": "particle-steps/s", "cores": cores, "kind": "port",
                                "sample": f"{cores} independent periodic boxes of {args.cpu_cells}^3 cells x {PPC} ppc, "
                                          f"full step, 3 steps after 1 warm-up ({s_per:.2f} s/step)"}
    print(json.dumps(line), flush=True)
    if main_res["parity_check"] is not None and not main_res["parity_check"]["ok"]:
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, default=512, help="cells per dimension per GPU")
    ap.add_argument("--cpu-cells", type=int, default=64, help="cells per dimension per CPU process (cpu arm)")
    ap.add_argument("--seed", type=int, default=12345)
    ap.add_argument("--shape", default="triangle", choices=["tophat", "triangle", "bspline3"])
    ap.add_argument("--scaling", default="auto", choices=["auto", "strong", "weak"],
                    help="N > 1: strong = the global cells^3 box cut into N slabs (the line printed; auto adds the weak "
                         "run as an extra key), weak = one cells^3 slab per GPU")
    ap.add_argument("--one-mode", action="store_true", help="N > 1, --scaling auto: skip the extra weak-scaling run")
    ap.add_argument("--no-check", action="store_true", help="skip the parity check (particle count, energy drift, Gauss residual)")
    ap.add_argument("--e2e-long", type=int, default=100, help="steps of the second e2e leg at N = 1 (0 = skip)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--host-gb", type=float, default=None, help="host memory per rank for the particle copy (default: from free memory)")
    ap.add_argument("--opt", action="append", default=[], help="library option key=value (tuning experiments)")
    args = ap.parse_args()
    global SHAPE
    SHAPE = args.shape
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
