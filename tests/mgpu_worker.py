"""Worker of tests/test_multi_gpu.py (launched with torchrun, one rank per GPU).

Runs the same seeded 3D plasma (i) decomposed over WORLD_SIZE GPUs along the axis
given on the command line and (ii) on rank 0 alone, and compares particles and
interior fields.  Exit code 0 = match."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from epoch_b200 import _abi  # noqa: E402
from epoch_b200.driver import Simulation, SpeciesSpec, Q0, M0, C  # noqa: E402

AXIS = int(sys.argv[1]) if len(sys.argv) > 1 else 2
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 4
SHAPE = sys.argv[3] if len(sys.argv) > 3 else "triangle"
WINDOW = len(sys.argv) > 4 and sys.argv[4] == "window"      # moving window at 0.9 c (x decomposition, open x boundaries)
CPML = len(sys.argv) > 4 and sys.argv[4] == "cpml"          # CPML layers on the decomposed axis and on x
BALANCE = len(sys.argv) > 4 and sys.argv[4] == "balance"    # lopsided plasma, re-sliced after 2 steps
SMALLBOX = len(sys.argv) > 4 and sys.argv[4] == "smallbox"  # migration mailboxes far too small at first: leavers wait, boxes grow
N = [16, 12, 20]
CELL = 1e-6


def fields_on(sim):
    """smooth periodic E,B evaluated on the rank's grid incl. ghosts"""
    ng = sim.ctx.ng
    ax = [sim.grid_min_local[d] + (np.arange(-ng, sim.n_local[d] + ng)) * sim.d[d] for d in range(3)]
    Z, Y, X = np.meshgrid(ax[2], ax[1], ax[0], indexing="ij")
    L = [N[d] * CELL for d in range(3)]
    ph = 2 * np.pi * (X / L[0] + 2 * Y / L[1] + Z / L[2])
    out = {}
    for i, k in enumerate(("ex", "ey", "ez")):
        out[k] = 3e10 * np.sin(ph + i)
    for i, k in enumerate(("bx", "by", "bz")):
        out[k] = 1e2 * np.cos(ph + 0.5 * i)
    return out


def build(abi, rank, nproc, nccl_id, device):
    nprocdir = [1, 1, 1]
    nprocdir[AXIS] = nproc
    sp = [SpeciesSpec("e", -Q0, M0, density=1e25, temperature=(1e6,) * 3),
          SpeciesSpec("p", Q0, 1836.2 * M0, density=1e25, temperature=(1e6,) * 3)]
    kw = {}
    if WINDOW:
        obc = ["open", "open"] + ["periodic"] * 4
        kw = dict(bc_field=obc)
        for s in sp:
            s.bcs = obc
    if CPML:
        fbc, pbc = ["periodic"] * 6, ["periodic"] * 6
        for d in {0, AXIS}:
            fbc[2 * d:2 * d + 2] = ["cpml_outflow"] * 2
            pbc[2 * d:2 * d + 2] = ["open"] * 2
        kw = dict(bc_field=fbc)
        for s in sp:
            s.bcs = pbc
    if SMALLBOX:
        kw["options"] = {"mig_capacity": 16}
    if os.environ.get("MGPU_OPTS"):          # bisecting aid: library options "k=v,k=v"
        kw.setdefault("options", {}).update(dict(kv.split("=", 1) for kv in os.environ["MGPU_OPTS"].split(",")))
    sim = Simulation(3, N, [0.0] * 3, [n * CELL for n in N], shape=SHAPE, species=sp, abi=abi, device=device,
                     rank=rank, nproc=nproc, nprocdir=tuple(nprocdir), nccl_id=nccl_id, **kw)
    rng = np.random.default_rng(11)
    npart = 6 * N[0] * N[1] * N[2]
    for isp, s in enumerate(sp):
        P = np.empty((npart, 7))
        for d in range(3):
            P[:, d] = rng.uniform(0.0, N[d] * CELL, npart)
        P[:, 3:6] = rng.normal(0.0, 0.6, (npart, 3)) * s.mass * C
        P[:, 6] = 1e25 * CELL ** 3 / 6 * (1.0 + np.arange(npart) / npart)      # unique weights
        own = np.ones(npart, bool)
        if BALANCE:       # most of the plasma sits in the low 40 % of the decomposed axis
            own &= (P[:, AXIS] < 0.4 * N[AXIS] * CELL) | (rng.random(npart) < 0.15)
        for d in range(3):
            own &= (P[:, d] >= sim.min_local[d]) & (P[:, d] < sim.max_local[d])
        sim.ctx.upload_particles(isp, P[own])
    sim.ctx.upload_fields(**fields_on(sim))
    if WINDOW:
        shifts = [0]

        def insert(s_, x0):
            # the incoming column over the full y-z extent (only the last x rank inserts): the same
            # particles whatever the decomposition along x
            shifts[0] += 1
            r = np.random.default_rng([5, shifts[0]])
            n = 3 * N[1] * N[2]
            out = []
            for isp, spc in enumerate(sp):
                P = np.empty((n, 7))
                P[:, 0] = x0 + r.random(n) * CELL
                P[:, 1] = r.uniform(0.0, N[1] * CELL, n)
                P[:, 2] = r.uniform(0.0, N[2] * CELL, n)
                P[:, 3:6] = r.normal(0.0, 0.3, (n, 3)) * spc.mass * C
                P[:, 6] = 1e25 * CELL ** 3 / 6 * (3.0 + shifts[0] + np.arange(n) / n + 0.5 * isp)
                out.append(P)
            return out
        sim.enable_window(v_x=0.9 * C, insert=insert)
    return sim


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    abi = _abi.product()
    import ctypes
    buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        raw = (ctypes.c_char * 128)()
        assert abi.nccl_unique_id(raw) == 0
        buf.copy_(torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8))
    dist.broadcast(buf, 0)
    sim = build(abi, rank, world, bytes(buf.cpu().numpy().tobytes()), local)
    def allreduce(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
        dist.all_reduce(t)
        return t.cpu().numpy()
    for it in range(STEPS):
        if BALANCE and it == 2 and not os.environ.get("MGPU_NO_REBALANCE"):
            before = [sim.ctx.particle_count(i) for i in range(2)]

            def snapshot():
                """global interior fields (assembled on rank 0) and the particle set"""
                ng_ = sim.ctx.ng
                f_ = sim.ctx.download_fields()
                loc = {k: v[ng_:-ng_, ng_:-ng_, ng_:-ng_].copy() for k, v in f_.items()}
                pp = [sim.ctx.download_particles(i) for i in range(2)]
                got = [None] * world
                dist.all_gather_object(got, (list(sim.cell_min), list(sim.cell_max), loc, pp))
                full = {k: np.zeros((N[2], N[1], N[0])) for k in loc}
                for (cmin, cmax, loc_r, _) in got:
                    sl = tuple(slice(cmin[d] - 1, cmax[d]) for d in (2, 1, 0))
                    for k in loc_r:
                        full[k][sl] = loc_r[k]
                parts = [np.concatenate([g_[3][i] for g_ in got]) for i in range(2)]
                parts = [P_[np.argsort(P_[:, 6])] for P_ in parts]
                return full, parts
            snap0 = snapshot()
            changed = sim.balance_workload(allreduce)
            snap1 = snapshot()
            for k in snap0[0]:
                assert np.array_equal(snap0[0][k], snap1[0][k]), "re-slicing changed the interior of " + k
            for i in range(2):
                assert np.array_equal(snap0[1][i], snap1[1][i]), "re-slicing changed the particles of species %d" % i
            after = [sim.ctx.particle_count(i) for i in range(2)]
            tot = allreduce(np.array([sum(before), sum(after), max(sum(after), 0)], dtype=np.int64))
            mx = torch.tensor([sum(before), sum(after)], device="cuda")
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            if rank == 0:
                print("balance: changed", changed, "tables", sim.cell_mins[AXIS], sim.cell_maxs[AXIS],
                      "particles total", int(tot[0]), "->", int(tot[1]), "max per rank", mx.tolist())
            assert changed and tot[0] == tot[1], "re-slicing must move the break and keep every particle"
            assert mx[1].item() < mx[0].item(), "the fullest rank must hold fewer particles afterwards"
        sim.step_once()
    ng = sim.ctx.ng
    parts = [sim.ctx.download_particles(i) for i in range(2)]
    f = sim.ctx.download_fields()
    # derived grid variables (calc_df.F90) take part in the comparison like fields
    DERIVED = [("charge_density", -1, 0), ("temperature", 0, 0), ("ekbar", 1, 0), ("species_current", -1, 1)]
    for kind, isp, direction in DERIVED:
        f["%s:%d:%d" % (kind, isp, direction)] = sim.ctx.calc_derived(kind, isp, direction)
    inter = {k: v[ng:-ng, ng:-ng, ng:-ng].copy() for k, v in f.items()}
    gathered = [None] * world
    dist.all_gather_object(gathered, (sim.cell_min, sim.cell_max, parts, inter))
    ok = True
    if rank == 0:
        ref = build(abi, 0, 1, None, local)
        for _ in range(STEPS):
            ref.step_once()
        rf = ref.ctx.download_fields()
        for kind, isp, direction in DERIVED:
            rf["%s:%d:%d" % (kind, isp, direction)] = ref.ctx.calc_derived(kind, isp, direction)
        for isp in range(2):
            A = np.concatenate([g[2][isp] for g in gathered])
            B = ref.ctx.download_particles(isp)
            if A.shape != B.shape:
                print("particle count mismatch", A.shape, B.shape); ok = False; continue
            A, B = A[np.argsort(A[:, 6])], B[np.argsort(B[:, 6])]
            # periodic wrap: compare positions modulo the box length
            for d in range(3):
                L = N[d] * CELL
                dx = np.abs(A[:, d] - B[:, d]); dx = np.minimum(dx, np.abs(L - dx))
                if dx.max() / L > 1e-11: print("pos", isp, d, dx.max() / L); ok = False
            e = np.abs(A[:, 3:6] - B[:, 3:6]).max() / np.abs(B[:, 3:6]).max()
            if e > 1e-11: print("mom", isp, e); ok = False
        for k in rf:
            full = rf[k][ng:-ng, ng:-ng, ng:-ng]
            scale = max(np.abs(full).max(), 1e-300)
            for (cmin, cmax, _, inter_r) in gathered:
                sl = tuple(slice(cmin[d] - 1, cmax[d]) for d in (2, 1, 0))
                e = np.abs(inter_r[k] - full[sl]).max() / scale
                if e > 1e-10: print("field", k, e); ok = False
        print("multi-gpu parity", "OK" if ok else "FAILED", "axis", AXIS, "shape", SHAPE, "ranks", world)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
