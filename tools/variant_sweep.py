"""Kernel-variant sweep on the bench deck (uniform thermal plasma, cells^3 x 16 ppc, triangle, periodic, one GPU).
Every variant library runs in its own process (EB200_LIB); to keep a variant's turn short the particle load is
generated for `--block` z planes only and uploaded nz/block times with the z coordinate shifted (the load is
statistically the bench's: uniform in the cell, Maxwellian momenta), so a 512^3 turn costs ~40 s instead of ~100 s.
usage: python tools/variant_sweep.py [--cells 512] [--steps 5] [--warmup 3] name=path/to/libepoch_b200.so ...
       (name=default takes the in-tree build).  One line per variant; device-timed phases in ms per step."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)


def worker(args):
    import numpy as np
    import bench
    from epoch_b200 import _abi
    abi = _abi.product()
    c = args.cells
    bench.SHAPE = args.shape
    cz = args.cells_z or c
    sim = bench.make_sim(abi, (c, c, cz))
    n_part = c * c * cz * bench.PPC
    sim.ctx.set_option("defer_window", max(n_part // 14, min(n_part // 2, 1 << 24)))
    for kv in args.opt:
        k, v = kv.split("=", 1)
        sim.ctx.set_option(k, v)
    sim.ctx.set_option("reserve", f"0:{n_part}")
    sim.ctx.set_option("timers", 1)
    blk = min(args.block, cz)
    per_plane = c * c * bench.PPC
    host = sim.ctx.host_empty((blk * per_plane, 7))
    bench.fill_particles(host, sim, 777, bench.host_cores(), 0, blk)
    dz = blk * sim.d[2]
    for k0 in range(0, cz, blk):
        m = min(blk, cz - k0) * per_plane
        if k0 > 0:
            host[:, 2] += dz
        sim.ctx.upload_particles(0, host[:m], append=k0 > 0)
    sim.ctx.sync()
    for _ in range(args.warmup):
        sim.step_once()
    sim.ctx.sync()
    names = ("push", "push_kernel", "push_A", "push_B", "sort", "fields", "halo", "bcs", "step")
    for ph in names:
        sim.ctx.timer(ph, reset=True)
    sim.ctx.mark("step", 1)
    for _ in range(args.steps):
        sim.step_once()
    sim.ctx.mark("step", 0)
    sim.ctx.sync()
    ms = sim.ctx.timer("step")[0] / args.steps
    out = {"variant": args.name, "value": n_part / (ms * 1e-3), "ms_per_step": ms,
           "phases": {ph: sim.ctx.timer(ph)[0] / args.steps for ph in names[:-1]},
           "particles": int(sim.ctx.particle_count(0)), "cell_mismatch": int(sim.ctx.timer("cell_mismatch")[1])}
    if args.check:
        res = {}
        sim.step_once(after_half=lambda s_: s_.ctx.set_option("capture_charge_reference", "1"))
        sim.step_once(after_half=lambda s_: res.setdefault("r", s_.ctx.charge_residual()))
        out["charge_residual_rel"] = res["r"][0] / max(res["r"][1], 1e-300)
    print(json.dumps(out), flush=True)
    sim.ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", type=int, default=512)
    ap.add_argument("--cells-z", type=int, default=0, help="z extent if not cubic (one slab of the bench box)")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--block", type=int, default=32)
    ap.add_argument("--shape", default="triangle")
    ap.add_argument("--opt", action="append", default=[])
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--name", default=None, help="(worker mode)")
    ap.add_argument("variants", nargs="*")
    args = ap.parse_args()
    if args.name is not None:
        return worker(args)
    for v in args.variants:
        name, path = v.split("=", 1)
        env = dict(os.environ)
        if path != "default":
            env["EB200_LIB"] = os.path.abspath(path)
        cmd = [sys.executable, os.path.abspath(__file__), "--name", name, "--cells", str(args.cells), "--cells-z", str(args.cells_z),
               "--steps", str(args.steps),
               "--warmup", str(args.warmup), "--block", str(args.block), "--shape", args.shape] + (["--check"] if args.check else [])
        for o in args.opt:
            cmd += ["--opt", o]
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
        line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ""
        if r.returncode != 0 or not line.startswith("{"):
            print(json.dumps({"variant": name, "error": (r.stderr or r.stdout)[-400:]}), flush=True)
        else:
            print(line, flush=True)


if __name__ == "__main__":
    main()
