"""Per-phase wall clock of the end-to-end path of bench.py (create, upload, each step, download)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import bench
from epoch_b200 import _abi

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 256
abi = _abi.product()
t = time.perf_counter()
sim = bench.make_sim(abi, (cells,) * 3)
n_part = cells ** 3 * bench.PPC
host = np.empty((n_part, 7))
bench.fill_particles(host, sim, 1, 16)
print("gen", time.perf_counter() - t)
for rep in range(2):
    t0 = time.perf_counter()
    if rep:
        sim = bench.make_sim(abi, (cells,) * 3)
    sim.ctx.set_option("reserve", f"0:{n_part}")
    fz = np.zeros(sim.ctx._np_shape())
    sim.ctx.upload_fields(ex=fz, ey=fz, ez=fz, bx=fz, by=fz, bz=fz)
    t1 = time.perf_counter()
    sim.ctx.upload_particles(0, host)
    t2 = time.perf_counter()
    ts = []
    for i in range(5):
        sim.step_once(); sim.ctx.sync(); ts.append(time.perf_counter())
    got = sim.ctx.download_particles(0, out=host)
    t3 = time.perf_counter()
    fld = sim.ctx.download_fields(); sim.ctx.sync()
    t4 = time.perf_counter()
    print(f"rep {rep}: create+fields {t1-t0:.3f} upload {t2-t1:.3f} ({host.nbytes/(t2-t1)/1e9:.1f} GB/s) steps",
          [round(b - a, 4) for a, b in zip([t2] + ts[:-1], ts)], f"dl_part {t3-ts[-1]:.3f} ({host.nbytes/(t3-ts[-1])/1e9:.1f} GB/s) dl_fields {t4-t3:.3f}")
    sim.ctx.close()
