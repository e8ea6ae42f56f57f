import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from epoch_b200 import _abi
from epoch_b200.driver import Simulation, SpeciesSpec, Q0, M0, C
abi = _abi.product()
res = {}
for kern in ("baseline", "tiled"):
    s = Simulation(3, [16, 12, 8], [0.0] * 3, [16e-6, 12e-6, 8e-6], shape="triangle", species=[SpeciesSpec("e", -Q0, M0)], abi=abi)
    P = np.array([[5.3e-6, 4.2e-6, 3.6e-6, 0.3 * M0 * C, -0.2 * M0 * C, 0.1 * M0 * C, 1e10]])
    s.ctx.upload_particles(0, P)
    s.ctx.set_option("push_kernel", kern)
    s.ctx.set_time(s.dt, 0.5 * s.dt, 0)
    s.ctx.push_particles()
    res[kern] = s.ctx.download_fields(("jx", "jy", "jz"))
for k in ("jx", "jy", "jz"):
    a, b = res["baseline"][k], res["tiled"][k]
    print(k, "sum", a.sum(), b.sum(), "absmax", np.abs(a).max(), np.abs(b).max())
    ia = np.argwhere(np.abs(a) > 1e-12 * np.abs(a).max()); ib = np.argwhere(b != 0)
    print(" baseline nz", len(ia), ia.min(0), ia.max(0), " tiled nz", len(ib), (ib.min(0), ib.max(0)) if len(ib) else None)
    for i in ia[:6]:
        print("   ", tuple(i), a[tuple(i)], b[tuple(i)])
