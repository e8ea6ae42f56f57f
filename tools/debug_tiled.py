"""Device-side A/B: tiled push kernel vs the general (baseline) kernel on the same deck."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from epoch_b200 import _abi
from helpers import make_plasma, match_particles, relerr

abi = _abi.product()
for nd, shape, n, uth in [(3, "triangle", [16, 12, 8], 0.05), (3, "triangle", [16, 12, 8], 0.5), (2, "triangle", [24, 20], 0.5),
                          (3, "tophat", [16, 12, 8], 0.5), (2, "tophat", [24, 20], 0.5)]:
    sims = []
    for kern in ("baseline", os.environ.get("KERN", "split")):
        s = make_plasma(abi, nd, shape, n, seed=7, ppc=8, u_th=uth)
        s.ctx.set_option("push_kernel", kern)
        s.ctx.set_time(s.dt, 0.5 * s.dt, 0)
        s.ctx.push_particles()
        sims.append(s)
    out = [nd, shape, uth, "mismatch", sims[1].ctx.timer("cell_mismatch")[1]]
    for isp in range(2):
        A, B = match_particles(sims[0].ctx.download_particles(isp), sims[1].ctx.download_particles(isp))
        out += ["pos", float(np.max(np.abs(A[:, :nd] - B[:, :nd])) / sims[0].length[0]), "mom", relerr(B[:, nd:nd + 3], A[:, nd:nd + 3])]
    fa, fb = sims[0].ctx.download_fields(("jx", "jy", "jz")), sims[1].ctx.download_fields(("jx", "jy", "jz"))
    for k in fa:
        out += [k, relerr(fb[k], fa[k])]
        if relerr(fb[k], fa[k]) > 1e-10:
            d = np.abs(fb[k] - fa[k])
            idx = np.unravel_index(np.argmax(d), d.shape)
            out += ["at", idx, "shape", d.shape, "ref", fa[k][idx], "got", fb[k][idx], "nbad", int((d > 1e-9 * np.abs(fa[k]).max()).sum())]
    print(*out)
