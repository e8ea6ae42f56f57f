"""Deterministic deposition at scale: two runs with deterministic_deposit=1 must agree bitwise; the default
mode is shown beside it (it usually differs in the last bits) together with the cost of the mode.
usage: python tools/det_check.py [cells]"""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from epoch_b200 import _abi
from helpers import make_plasma

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 64
abi = _abi.product()


def run(det):
    s = make_plasma(abi, 3, "triangle", [cells] * 3, seed=3, ppc=16, u_th=0.06, e_amp=1e8, b_amp=1.0, two_species=False,
                    options={"deterministic_deposit": det, "timers": 1})
    s.step_once()
    s.ctx.sync()
    t = time.perf_counter()
    for _ in range(4):
        s.step_once()
    s.ctx.sync()
    dt = (time.perf_counter() - t) / 4
    f = s.ctx.download_fields(("jx", "jy", "jz", "ex"))
    return f, dt

for det in (1, 0):
    (a, ta), (b, tb) = run(det), run(det)
    same = all(np.array_equal(a[k], b[k]) for k in a)
    diff = max(float(np.max(np.abs(a[k] - b[k])) / max(np.max(np.abs(a[k])), 1e-300)) for k in a)
    print(f"deterministic_deposit={det}: {cells}^3 x 16 ppc, {1e3 * min(ta, tb):.2f} ms/step, two runs bitwise equal: {same}, max rel diff {diff:.2e}")
    if det:
        assert same
