"""Print measured vs recorded group velocities of the reference's custom-stencil decks (CUDA library)."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from epoch_b200 import _abi
import test_maxwell_solvers as T
abi = _abi.product()
for name, (nd, _, vgc, rtol, rec) in T.CUSTOM_DECKS.items():
    sim = T.custom_deck(abi, name)
    vg, pts = T.centroid_velocity(sim, 12e-15 if nd == 1 else 25e-15)
    exp = vgc * T.C if vgc else float("nan")
    print(f"{name:26s} vg {vg:.3f} recorded {rec:.3f} rel {abs(vg / rec - 1):.3e}  expected {exp:.3f} rel {abs(vg / exp - 1):.3e}")
