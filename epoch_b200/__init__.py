"""epoch_b200 -- B200-native (sm_100a) implementation of EPOCH's per-timestep PIC
hot path behind a C ABI (include/epoch_b200.h).  Python here is harness glue:
ctypes binding (`_abi`), the host-side mirror of the reference entry points
(`context.PicContext`), the driver-loop mirror (`driver.Simulation`) and the
domain decomposition math (`decomp`); `balance` restates the reference's host-side
load-balancing decisions for the same harness."""
from .context import PicContext, EpochError, BC, SHAPES  # noqa: F401
