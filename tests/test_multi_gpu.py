"""-m gpu, needs >= 2 GPUs (skipped otherwise): the domain-decomposed run over NCCL
(field halos, current ghost sums, particle migration) equals the single-GPU run."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("axis,shape", [(2, "triangle"), (0, "triangle"), (1, "tophat"), (2, "bspline3")])
def test_two_ranks_equal_one(axis, shape):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(29611 + axis), os.path.join(ROOT, "tests", "mgpu_worker.py"), str(axis), "4", shape]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("axis", [0, 2])
def test_two_ranks_small_mailboxes(axis):
    """migration mailboxes of 16 particles against hundreds of leavers per step: the surplus waits a step in the
    ghost cells of the rank it left (gather and deposit still valid there), the boxes grow by the agreed rule --
    no error, no deadlock, and the run still equals the single-GPU run"""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(29651 + axis), os.path.join(ROOT, "tests", "mgpu_worker.py"), str(axis), "6", "triangle", "smallbox"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_two_ranks_moving_window():
    """moving window with the domain cut along x: insertion on the last rank, removal on the first,
    shifted fields exchanged through the halos -- equals the single-GPU run"""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29621", os.path.join(ROOT, "tests", "mgpu_worker.py"), "0", "8", "triangle", "window"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("axis", [0, 2])
def test_two_ranks_cpml(axis):
    """CPML layers on x and on the decomposed axis: each rank builds the profiles of its own part
    of the layer (set_cpml_helpers with its global offsets) -- equals the single-GPU run"""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(29631 + axis), os.path.join(ROOT, "tests", "mgpu_worker.py"), str(axis), "6", "triangle", "cpml"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("axis", [0, 2])
def test_two_ranks_rebalance(axis):
    """a lopsided plasma is re-sliced mid-run (get_load -> calculate_breaks -> redistribute of
    fields and particles over NCCL): the break moves, no particle is lost, the fullest rank gets
    lighter and the run still equals the single-GPU run"""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(29641 + axis), os.path.join(ROOT, "tests", "mgpu_worker.py"), str(axis), "6", "triangle", "balance"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
