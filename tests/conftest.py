import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_addoption(parser):
    parser.addoption("--ref-dump", action="store", default=os.environ.get("EB200_REF_DUMP"),
                     help="directory of reference dumps made by tools/ref_build/build_and_run.sh (EPOCH's own output)")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _oracle_path():
    p = os.path.join(ROOT, "oracle", "liborc.so")
    if not os.path.exists(p):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    return p


@pytest.fixture(scope="session")
def oracle_abi():
    """The CPU oracle (checker only)."""
    from epoch_b200 import _abi
    return _abi.Abi(_oracle_path(), "orc_")


@pytest.fixture(scope="session")
def product_abi():
    from epoch_b200 import _abi
    return _abi.product()
