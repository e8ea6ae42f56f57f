"""-m "not gpu": the C-ABI library loads and exports every symbol include/epoch_b200.h declares."""
import ctypes
import os
import re

from epoch_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "epoch_b200.h")).read()
    return sorted(set(re.findall(r"EB200_NAME\((\w+)\)\(", src)))


def test_header_and_binding_table_agree():
    assert _declared() == sorted(_abi.SIGNATURES)


def test_product_library_exports_all_symbols():
    lib = ctypes.CDLL(_abi.PRODUCT_LIB)
    for name in _declared():
        assert hasattr(lib, "eb200_" + name), name
    assert b"sm_100a" in ctypes.cast(ctypes.CDLL(_abi.PRODUCT_LIB).eb200_version, ctypes.CFUNCTYPE(ctypes.c_char_p))()


def test_oracle_exports_all_symbols(oracle_abi):
    for name in _declared():
        assert hasattr(oracle_abi, name)


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to create a context."""
    import torch
    if torch.cuda.is_available():
        return
    from epoch_b200 import PicContext, EpochError
    try:
        PicContext(1)
    except EpochError as e:
        assert "no CUDA device" in str(e) or "rc=" in str(e)
    else:
        raise AssertionError("context creation must fail without a GPU")
