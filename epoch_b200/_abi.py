"""ctypes binding of the C ABI in include/epoch_b200.h.

The same signature table binds the product (`libepoch_b200.so`, prefix
``eb200_``) and -- from tests/, smoke() and bench.py's cpu_baseline leg ONLY --
the CPU oracle (`oracle/liborc.so`, prefix ``orc_``).  There is no fallback: if
the CUDA library is missing, loading it raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
PRODUCT_LIB = os.environ.get("EB200_LIB", os.path.join(_HERE, "csrc", "libepoch_b200.so"))

c_ctx_p = C.c_void_p
i32, i64, f64 = C.c_int32, C.c_int64, C.c_double
pi32, pi64, pf64 = C.POINTER(i32), C.POINTER(i64), C.POINTER(f64)

# name -> (restype, argtypes)   (ctx first unless stated)
SIGNATURES = {
    "create": (i32, [C.POINTER(c_ctx_p), i32, i32, i32, i32, i32, pi32, pi32, C.c_void_p]),
    "nccl_unique_id": (i32, [C.c_void_p]),
    "destroy": (i32, [c_ctx_p]),
    "host_alloc": (i32, [C.POINTER(C.c_void_p), C.c_int64]),
    "host_free": (i32, [C.c_void_p]),
    "set_grid": (i32, [c_ctx_p, pi32, pi32, pi32, pf64, pf64, pf64, pf64, pf64, pf64, i32]),
    "set_time": (i32, [c_ctx_p, f64, f64, i32]),
    "set_field_bcs": (i32, [c_ctx_p, pi32]),
    "set_solver": (i32, [c_ctx_p, i32, i32, pf64]),
    "set_cpml": (i32, [c_ctx_p, i32, pf64, pf64, pf64, pf64, pf64, pf64, i32, i32, i32, i32, i32, i32]),
    "set_window_fields": (i32, [c_ctx_p, pf64, pf64, pf64, pf64, pf64, pf64]),
    "setup_field_boundaries": (i32, [c_ctx_p]),
    "set_boundary_fields": (i32, [c_ctx_p, i32, pf64, pf64, pf64, pf64, pf64, pf64]),
    "shift_window": (i32, [c_ctx_p, f64, f64, f64, f64, f64]),
    "set_current_options": (i32, [c_ctx_p, i32, i32, i32, i32, pi32, i32, pf64]),
    "set_species": (i32, [c_ctx_p, i32, f64, f64, f64, i32, i32, i32, pi32]),
    "upload_particles": (i32, [c_ctx_p, i32, i64, C.c_void_p, i32]),
    "append_particles": (i32, [c_ctx_p, i32, i64, C.c_void_p, i32]),
    "particle_count": (i32, [c_ctx_p, i32, pi64]),
    "download_particles": (i32, [c_ctx_p, i32, i64, i64, C.c_void_p, i32, pi64]),
    "upload_fields": (i32, [c_ctx_p] + [C.c_void_p] * 9),
    "download_fields": (i32, [c_ctx_p] + [C.c_void_p] * 9),
    "set_laser": (i32, [c_ctx_p, i32, i32, f64, f64, f64, f64, C.c_void_p, C.c_void_p]),
    "set_laser_step": (i32, [c_ctx_p, i32, i32, f64, f64]),
    "update_eb_fields_half": (i32, [c_ctx_p]),
    "push_particles": (i32, [c_ctx_p]),
    "current_finish": (i32, [c_ctx_p]),
    "update_eb_fields_final": (i32, [c_ctx_p]),
    "particle_bcs": (i32, [c_ctx_p]),
    "efield_bcs": (i32, [c_ctx_p]),
    "bfield_final_bcs": (i32, [c_ctx_p]),
    "energy": (i32, [c_ctx_p, pf64, pf64]),
    "charge_residual": (i32, [c_ctx_p, pf64, pf64]),
    "charge_residual_static": (i32, [c_ctx_p, pf64, pf64]),
    "calc_derived": (i32, [c_ctx_p, i32, i32, i32, pf64]),
    "set_thermal_boundary": (i32, [c_ctx_p, i32, i32, pf64, i64]),
    "set_probe": (i32, [c_ctx_p, i32, i32, pf64, pf64, f64, f64]),
    "download_probe": (i32, [c_ctx_p, i32, i32, C.c_int64, pf64, i32, pi64, i32]),
    "dist_fn": (i32, [c_ctx_p, i32, i32, pi32, pf64, pi32, pf64, pf64, pi32, pf64]),
    "get_load": (i32, [c_ctx_p, pf64, pi64, pi64, pi64]),
    "redistribute": (i32, [c_ctx_p, pi32, pi32, pi32, pi32, pf64, pf64, pf64, pf64]),
    "set_option": (i32, [c_ctx_p, C.c_char_p, C.c_char_p]),
    "sync": (i32, [c_ctx_p]),
    "timer": (i32, [c_ctx_p, C.c_char_p, i32, pf64, pi64]),
    "mark": (i32, [c_ctx_p, C.c_char_p, i32]),
    "last_error": (C.c_char_p, [c_ctx_p]),
    "version": (C.c_char_p, []),
}


class Abi:
    """A loaded shared library exposing the ABI under ``prefix``."""

    def __init__(self, path, prefix):
        if not os.path.exists(path):
            raise OSError(
                f"{path} not found: build it first (python -c 'import __graft_entry__ as g; g.build()')")
        self.path = path
        self.prefix = prefix
        self.lib = C.CDLL(path, mode=C.RTLD_GLOBAL if prefix == "eb200_" else C.RTLD_LOCAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(self.lib, prefix + name)  # AttributeError if a symbol is missing
            fn.restype = res
            fn.argtypes = args
            setattr(self, name, fn)


_product = None


def product():
    """The CUDA library.  Fails loudly when it has not been built."""
    global _product
    if _product is None:
        _product = Abi(PRODUCT_LIB, "eb200_")
    return _product
