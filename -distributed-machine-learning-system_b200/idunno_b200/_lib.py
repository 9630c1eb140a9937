"""ctypes binding of csrc/libidunno_b200.so (the C ABI declared in include/idunno_b200.h).

The library is the only implementation: when it is missing, or when it cannot run (no sm_100
device), every call raises -- there is no CPU or PyTorch fallback behind this module.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "csrc", "libidunno_b200.so")

IDN_MODEL_ALEXNET, IDN_MODEL_RESNET18 = 0, 1
IDN_IMPL_TC, IDN_IMPL_SIMT = 0, 1

_lib = None

# name -> (restype, argtypes); must list every symbol of include/idunno_b200.h
SIGNATURES = {
    "idn_create": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)]),
    "idn_destroy": (C.c_int, [C.c_void_p]),
    "idn_weight_count": (C.c_int64, [C.c_int]),
    "idn_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "idn_set_impl": (C.c_int, [C.c_void_p, C.c_int]),
    "idn_infer_u8hwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p]),
    "idn_submit_u8hwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                   C.POINTER(C.c_int)]),
    "idn_collect": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "idn_infer_u8hwc_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    "idn_sync": (C.c_int, [C.c_void_p]),
    "idn_abort": (C.c_int, [C.c_void_p]),
    "idn_preprocess_u8hwc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "idn_preprocess_u8hwc_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                               C.c_void_p, C.c_void_p]),
    "idn_num_steps": (C.c_int, [C.c_void_p]),
    "idn_step_info": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.c_void_p]),
    "idn_read_activation": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "idn_step_kernel": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p, C.c_int]),
    "idn_debug_stop_after": (C.c_int, [C.c_void_p, C.c_int]),
    "idn_kernel_launches": (C.c_int64, [C.c_void_p]),
    "idn_h2d_bytes": (C.c_int64, [C.c_void_p]),
    "idn_debug_needed_rows": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "idn_set_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "idn_step_wait_cycles": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "idn_step_time_ms": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float)]),
    "idn_conv2d_nchw": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "idn_host_alloc": (C.c_int, [C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]),
    "idn_host_free": (C.c_int, [C.c_void_p]),
    "idn_last_error": (C.c_char_p, []),
    "idn_version": (C.c_char_p, []),
}


class IdunnoError(RuntimeError):
    pass


def load_library():
    """dlopen the in-tree shared library and type its entry points.  Needs no GPU."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise IdunnoError(
            f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
            "(idunno_b200 has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load_library().idn_last_error()
        raise IdunnoError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
