"""ctypes binding of libb200slu.so (the C ABI declared in include/b200slu.h).

The library is built in-tree by ``espnet_slurp_b200.build`` and loaded from this
directory.  There is no fallback: if the shared object is missing or fails to load the
import raises, and every op raises ``RuntimeError`` with the library's own message on a
non-zero status.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200slu.so")


class Epilogue(C.Structure):
    """b200slu_epilogue_t"""

    _fields_ = [
        ("bias", C.c_void_p),
        ("act", C.c_int),
        ("pre_act_out", C.c_void_p),
        ("alpha", C.c_float),
        ("dropout_p", C.c_float),
        ("dropout_seed", C.c_ulonglong),
        ("dropout_site", C.c_uint),
        ("act_grad_input", C.c_void_p),
        ("ld_act_grad_input", C.c_longlong),
        ("act_grad", C.c_int),
        ("residual", C.c_void_p),
        ("ld_residual", C.c_longlong),
        ("out_f32", C.c_int),
        ("accumulate", C.c_int),
    ]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -m espnet_slurp_b200.build` "
                "(there is no CPU or PyTorch fallback for the B200 kernels)")
        _lib = C.CDLL(LIB_PATH)
        _lib.b200slu_last_error.restype = C.c_char_p
        _lib.b200slu_prof_family_name.restype = C.c_char_p
        _lib.b200slu_prof_total_launches.restype = C.c_longlong
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().b200slu_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libb200slu {what} failed ({rc}): {msg}")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
