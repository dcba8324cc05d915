"""ctypes binding of libact_b200.so (the C ABI declared in include/act_b200.h).

Fails loudly when the library is missing: there is no CPU or PyTorch fallback for the compute path."""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libact_b200.so")


class ActB200Error(RuntimeError):
    pass


class Epilogue(C.Structure):
    _fields_ = [
        ("flags", C.c_uint32), ("alpha", C.c_float),
        ("bias", C.c_void_p), ("scale", C.c_void_p), ("res_f32", C.c_void_p), ("res_bf16", C.c_void_p),
        ("aux_bf16", C.c_void_p), ("out_f32", C.c_void_p), ("out_bf16", C.c_void_p),
        ("ldc", C.c_int64), ("ldr", C.c_int64), ("osz0", C.c_int64), ("osz1", C.c_int64),
        ("rsz0", C.c_int64), ("rsz1", C.c_int64),
    ]


EPI_BIAS, EPI_SCALE, EPI_RES_F32, EPI_RES_BF16, EPI_RELU, EPI_MASK_POS, EPI_OUT_F32, EPI_OUT_BF16, EPI_ATOMIC = (
    1, 2, 4, 8, 16, 32, 64, 128, 256)

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ActB200Error(
                f"{LIB_PATH} is missing: build it with `python -m act_plus_plus_b200._build` "
                "(or __graft_entry__.build()).  There is no CPU fallback.")
        _lib = C.CDLL(LIB_PATH)
        _lib.act_last_error.restype = C.c_char_p
        _declare(_lib)
    return _lib


def _declare(l):
    from ._abi import PROTOTYPES
    for name, (restype, argtypes) in PROTOTYPES.items():
        fn = getattr(l, name)        # AttributeError if the .so does not export a declared symbol
        fn.restype = restype
        fn.argtypes = argtypes


def last_error() -> str:
    return lib().act_last_error().decode(errors="replace")


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().act_last_error().decode(errors="replace")
        raise ActB200Error(f"{what} failed (code {rc}): {msg}")
