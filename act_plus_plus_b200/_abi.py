"""Prototype table of every symbol include/act_b200.h declares (used by the loader and by the CPU test
that checks the library exports all of them)."""
import ctypes as C

P, I, I64, F = C.c_void_p, C.c_int, C.c_int64, C.c_float

PROTOTYPES = {
    "act_abi_version": (I, []),
    "act_tc_gemm": (I, [I, I, I, I, I, P, I, I64, I64, I64, P, I, I64, I64, I64, P, I, I, I, P]),
    "act_cast_bf16": (I, [P, P, I64, P]),
    "act_pack_conv_weights": (I, [P, P, P, P, I, I, I, P]),
    "act_conv_fwd": (I, [I, I, I, I, I, I, I, I, P, P, P, P]),
    "act_conv_dgrad": (I, [I, I, I, I, I, I, I, I, P, P, P, P, P]),
    "act_conv_wgrad": (I, [I, I, I, I, I, I, I, I, P, P, P, F, P, P]),
}
