"""Prototype table of every symbol include/act_b200.h declares (used by the loader and by the CPU test
that checks the library exports all of them)."""
import ctypes as C

P, I, I64, F = C.c_void_p, C.c_int, C.c_int64, C.c_float

PROTOTYPES = {
    "act_abi_version": (I, []),
    "act_tc_gemm": (I, [I, I, I, I, I, P, I, I64, I64, I64, P, I, I64, I64, I64, P, I, I, I, P]),
}
