"""Prototype table of every symbol include/act_b200.h declares (used by the loader and by the CPU test
that checks the library exports all of them)."""
import ctypes as C

P, I, I64, F, D = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double

PROTOTYPES = {
    "act_abi_version": (I, []),
    "act_tc_gemm": (I, [I, I, I, I, I, P, I, I64, I64, I64, P, I, I64, I64, I64, P, I, I, I, P]),
    "act_cast_bf16": (I, [P, P, I64, P]),
    "act_pack_conv_weights": (I, [P, P, P, P, I, I, I, P]),
    "act_conv_fwd": (I, [I, I, I, I, I, I, I, I, P, P, P, P]),
    "act_conv_dgrad": (I, [I, I, I, I, I, I, I, I, P, P, P, P, P]),
    "act_conv_wgrad": (I, [I, I, I, I, I, I, I, I, P, P, P, F, P, P]),
    "act_ln_fwd": (I, [P, P, P, P, I, P, P, P, P, P, I, I, P]),
    "act_ln_bwd": (I, [P, P, P, P, P, P, P, P, P, P, I, I, P]),
    "act_add_pos_cast": (I, [P, P, I, P, P, I, I, P]),
    "act_softmax_fwd": (I, [P, P, P, I, I, I, I, P]),
    "act_softmax_bwd": (I, [P, P, P, I, I, I, F, P]),
    "act_colsum_bf16": (I, [P, I64, I, I, P, P]),
    "act_sgemm": (I, [I, I, I, P, P, P, P, P, P, P, P, P, P, F, I, P]),
    "act_loss": (I, [P, P, I64, P, I64, P, P, I, I, I, I, F, F, P, P, P, P, P]),
    "act_reparam_fwd": (I, [P, P, P, P, P, I, I, P]),
    "act_reparam_bwd": (I, [P, P, P, P, P, P, I, I, P]),
    "act_maxpool_fwd": (I, [P, P, P, I, I, I, I, P]),
    "act_maxpool_bwd": (I, [P, P, P, P, I, I, I, I, P]),
    "act_im2col_conv1": (I, [P, I64, P, I, I, I, I, P]),
    "act_adamw_step": (I, [P, P, P, P, P, I64, F, F, F, F, F, I, I, P]),
    "act_temporal_agg_step": (I, [P, P, P, P, I, I, I, D, P]),
}
