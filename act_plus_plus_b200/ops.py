"""Per-kernel Python entry points over the C ABI (torch tensors in, raw pointers out).

Used by the parity tests and by host code that needs a single kernel; the full policy step goes through the
engine entry points (``act_engine_*``) instead.  Everything here requires CUDA tensors."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import (EPI_ATOMIC, EPI_BIAS, EPI_MASK_POS, EPI_OUT_BF16, EPI_OUT_F32, EPI_RELU, EPI_RES_BF16,
                   EPI_RES_F32, EPI_SCALE, Epilogue, check)


def _ptr(t: Optional[torch.Tensor]):
    if t is None:
        return None
    assert t.is_cuda, "act_plus_plus_b200 ops need CUDA tensors (no CPU fallback)"
    return t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def make_epilogue(out_f32=None, out_bf16=None, bias=None, scale=None, res_f32=None, res_bf16=None, aux_bf16=None,
                  relu=False, alpha=1.0, atomic=False, ldc=0, ldr=0, osz0=0, osz1=0, rsz0=0, rsz1=0) -> Epilogue:
    e = Epilogue()
    f = 0
    if out_f32 is not None: f |= EPI_OUT_F32
    if out_bf16 is not None: f |= EPI_OUT_BF16
    if bias is not None: f |= EPI_BIAS
    if scale is not None: f |= EPI_SCALE
    if res_f32 is not None: f |= EPI_RES_F32
    if res_bf16 is not None: f |= EPI_RES_BF16
    if aux_bf16 is not None: f |= EPI_MASK_POS
    if relu: f |= EPI_RELU
    if atomic: f |= EPI_ATOMIC
    e.flags, e.alpha = f, alpha
    e.bias, e.scale, e.res_f32, e.res_bf16 = _ptr(bias), _ptr(scale), _ptr(res_f32), _ptr(res_bf16)
    e.aux_bf16, e.out_f32, e.out_bf16 = _ptr(aux_bf16), _ptr(out_f32), _ptr(out_bf16)
    e.ldc, e.ldr, e.osz0, e.osz1, e.rsz0, e.rsz1 = ldc, ldr, osz0, osz1, rsz0, rsz1
    return e


def tc_gemm_raw(M, N, K, nz0, nz1, A, a_mn, lda, a_sz0, a_sz1, B, b_mn, ldb, b_sz0, b_sz1, epi: Epilogue,
                ksplit=1, bn=0, stages=0):
    rc = _lib.lib().act_tc_gemm(M, N, K, nz0, nz1, _ptr(A), a_mn, lda, a_sz0, a_sz1, _ptr(B), b_mn, ldb, b_sz0,
                                b_sz1, C.byref(epi), ksplit, bn, stages, _stream())
    check(rc, "act_tc_gemm")


def tc_gemm(a: torch.Tensor, b: torch.Tensor, a_mn=False, b_mn=False, out_dtype=torch.float32, bn=0, stages=0,
            ksplit=1, **epi_kw) -> torch.Tensor:
    """Batched D = A @ B^T.  a: [Z,M,K] (or [Z,K,M] if a_mn), b: [Z,N,K] (or [Z,K,N] if b_mn); bf16, contiguous."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    if a.dim() == 2: a = a[None]
    if b.dim() == 2: b = b[None]
    Z = a.shape[0]
    M, K = (a.shape[2], a.shape[1]) if a_mn else (a.shape[1], a.shape[2])
    N = b.shape[2] if b_mn else b.shape[1]
    out = torch.zeros((Z, M, N), dtype=out_dtype, device=a.device)
    kw = dict(epi_kw)
    if out_dtype == torch.float32: kw["out_f32"] = out
    else: kw["out_bf16"] = out
    epi = make_epilogue(ldc=N, osz0=M * N, ldr=kw.pop("ldr", N), rsz0=kw.pop("rsz0", M * N), **kw)
    tc_gemm_raw(M, N, K, Z, 1, a, int(a_mn), a.stride(1), a.stride(0), 0, b, int(b_mn), b.stride(1), b.stride(0), 0,
                epi, ksplit, bn, stages)
    return out


# ------------------------------------------------------------------------------------------- convolutions
def pack_conv_weights(w: torch.Tensor, scale: Optional[torch.Tensor] = None):
    """fp32 [Cout,Cin,KH,KW] -> (wpk [Cout, KK*Cin], wpkt [Cin, KK*Cout] * scale) bf16."""
    Cout, Cin, KH, KW = w.shape
    w = w.contiguous()
    wpk = torch.empty((Cout, KH * KW * Cin), dtype=torch.bfloat16, device=w.device)
    wpkt = torch.empty((Cin, KH * KW * Cout), dtype=torch.bfloat16, device=w.device)
    check(_lib.lib().act_pack_conv_weights(_ptr(w), _ptr(scale), _ptr(wpk), _ptr(wpkt), Cout, Cin, KH * KW, _stream()),
          "act_pack_conv_weights")
    return wpk, wpkt


def conv_fwd(x, wpk, Cout, KH, stride, pad, out_dtype=torch.bfloat16, **epi_kw):
    """x NHWC bf16 -> y NHWC."""
    N, H, W, Cin = x.shape
    Ho, Wo = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KH) // stride + 1
    y = torch.empty((N, Ho, Wo, Cout), dtype=out_dtype, device=x.device)
    kw = dict(epi_kw)
    kw["out_bf16" if out_dtype == torch.bfloat16 else "out_f32"] = y
    epi = make_epilogue(**kw)
    check(_lib.lib().act_conv_fwd(N, H, W, Cin, Cout, KH, stride, pad, _ptr(x), _ptr(wpk), C.byref(epi), _stream()),
          "act_conv_fwd")
    return y


def conv_dgrad(dy, wpkt, H, W, Cin, KH, stride, pad, res_compact=None, **epi_kw):
    """dy NHWC bf16 [N,Ho,Wo,Cout] -> dx NHWC bf16 [N,H,W,Cin]."""
    N, Ho, Wo, Cout = dy.shape
    dx = torch.zeros((N, H, W, Cin), dtype=torch.bfloat16, device=dy.device)
    epi = make_epilogue(out_bf16=dx, **epi_kw)
    check(_lib.lib().act_conv_dgrad(N, H, W, Cin, Cout, KH, stride, pad, _ptr(dy), _ptr(wpkt), C.byref(epi),
                                    _ptr(res_compact), _stream()), "act_conv_dgrad")
    return dx


def conv_wgrad(x, dy, KH, stride, pad, alpha=1.0, scale=None, dw=None):
    N, H, W, Cin = x.shape
    Cout = dy.shape[3]
    if dw is None:
        dw = torch.zeros((Cout, Cin, KH, KH), dtype=torch.float32, device=x.device)
    check(_lib.lib().act_conv_wgrad(N, H, W, Cin, Cout, KH, stride, pad, _ptr(x), _ptr(dy), _ptr(dw), alpha,
                                    _ptr(scale), _stream()), "act_conv_wgrad")
    return dw
