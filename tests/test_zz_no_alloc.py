"""act_conv_wgrad_noalloc (include/act_b200.h): the allocation-free variant of act_conv_wgrad whose staging tile is a static
device array.  No counterpart in the reference -- added for callers that cannot call
cudaMalloc.  Checked through the C-ABI against the existing entry point, on consecutive calls (the shared scratch keeps the
previous call's contents, so a missing clear shows up on the second call).  Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 9 * 256 * 256          # ACT_CONV_WGRAD_NOALLOC_MAX_FLOATS


def _rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-20))


def _case(shape, seed):
    N, H, W, Cin, Cout, KH, s, p = shape
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn((N, H, W, Cin), generator=g, device="cuda").bfloat16()
    Ho, Wo = (H + 2 * p - KH) // s + 1, (W + 2 * p - KH) // s + 1
    dy = torch.randn((N, Ho, Wo, Cout), generator=g, device="cuda").bfloat16()
    scale = torch.rand(Cout, generator=g, device="cuda") + 0.5
    dw0 = torch.randn((Cout, Cin, KH, KH), generator=g, device="cuda")      # both entry points ACCUMULATE into dw
    return x, dy, scale, dw0


def _call(fn, shape, x, dy, scale, dw, alpha):
    N, H, W, Cin, Cout, KH, s, p = shape
    return fn(N, H, W, Cin, Cout, KH, s, p, x.data_ptr(), dy.data_ptr(), dw.data_ptr(), C.c_float(alpha), scale.data_ptr(),
              torch.cuda.current_stream().cuda_stream)


def test_conv_wgrad_noalloc_matches_allocating_entry_point():
    from act_plus_plus_b200 import _lib
    lib = _lib.lib()
    # largest supported tile first (9*256*256 floats, exactly the limit), then smaller ones that reuse its dirty prefix
    shapes = [(1, 30, 40, 256, 256, 3, 1, 1), (2, 24, 32, 64, 128, 3, 2, 1), (2, 24, 32, 64, 128, 1, 2, 0),
              (1, 30, 40, 256, 256, 3, 1, 1)]
    cases = [_case(sh, 40 + i) for i, sh in enumerate(shapes)]
    refs, outs = [], []
    for sh, (x, dy, scale, dw0) in zip(shapes, cases):
        ref = dw0.clone()
        assert _call(lib.act_conv_wgrad, sh, x, dy, scale, ref, 0.5) == 0, _lib.last_error()
        refs.append(ref)
        outs.append(dw0.clone())
    torch.cuda.synchronize()
    free_before = None
    for i, (sh, (x, dy, scale, _)) in enumerate(zip(shapes, cases)):       # consecutive calls, same stream, no sync between
        if i == len(shapes) - 1:                                             # a shape already run once: nothing lazy is left
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        assert _call(lib.act_conv_wgrad_noalloc, sh, x, dy, scale, outs[i], 0.5) == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before, "the allocation-free entry point took device memory"
    for sh, out, ref, (_, _, _, dw0) in zip(shapes, outs, refs, cases):
        assert torch.isfinite(out).all()
        assert _rel(out - dw0, ref - dw0) < 2e-3, f"noalloc wgrad {sh}: rel {_rel(out - dw0, ref - dw0)}"


def test_conv_wgrad_noalloc_rejects_oversize():
    from act_plus_plus_b200 import _lib
    lib = _lib.lib()
    sh = (1, 15, 20, 512, 512, 3, 1, 1)                                      # 9*512*512 floats > the static scratch
    assert sh[3] * sh[4] * 9 > MAX_FLOATS
    x, dy, scale, dw0 = _case(sh, 7)
    dw = dw0.clone()
    assert _call(lib.act_conv_wgrad_noalloc, sh, x, dy, scale, dw, 1.0) != 0
    msg = _lib.last_error()
    assert "act_conv_wgrad_noalloc" in msg and "exceeds" in msg, msg
    torch.cuda.synchronize()
    assert torch.equal(dw, dw0), "a rejected call must not touch its output"
    # and the supported path still works afterwards
    sh2 = (2, 12, 16, 128, 256, 3, 2, 1)
    x, dy, scale, dw0 = _case(sh2, 8)
    a, b = dw0.clone(), dw0.clone()
    assert _call(lib.act_conv_wgrad, sh2, x, dy, scale, a, 1.0) == 0, _lib.last_error()
    assert _call(lib.act_conv_wgrad_noalloc, sh2, x, dy, scale, b, 1.0) == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert _rel(b - dw0, a - dw0) < 2e-3
