"""Implicit-GEMM convolution kernels (fwd / dgrad / wgrad) vs torch conv2d + autograd on the same bf16 operands."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# (N, H, W, Cin, Cout, KH, stride, pad): every resnet18 layer1-4 flavour, at reduced and at full ACT resolution
SHAPES = [
    (2, 24, 32, 64, 64, 3, 1, 1), (2, 24, 32, 64, 128, 3, 2, 1), (2, 24, 32, 64, 128, 1, 2, 0),
    (3, 12, 16, 128, 128, 3, 1, 1), (2, 12, 16, 128, 256, 3, 2, 1), (2, 6, 8, 256, 512, 3, 2, 1),
    (2, 6, 8, 256, 512, 1, 2, 0), (3, 3, 4, 512, 512, 3, 1, 1),
    (2, 15, 20, 512, 512, 3, 1, 1), (1, 30, 40, 256, 256, 3, 1, 1), (1, 120, 160, 64, 64, 3, 1, 1),
    (1, 120, 160, 64, 128, 3, 2, 1), (2, 30, 40, 256, 512, 3, 2, 1),
]


def _mk(shape, seed=0):
    N, H, W, Cin, Cout, KH, s, p = shape
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn((N, H, W, Cin), generator=g, device="cuda").bfloat16()
    w = (torch.randn((Cout, Cin, KH, KH), generator=g, device="cuda") * (2.0 / (Cin * KH * KH)) ** 0.5)
    return x, w


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


@pytest.mark.parametrize("shape", SHAPES)
def test_conv_fwd(shape):
    from act_plus_plus_b200 import ops
    N, H, W, Cin, Cout, KH, s, p = shape
    x, w = _mk(shape)
    g = torch.Generator(device="cuda").manual_seed(9)
    scale = torch.rand(Cout, generator=g, device="cuda") + 0.5
    bias = torch.randn(Cout, generator=g, device="cuda") * 0.1
    wpk, _ = ops.pack_conv_weights(w)
    wb = w.bfloat16().float()
    ref0 = F.conv2d(x.float().permute(0, 3, 1, 2), wb, stride=s, padding=p).permute(0, 2, 3, 1)
    y = ops.conv_fwd(x, wpk, Cout, KH, s, p, out_dtype=torch.float32)
    torch.cuda.synchronize()
    assert (y - ref0).abs().max().item() < 2e-2, f"plain conv {shape}: {(y - ref0).abs().max().item()}"
    res = torch.randn(ref0.shape, generator=g, device="cuda").bfloat16()
    y = ops.conv_fwd(x, wpk, Cout, KH, s, p, scale=scale, bias=bias, res_bf16=res, relu=True)
    ref = torch.relu(ref0 * scale + bias + res.float())
    torch.cuda.synchronize()
    assert _rel(y, ref) < 6e-3, f"fused conv {shape}: rel {_rel(y, ref)}"


@pytest.mark.parametrize("shape", SHAPES)
def test_conv_dgrad(shape):
    from act_plus_plus_b200 import ops
    N, H, W, Cin, Cout, KH, s, p = shape
    x, w = _mk(shape, 1)
    g = torch.Generator(device="cuda").manual_seed(10)
    scale = torch.rand(Cout, generator=g, device="cuda") + 0.5
    Ho, Wo = (H + 2 * p - KH) // s + 1, (W + 2 * p - KH) // s + 1
    dy = torch.randn((N, Ho, Wo, Cout), generator=g, device="cuda").bfloat16()
    _, wpkt = ops.pack_conv_weights(w, scale)
    ws = (w * scale[:, None, None, None]).bfloat16().float()
    ref = F.conv_transpose2d(dy.float().permute(0, 3, 1, 2), ws, stride=s, padding=p,
                             output_padding=(H - ((Ho - 1) * s - 2 * p + KH), W - ((Wo - 1) * s - 2 * p + KH))).permute(0, 2, 3, 1)
    mask = torch.randn((N, H, W, Cin), generator=g, device="cuda").bfloat16()
    if s == 1:
        res = torch.randn((N, H, W, Cin), generator=g, device="cuda").bfloat16()
        dx = ops.conv_dgrad(dy, wpkt, H, W, Cin, KH, s, p, res_bf16=res, aux_bf16=mask)
        ref = (ref + res.float()) * (mask.float() > 0)
    else:
        resc = torch.randn((N, H // 2, W // 2, Cin), generator=g, device="cuda").bfloat16()
        dx = ops.conv_dgrad(dy, wpkt, H, W, Cin, KH, s, p, res_compact=resc, aux_bf16=mask)
        ref = ref.clone()
        ref[:, ::2, ::2] += resc.float()
        ref = ref * (mask.float() > 0)
    torch.cuda.synchronize()
    assert _rel(dx, ref) < 6e-3, f"dgrad {shape}: rel {_rel(dx, ref)}"


@pytest.mark.parametrize("shape", SHAPES)
def test_conv_wgrad(shape):
    from act_plus_plus_b200 import ops
    N, H, W, Cin, Cout, KH, s, p = shape
    x, w = _mk(shape, 2)
    g = torch.Generator(device="cuda").manual_seed(11)
    scale = torch.rand(Cout, generator=g, device="cuda") + 0.5
    Ho, Wo = (H + 2 * p - KH) // s + 1, (W + 2 * p - KH) // s + 1
    dy = torch.randn((N, Ho, Wo, Cout), generator=g, device="cuda").bfloat16()
    dw = ops.conv_wgrad(x, dy, KH, s, p, alpha=0.5, scale=scale)
    ref = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), w.shape, dy.float().permute(0, 3, 1, 2), stride=s, padding=p)
    ref = ref * 0.5 * scale[:, None, None, None]
    torch.cuda.synchronize()
    assert _rel(dw, ref) < 2e-3, f"wgrad {shape}: rel {_rel(dw, ref)}"
