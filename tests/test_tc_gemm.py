"""tcgen05 GEMM kernel vs torch fp32 matmul of the same bf16-rounded operands (GPU)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = a.float().transpose(1, 2) if a_mn else a.float()
    B = b.float().transpose(1, 2) if b_mn else b.float()
    return torch.bmm(A, B.transpose(1, 2))


def _mk(Z, M, N, K, a_mn, b_mn, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = torch.randn((Z, K, M) if a_mn else (Z, M, K), generator=g, device="cuda").bfloat16()
    b = torch.randn((Z, K, N) if b_mn else (Z, N, K), generator=g, device="cuda").bfloat16()
    return a, b


def _check(out, ref, K, what):
    err = (out.float() - ref).abs().max().item()
    tol = 2e-3 * (K ** 0.5) if out.dtype == torch.float32 else 3e-2 * (K ** 0.5)
    assert err <= tol, f"{what}: max abs err {err:.4g} > {tol:.4g}"


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("shape", [(1, 128, 128, 64), (1, 256, 512, 512), (3, 304, 304, 64), (2, 304, 64, 304),
                                   (1, 2416, 3200, 512), (1, 520, 512, 3200)])
def test_gemm_majors(a_mn, b_mn, shape):
    from act_plus_plus_b200 import ops
    Z, M, N, K = shape
    a, b = _mk(Z, M, N, K, a_mn, b_mn)
    out = ops.tc_gemm(a, b, bool(a_mn), bool(b_mn))
    torch.cuda.synchronize()
    _check(out, _ref(a, b, a_mn, b_mn), K, f"majors a_mn={a_mn} b_mn={b_mn} {shape}")


@pytest.mark.parametrize("bn", [64, 128, 256])
@pytest.mark.parametrize("stages", [2, 4])
def test_gemm_tiles(bn, stages):
    from act_plus_plus_b200 import ops
    a, b = _mk(2, 300, 520, 200, 0, 0)
    out = ops.tc_gemm(a, b, bn=bn, stages=stages)
    torch.cuda.synchronize()
    _check(out, _ref(a, b, 0, 0), 200, f"bn={bn} stages={stages}")


def test_gemm_ragged_attention_shapes():
    """Strided head views as used by attention: Q/K [B*L, 1024] -> S [B,h,L,Lpad]; P.V with K = L (not /64)."""
    from act_plus_plus_b200 import ops
    B, h, L, d = 2, 8, 302, 64
    Lp = 304
    g = torch.Generator(device="cuda").manual_seed(1)
    qk = torch.randn((B * L, 2 * h * d), generator=g, device="cuda").bfloat16()
    S = torch.zeros((B, h, L, Lp), device="cuda")
    epi = ops.make_epilogue(out_f32=S, alpha=0.125, ldc=Lp, osz0=L * Lp, osz1=h * L * Lp)
    ops.tc_gemm_raw(L, L, d, h, B, qk, 0, 2 * h * d, d, L * 2 * h * d, qk[:, h * d:], 0, 2 * h * d, d, L * 2 * h * d, epi)
    torch.cuda.synchronize()
    q = qk[:, :h * d].float().reshape(B, L, h, d).permute(0, 2, 1, 3)
    k = qk[:, h * d:].float().reshape(B, L, h, d).permute(0, 2, 1, 3)
    ref = torch.matmul(q, k.transpose(2, 3)) * 0.125
    assert (S[..., :L] - ref).abs().max().item() < 2e-2
    assert S[..., L:].abs().max().item() == 0.0
    # O = P V : A = P [L, Lp] K-major with K = L, B = V MN-major strided head view
    P = torch.softmax(S[..., :L], -1)
    Pb = torch.zeros((B, h, L, Lp), device="cuda", dtype=torch.bfloat16)
    Pb[..., :L] = P.bfloat16()
    v = torch.randn((B * L, h * d), generator=g, device="cuda").bfloat16()
    O = torch.zeros((B * L, h * d), device="cuda", dtype=torch.bfloat16)
    epi = ops.make_epilogue(out_bf16=O, ldc=h * d, osz0=d, osz1=L * h * d)
    ops.tc_gemm_raw(L, d, L, h, B, Pb, 0, Lp, L * Lp, h * L * Lp, v, 1, h * d, d, L * h * d, epi)
    torch.cuda.synchronize()
    vr = v.float().reshape(B, L, h, d).permute(0, 2, 1, 3)
    ref = torch.matmul(Pb[..., :L].float(), vr).permute(0, 2, 1, 3).reshape(B * L, h * d)
    assert (O.float() - ref).abs().max().item() < 2e-2


def test_gemm_epilogue():
    from act_plus_plus_b200 import ops
    Z, M, N, K = 1, 300, 200, 128
    a, b = _mk(Z, M, N, K, 0, 0, seed=3)
    g = torch.Generator(device="cuda").manual_seed(4)
    bias = torch.randn(N, generator=g, device="cuda")
    scale = torch.rand(N, generator=g, device="cuda") + 0.5
    res = torch.randn((Z, M, N), generator=g, device="cuda")
    aux = torch.randn((Z, M, N), generator=g, device="cuda").bfloat16()
    base = _ref(a, b, 0, 0)
    out = ops.tc_gemm(a, b, bias=bias, scale=scale, res_f32=res, relu=True, alpha=0.5)
    ref = torch.relu(base * 0.5 * scale + bias + res)
    _check(out, ref, K, "bias/scale/res/relu f32")
    out = ops.tc_gemm(a, b, out_dtype=torch.bfloat16, bias=bias, res_bf16=res.bfloat16(), aux_bf16=aux)
    ref = (base + bias + res.bfloat16().float()) * (aux.float() > 0)
    _check(out, ref, K, "bf16 out / bf16 res / mask")
    # split-K with atomics accumulates onto the existing output
    Z, M, N, K = 1, 128, 192, 2416
    a, b = _mk(Z, M, N, K, 1, 1, seed=5)
    out = torch.ones((Z, M, N), device="cuda")
    epi = ops.make_epilogue(out_f32=out, atomic=True, ldc=N, osz0=M * N)
    ops.tc_gemm_raw(M, N, K, 1, 1, a, 1, a.stride(1), 0, 0, b, 1, b.stride(1), 0, 0, epi, ksplit=5)
    torch.cuda.synchronize()
    _check(out, _ref(a, b, 1, 1) + 1.0, K, "split-K atomic")
