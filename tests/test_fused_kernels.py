"""Fused bandwidth kernels through the C ABI vs torch fp32 on the GPU; temporal ensemble vs the literal oracle."""
import ctypes as C
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def L():
    from act_plus_plus_b200 import _lib
    return _lib.lib()


def ck(rc):
    from act_plus_plus_b200 import _lib
    _lib.check(rc, "kernel")


def p(t):
    return None if t is None else t.data_ptr()


def st():
    return torch.cuda.current_stream().cuda_stream


def rnd(*shape, seed=0, dtype=torch.float32):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(shape, generator=g, device="cuda").to(dtype)


@pytest.mark.parametrize("M,Lr", [(7, 3), (2416, 302), (816, 102)])
def test_layernorm_fwd_bwd(M, Lr):
    D = 512
    t, gamma, beta, pos = rnd(M, D, seed=1) * 2 + 0.3, rnd(D, seed=2) * 0.1 + 1, rnd(D, seed=3) * 0.1, rnd(Lr, D, seed=4)
    y, yb, ypb = torch.empty(M, D, device="cuda"), torch.empty(M, D, device="cuda", dtype=torch.bfloat16), torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
    ck(L().act_ln_fwd(p(t), p(gamma), p(beta), p(pos), Lr, p(y), p(yb), p(ypb), p(mean), p(rstd), M, D, st()))
    tr = t.clone().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    ref = F.layer_norm(tr, (D,), gr, br, 1e-5)
    assert (y - ref).abs().max().item() < 2e-5
    assert (yb.float() - ref).abs().max().item() < 3e-2
    posr = pos.repeat((M + Lr - 1) // Lr, 1)[:M]
    assert (ypb.float() - (ref + posr)).abs().max().item() < 5e-2
    dy, dy2 = rnd(M, D, seed=5), rnd(M, D, seed=6)
    ref.backward(dy + dy2)
    dt, dtb = torch.empty(M, D, device="cuda"), torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    dg, db = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
    ck(L().act_ln_bwd(p(dy), p(dy2), p(t), p(mean), p(rstd), p(gamma), p(dt), p(dtb), p(dg), p(db), M, D, st()))
    assert (dt - tr.grad).abs().max().item() < 1e-4
    assert (dg - gr.grad).abs().max().item() < 2e-3 * max(1.0, gr.grad.abs().max().item())
    assert (db - br.grad).abs().max().item() < 2e-3 * max(1.0, br.grad.abs().max().item())
    assert (dtb.float() - tr.grad).abs().max().item() < 5e-2


def test_add_pos_cast():
    M, D, Lr = 604, 512, 302
    x, pos = rnd(M, D, seed=1), rnd(Lr, D, seed=2)
    xb, xpb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16), torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    ck(L().act_add_pos_cast(p(x), p(pos), Lr, p(xb), p(xpb), M, D, st()))
    assert torch.equal(xb, x.bfloat16())
    assert torch.equal(xpb, (x + pos.repeat(2, 1)).bfloat16())


@pytest.mark.parametrize("B,h,Lq,Lk", [(2, 8, 102, 102), (2, 8, 100, 302), (1, 8, 50, 902)])
def test_softmax_fwd_bwd(B, h, Lq, Lk):
    Lp = (Lk + 7) // 8 * 8
    R = B * h * Lq
    S = torch.zeros(R, Lp, device="cuda")
    S[:, :Lk] = rnd(R, Lk, seed=1) * 3
    S[:, Lk:] = 1e9   # pad garbage must be ignored
    mask = torch.zeros(B, Lk, dtype=torch.uint8, device="cuda")
    mask[0, Lk // 2:] = 1
    mask[-1, 5:9] = 1
    P = torch.empty(R, Lp, device="cuda", dtype=torch.bfloat16)
    ck(L().act_softmax_fwd(p(S), p(mask), p(P), R, Lk, Lp, h * Lq, st()))
    mb = mask.bool()[:, None, :].expand(B, h * Lq, Lk).reshape(R, Lk)
    Sr = S[:, :Lk].masked_fill(mb, float("-inf")).clone().requires_grad_(True)
    ref = torch.softmax(Sr, -1)
    assert (P[:, :Lk].float() - ref).abs().max().item() < 4e-3
    assert P[:, Lk:].abs().max().item() == 0 if Lp > Lk else True
    dP = torch.zeros(R, Lp, device="cuda")
    dP[:, :Lk] = rnd(R, Lk, seed=2)
    dS = torch.empty(R, Lp, device="cuda", dtype=torch.bfloat16)
    ck(L().act_softmax_bwd(p(P), p(dP), p(dS), R, Lk, Lp, 0.125, st()))
    Pf = P[:, :Lk].float()
    refd = 0.125 * Pf * (dP[:, :Lk] - (Pf * dP[:, :Lk]).sum(-1, keepdim=True))
    assert (dS[:, :Lk].float() - refd).abs().max().item() < 1e-2 * max(1.0, refd.abs().max().item())


def test_colsum():
    M, N = 2416, 3200
    x = rnd(M, N, seed=1, dtype=torch.bfloat16)
    out = torch.ones(N, device="cuda")
    ck(L().act_colsum_bf16(p(x), N, M, N, p(out), st()))
    ref = x.float().sum(0) + 1
    assert (out - ref).abs().max().item() < 2e-2


def _map(rpb, bs, s):
    return (C.c_int64 * 3)(rpb, bs, s)


def test_sgemm_token_rows_and_reduction():
    B, nq, A, D = 3, 100, 16, 512
    x, w, b = rnd(B * nq, A, seed=1), rnd(D, A, seed=2), rnd(D, seed=3)
    tok = torch.zeros(B, nq + 2, D, device="cuda")
    BIG = 1 << 30
    ck(L().act_sgemm(B * nq, D, A, p(x), _map(BIG, 0, A), _map(BIG, 0, 1), p(w), _map(BIG, 0, 1), _map(BIG, 0, A),
                     tok.data_ptr() + 2 * D * 4, _map(nq, (nq + 2) * D, D), _map(BIG, 0, 1), p(b), 1.0, 0, st()))
    ref = F.linear(x, w, b).reshape(B, nq, D)
    assert (tok[:, 2:] - ref).abs().max().item() < 1e-4
    assert tok[:, :2].abs().max().item() == 0
    # dW[n,k] += sum_m dY(m,n) X(m,k) with dY rows living in the token buffer, split reduction
    dtok = rnd(B, nq + 2, D, seed=4)
    dw = torch.zeros(D, A, device="cuda")
    ck(L().act_sgemm(D, A, B * nq, dtok.data_ptr() + 2 * D * 4, _map(BIG, 0, 1), _map(nq, (nq + 2) * D, D), p(x),
                     _map(BIG, 0, A), _map(BIG, 0, 1), p(dw), _map(BIG, 0, A), _map(BIG, 0, 1), None, 1.0, 1, st()))
    ref = dtok[:, 2:].reshape(-1, D).t() @ x
    assert (dw - ref).abs().max().item() < 1e-3


def test_loss_and_reparam():
    B, nq, A, Z, T = 5, 100, 16, 32, 120
    a_hat, actions = rnd(B, nq, A, seed=1), rnd(B, T, A, seed=2)
    actions[0, 3, 5] = a_hat[0, 3, 5]            # exact tie -> zero subgradient
    is_pad = (torch.arange(T, device="cuda")[None] >= torch.tensor([100, 37, 1, 64, 99], device="cuda")[:, None])
    pad8 = is_pad.to(torch.uint8)
    info, eps = rnd(B, 2 * Z, seed=3) * 0.5, rnd(B, Z, seed=4)
    mu, lv, z = (torch.empty(B, Z, device="cuda") for _ in range(3))
    ck(L().act_reparam_fwd(p(info), p(eps), p(mu), p(lv), p(z), B, Z, st()))
    assert (z - (info[:, :Z] + (info[:, Z:] / 2).exp() * eps)).abs().max().item() < 1e-5
    losses = torch.empty(3, device="cuda")
    da, dmu, dlv = torch.empty(B, nq, A, device="cuda"), torch.empty(B, Z, device="cuda"), torch.empty(B, Z, device="cuda")
    ck(L().act_loss(p(a_hat), p(actions), T * A, p(pad8), T, p(mu), p(lv), B, nq, A, Z, 10.0, 1.0, p(losses), p(da), p(dmu), p(dlv), st()))
    ar, ir = a_hat.clone().requires_grad_(True), info.clone().requires_grad_(True)
    mur, lvr = ir[:, :Z], ir[:, Z:]
    klds = -0.5 * (1 + lvr - mur.pow(2) - lvr.exp())
    kl = klds.sum(1).mean(0, True)[0]
    l1 = (F.l1_loss(actions[:, :nq], ar, reduction="none") * ~is_pad[:, :nq].unsqueeze(-1)).mean()
    loss = l1 + kl * 10
    loss.backward()
    ref = torch.stack([l1, kl, loss]).detach()
    assert ((losses - ref).abs() / ref.abs()).max().item() < 1e-5
    assert (da - ar.grad).abs().max().item() < 1e-9
    assert (torch.cat([dmu, dlv], 1) - ir.grad).abs().max().item() < 1e-5
    dz = rnd(B, Z, seed=5)
    dinfo = torch.empty(B, 2 * Z, device="cuda")
    ck(L().act_reparam_bwd(p(dz), p(eps), p(lv), p(dmu), p(dlv), p(dinfo), B, Z, st()))
    ir2 = info.clone().requires_grad_(True)
    zz = ir2[:, :Z] + (ir2[:, Z:] / 2).exp() * eps
    (zz * dz).sum().backward()
    assert (dinfo - (ir2.grad + ir.grad)).abs().max().item() < 1e-5


def test_maxpool_fwd_bwd():
    N, H, W, Cc = 2, 48, 64, 64
    x = torch.relu(rnd(N, H, W, Cc, seed=1)).bfloat16()
    y = torch.empty(N, H // 2, W // 2, Cc, device="cuda", dtype=torch.bfloat16)
    idx = torch.empty(N, H // 2, W // 2, Cc, device="cuda", dtype=torch.uint8)
    ck(L().act_maxpool_fwd(p(x), p(y), p(idx), N, H, W, Cc, st()))
    xr = x.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
    ref = F.max_pool2d(xr, 3, 2, 1)
    assert torch.equal(y.float(), ref.permute(0, 2, 3, 1))
    dy = rnd(N, H // 2, W // 2, Cc, seed=2).bfloat16()
    dx = torch.empty_like(x)
    ck(L().act_maxpool_bwd(p(dy), p(idx), p(x), p(dx), N, H, W, Cc, st()))
    ref.backward(dy.float().permute(0, 3, 1, 2))
    refdx = (xr.grad * (xr > 0)).permute(0, 2, 3, 1)
    # ties between equal positive bf16 values may resolve to a different tap than torch; they are rare
    bad = ((dx.float() - refdx).abs() > 2e-2 * (1 + refdx.abs())).float().mean().item()
    assert bad < 2e-3, bad
    # x == NULL: the ReLU mask comes from the arg-max record (tap 9 = window maximum not positive) -- same result
    dx2 = torch.full_like(x, 7.0)
    ck(L().act_maxpool_bwd(p(dy), p(idx), None, p(dx2), N, H, W, Cc, st()))
    assert torch.equal(dx2, dx)


@pytest.mark.parametrize("shape", [(1, 30, 50, 64), (2, 17, 23, 8)])
def test_maxpool_bwd_odd_sizes(shape):
    N, H, W, Cc = shape
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    x = torch.relu(rnd(N, H, W, Cc, seed=5)).bfloat16()
    y = torch.empty(N, Ho, Wo, Cc, device="cuda", dtype=torch.bfloat16)
    idx = torch.empty(N, Ho, Wo, Cc, device="cuda", dtype=torch.uint8)
    ck(L().act_maxpool_fwd(p(x), p(y), p(idx), N, H, W, Cc, st()))
    xr = x.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
    ref = F.max_pool2d(xr, 3, 2, 1)
    assert torch.equal(y.float(), ref.permute(0, 2, 3, 1))
    dy = rnd(N, Ho, Wo, Cc, seed=6).bfloat16()
    for xin in (x, None):
        dx = torch.full_like(x, 7.0)
        ck(L().act_maxpool_bwd(p(dy), p(idx), p(xin) if xin is not None else None, p(dx), N, H, W, Cc, st()))
        xr.grad = None
        ref.backward(dy.float().permute(0, 3, 1, 2), retain_graph=True)
        refdx = (xr.grad * (xr > 0)).permute(0, 2, 3, 1)
        bad = ((dx.float() - refdx).abs() > 2e-2 * (1 + refdx.abs())).float().mean().item()
        assert bad < 4e-3, bad


@pytest.mark.parametrize("shape", [(2, 48, 64, 64), (1, 30, 50, 64), (3, 17, 23, 128)])
def test_maxpool_fwd_inference_form(shape):
    """idx == NULL selects the inference kernel (packed bf16 maxima, two outputs per thread): bit-equal to torch, also
    for odd sizes (odd output width: the second output of the last pair does not exist) and signed inputs."""
    N, H, W, Cc = shape
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    x = rnd(N, H, W, Cc, seed=3).bfloat16()
    y = torch.full((N, Ho, Wo, Cc), 9.0, device="cuda", dtype=torch.bfloat16)
    ck(L().act_maxpool_fwd(p(x), p(y), None, N, H, W, Cc, st()))
    ref = F.max_pool2d(x.float().permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1)
    assert torch.equal(y.float(), ref)


def test_adamw_matches_torch():
    n = 100003
    p0, g = rnd(n + 1, seed=1)[:n + 1], rnd(n + 1, seed=2)
    pt = torch.nn.Parameter(p0[:n].clone())
    opt = torch.optim.AdamW([pt], lr=1e-3, weight_decay=1e-4)
    pp, m, v = p0[:n].clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    sh = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for step in range(1, 4):
        gi = g[:n] * step
        pt.grad = gi.clone()
        opt.step()
        ck(L().act_adamw_step(p(pp), p(gi), p(m), p(v), p(sh), n, 1e-3, 0.9, 0.999, 1e-8, 1e-4, step, 0, st()))
    assert (pp - pt.data).abs().max().item() < 2e-6
    assert torch.equal(sh, pp.bfloat16())
    q = p0[:n].clone()
    ck(L().act_adamw_step(p(q), None, None, None, None, n, 1e-3, 0.9, 0.999, 1e-8, 1e-4, 1, 1, st()))
    assert (q - p0[:n] * (1 - 1e-7)).abs().max().item() < 1e-6


def test_temporal_agg_bit_exact_vs_reference_lines():
    """Golden produced by the literal lines imitate_episodes.py:403-411 (oracle/make_golden.py)."""
    gold = np.load(os.path.join(GOLDEN, "temporal_agg.npz"))
    T, nq, A = [int(x) for x in gold["meta"]]
    chunks = torch.from_numpy(gold["chunks"]).cuda()          # [T, nq, A]
    E = 3                                                      # run the same stream in several envs at once
    ring = torch.zeros(E, nq, nq, A, device="cuda")
    out = torch.empty(E, A, device="cuda", dtype=torch.float64)
    tt = torch.zeros(E, dtype=torch.int32, device="cuda")
    for t in range(T):
        tt.fill_(t)
        c = chunks[t][None].repeat(E, 1, 1).contiguous()
        ck(L().act_temporal_agg_step(p(ring), p(c), p(tt), p(out), E, nq, A, 0.01, st()))
        ref = torch.from_numpy(gold["out"][t]).cuda()
        for e in range(E):
            err = ((out[e] - ref).abs() / (ref.abs() + 1e-300)).max().item()
            assert err < 1e-12, (t, e, err)


def test_temporal_agg_full_size_ring_wraps_bit_exact():
    """nq = 100, T = 400 (BASELINE rollout shape): the [nq, nq, A] ring wraps four times; every step's float64 output
    equals the literal imitate_episodes.py:403-411 lines on the [T, T+nq, A] buffer (golden from oracle/make_golden.py).
    Environments run the same stream at different phases (env e starts e*37 steps late) to exercise per-env t."""
    from oracle import act_oracle as O
    gold = np.load(os.path.join(GOLDEN, "temporal_agg_full.npz"))
    T, nq, A, seed = [int(x) for x in gold["meta"]]
    chunks = O.temporal_agg_stream(T, nq, A, seed)[:, 0].cuda()
    ref = torch.from_numpy(gold["out"]).cuda()
    E, lag = 3, 37
    ring = torch.zeros(E, nq, nq, A, device="cuda")
    out = torch.empty(E, A, device="cuda", dtype=torch.float64)
    tt = torch.zeros(E, dtype=torch.int32, device="cuda")
    worst = 0.0
    for step in range(T + (E - 1) * lag):
        te = [min(max(step - e * lag, 0), T - 1) for e in range(E)]
        live = [0 <= step - e * lag < T for e in range(E)]
        tt.copy_(torch.tensor(te, dtype=torch.int32))
        c = torch.stack([chunks[t] for t in te]).contiguous()
        if not all(live):                     # idle envs: run them on a scratch ring so their real ring is untouched
            ring_s, out_s = ring.clone(), out.clone()
            ck(L().act_temporal_agg_step(p(ring_s), p(c), p(tt), p(out_s), E, nq, A, 0.01, st()))
            for e in range(E):
                if live[e]:
                    ring[e].copy_(ring_s[e]); out[e].copy_(out_s[e])
        else:
            ck(L().act_temporal_agg_step(p(ring), p(c), p(tt), p(out), E, nq, A, 0.01, st()))
        for e in range(E):
            if live[e]:
                r = ref[te[e]]
                worst = max(worst, ((out[e] - r).abs() / (r.abs() + 1e-300)).max().item())
    assert worst < 1e-12, worst


@pytest.mark.parametrize("B,h,Lq,Lk,masked", [(2, 8, 302, 302, False), (2, 8, 102, 102, True), (3, 8, 100, 302, False),
                                              (1, 8, 100, 100, False), (2, 8, 130, 320, True),
                                              # long rows (3 cameras: 902 keys) -> chunked two-pass kernel, several query tiles per item
                                              (2, 8, 902, 902, False), (1, 8, 100, 902, False), (1, 8, 300, 520, True),
                                              (20, 8, 902, 902, False)])
def test_attn_scores_fused_matches_torch(B, h, Lq, Lk, masked):
    """act_attn_scores: softmax(q k^T / 8 + key_padding_mask) and its backward with the score tile kept in TMEM, against
    plain fp32 torch on the same bf16 operands (nn.MultiheadAttention math, transformer.py:194,254-255)."""
    torch.manual_seed(3)
    d, hd = h * 64, 64
    Lkp = (Lk + 7) // 8 * 8
    q = (torch.randn(B, Lq, d, device="cuda") * 0.7).bfloat16()
    k = (torch.randn(B, Lk, d, device="cuda") * 0.7).bfloat16()
    mask = None
    if masked:
        mask = torch.zeros(B, Lk, dtype=torch.uint8, device="cuda")
        mask[:, Lk - 17:] = 1
        mask[0, 5] = 1
    P = torch.full((B, h, Lq, Lkp), 7.0, device="cuda", dtype=torch.bfloat16)
    ck(L().act_attn_scores(0, p(q), d, hd, Lq * d, p(k), d, hd, Lk * d, Lq, Lk, Lkp, h, B, C.c_float(0.125),
                           p(mask) if mask is not None else None, None, p(P), st()))
    qf = q.float().view(B, Lq, h, hd).transpose(1, 2)
    kf = k.float().view(B, Lk, h, hd).transpose(1, 2)
    S = qf @ kf.transpose(2, 3) * 0.125
    if mask is not None:
        S = S.masked_fill(mask.bool()[:, None, None, :], float("-inf"))
    ref = torch.softmax(S, dim=-1)
    torch.cuda.synchronize()
    assert (P[..., :Lk].float() - ref).abs().max().item() <= 4e-3
    assert (P[..., Lk:].float() == 0).all()
    # backward: dS = scale * P o (dP - rowsum(dP o P)), dP = dO V^T, with the kernel's own bf16 P
    dO = torch.randn(B, Lq, d, device="cuda").bfloat16()
    v = torch.randn(B, Lk, d, device="cuda").bfloat16()
    dS = torch.full((B, h, Lq, Lkp), 7.0, device="cuda", dtype=torch.bfloat16)
    ck(L().act_attn_scores(1, p(dO), d, hd, Lq * d, p(v), d, hd, Lk * d, Lq, Lk, Lkp, h, B, C.c_float(0.125), None, p(P), p(dS), st()))
    dP = dO.float().view(B, Lq, h, hd).transpose(1, 2) @ v.float().view(B, Lk, h, hd).transpose(1, 2).transpose(2, 3)
    Pf = P[..., :Lk].float()
    ref_dS = 0.125 * Pf * (dP - (dP * Pf).sum(-1, keepdim=True))
    torch.cuda.synchronize()
    err = (dS[..., :Lk].float() - ref_dS).abs().max().item()
    assert err <= 2e-2 * max(1.0, ref_dS.abs().max().item()), err
    assert (dS[..., Lk:].float() == 0).all()
    # the same through act_attn_scores_bwd with the forward's output O = P V given: rowsum(dP o P) = dO . O, one pass over
    # the keys for rows of more than 320 keys (shorter rows ignore O).  O is rounded to bf16 like the engine's buffer.
    Of = (Pf @ v.float().view(B, Lk, h, hd).transpose(1, 2)).transpose(1, 2).reshape(B, Lq, d)
    O = Of.bfloat16().contiguous()
    dS2 = torch.full((B, h, Lq, Lkp), 7.0, device="cuda", dtype=torch.bfloat16)
    ck(L().act_attn_scores_bwd(p(dO), d, hd, Lq * d, p(v), d, hd, Lk * d, p(O), d, hd, Lq * d, Lq, Lk, Lkp, h, B, C.c_float(0.125),
                               p(P), p(dS2), st()))
    torch.cuda.synchronize()
    err2 = (dS2[..., :Lk].float() - ref_dS).abs().max().item()
    assert err2 <= 2e-2 * max(1.0, ref_dS.abs().max().item()), err2
    assert (dS2[..., Lk:].float() == 0).all()
