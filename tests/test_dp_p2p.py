"""Data-parallel optimiser step over peer memory (csrc/dp_p2p.cu, act_engine_backward_and_step_p2p) on ONE GPU: the
engine is rank r of a world of 2 whose other member is a "ghost" -- plain device buffers holding that rank's gradients
and receiving the parameters this rank broadcasts, with its barrier flags pre-set as arrived.  Checks the slice
ownership, the averaged-gradient AdamW arithmetic (torch.optim.AdamW semantics, detr/main.py:102-110) and the stores
into the peer's parameter / bf16-shadow buffers.  The 2-process run over NVLink is tools/dp_check.py (ACT_DP_P2P=1)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _engine_and_batch(seed=5, B=2, H=96, W=128):
    from act_plus_plus_b200.engine import ActEngine
    from act_plus_plus_b200.policy import _init_like_reference
    eng = ActEngine(batch_max=B, img_h=H, img_w=W, train=True)
    _init_like_reference(eng, seed=seed)
    eng.set_lr(1e-3, 1e-3)
    g = torch.Generator(device="cpu").manual_seed(11)
    b = dict(image=torch.rand((B, 1, 3, H, W), generator=g), qpos=torch.randn((B, 14), generator=g),
             actions=torch.randn((B, 100, 16), generator=g), is_pad=(torch.arange(100)[None].repeat(B, 1) >= 70),
             eps=torch.randn((B, 32), generator=g))
    return eng, {k: v.cuda() for k, v in b.items()}


@pytest.mark.parametrize("rank", [0, 1])
@pytest.mark.parametrize("concurrent", [True, False])
def test_p2p_step_with_ghost_peer(rank, concurrent):
    from act_plus_plus_b200.dp import owner_slices
    from act_plus_plus_b200.engine import R_BACKBONE, R_CVAE, R_DEAD, R_MAIN
    eng, b = _engine_and_batch()
    eng.set_concurrent(concurrent)
    n = eng.flat_params.numel()
    gen = torch.Generator(device="cuda").manual_seed(3 + rank)
    ghost_grads = torch.randn(n, device="cuda", generator=gen) * 1e-3
    ghost_params = torch.full((n,), 7.0, device="cuda")
    ghost_shadow = torch.full((n,), 7.0, device="cuda", dtype=torch.bfloat16)
    ghost_signal = torch.zeros(64 * 8, device="cuda", dtype=torch.int32)
    mine = eng.dp_local_ptrs()
    ghost = (ghost_grads.data_ptr(), ghost_params.data_ptr(), ghost_shadow.data_ptr(), ghost_signal.data_ptr())
    table = [[mine[i], ghost[i]] if rank == 0 else [ghost[i], mine[i]] for i in range(4)]
    eng.dp_set_peers(rank, 2, table)
    from act_plus_plus_b200 import _lib
    _lib.check(eng.lib.act_engine_dp_preset_signal(eng._h, 1 - rank, 1 << 30), "preset")
    p0 = eng.flat_params.clone()
    eng.zero_grad()
    eng.forward(b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
    eng.backward_and_step_p2p(1.0)
    torch.cuda.synchronize()
    assert eng.dp_status() == 0
    g_local = eng.flat_grads.clone()                      # the local buffer keeps this rank's un-averaged gradients
    p1 = eng.flat_params
    ranges = eng.dp_slices()
    covered = sum(c for _, c in ranges)
    want = sum(eng.region(r)[1] for r in (R_MAIN, R_CVAE, R_BACKBONE))
    assert covered == want, (covered, want, ranges)
    # every barrier of the step published this rank's arrival in the ghost's flag block (epoch 1)
    sig = ghost_signal.view(64, 8)[:, rank].cpu().numpy()
    assert (sig[:3] == 1).all() and sig.max() == 1 and int((sig == 1).sum()) >= 3
    lr, wd, b1, b2, eps = 1e-3, 1e-4, 0.9, 0.999, 1e-8
    own = torch.zeros(n, dtype=torch.bool, device="cuda")
    for off, cnt in ranges:
        for r, lo, hi in owner_slices(cnt, 2):
            if r == rank:
                own[off + lo:off + hi] = True
    exchanged = torch.zeros(n, dtype=torch.bool, device="cuda")
    for off, cnt in ranges:
        exchanged[off:off + cnt] = True
    # expected AdamW (step 1, zero moments) on the averaged gradient
    g = 0.5 * (g_local + ghost_grads)
    m = (1 - b1) * g
    v = (1 - b2) * g * g
    exp = p0 * (1 - lr * wd) - (lr / (1 - b1)) * (m / (v.sqrt() / np.sqrt(1 - b2) + eps))
    d = (p1 - exp).abs()
    assert float(d[own].max()) < 2e-6, float(d[own].max())
    # slices owned by the ghost are left for it to write; the dead decoder layers are decayed locally
    theirs = exchanged & ~own
    assert torch.equal(p1[theirs], p0[theirs])
    off_d, n_d = eng.region(R_DEAD)
    assert torch.allclose(p1[off_d:off_d + n_d], p0[off_d:off_d + n_d] * (1 - lr * wd), rtol=0, atol=1e-7)
    # broadcast: the ghost's parameter buffer received exactly this rank's slices, bit for bit, and nothing else
    assert torch.equal(ghost_params[own], p1[own])
    assert bool((ghost_params[~own] == 7.0).all())
    # bf16 shadow for the transformer / CVAE regions only (conv operands are re-packed from the fp32 masters)
    sh = torch.zeros(n, dtype=torch.bool, device="cuda")
    for r in (R_MAIN, R_CVAE):
        o, c = eng.region(r)
        sh[o:o + c] = True
    assert torch.equal(ghost_shadow[own & sh], p1[own & sh].to(torch.bfloat16))
    assert bool((ghost_shadow[~(own & sh)] == 7.0).all())
    # moments: current only on the owned slices
    assert torch.allclose(eng.exp_avg[own], m[own], rtol=1e-5, atol=1e-10)
    assert bool((eng.exp_avg[theirs] == 0).all())
    # a second step runs (epochs advance, ghost flags still satisfied) and keeps the loss finite
    eng.zero_grad()
    out = eng.forward(b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
    eng.backward_and_step_p2p(1.0)
    torch.cuda.synchronize()
    assert eng.dp_status() == 0 and bool(torch.isfinite(out["losses"]).all())


def test_p2p_step_requires_connected_peers():
    eng, b = _engine_and_batch()
    eng.zero_grad()
    eng.forward(b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
    with pytest.raises(RuntimeError):
        eng.backward_and_step_p2p(1.0)
