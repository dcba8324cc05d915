"""Train-mode dropout (transformer.py:194,197,202-203,219-222,254-266,284-293; p = 0.1 at detr/main.py:45): the engine
draws counter-based Philox masks inside its kernels; with the SAME masks injected into the oracle's dropout sites the
training call, its loss and its gradients must agree within the bf16 budget of the deterministic path."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _engine(case="small_k1"):
    from oracle import act_oracle as O
    from act_plus_plus_b200.engine import ActEngine
    gold = np.load(os.path.join(GOLDEN, f"act_{case}.npz"))
    B, H, W, K, seed = [int(x) for x in gold["meta"]]
    cfg = O.make_cfg(num_cameras=K)
    sd = O.init_state_dict(cfg, seed)
    batch = O.make_inputs(cfg, B, H, W, seed)
    eng = ActEngine(batch_max=B, num_cameras=K, img_h=H, img_w=W, train=True)
    eng.load_state_dict(sd)
    return O, gold, cfg, sd, batch, eng


def test_mask_stream_statistics_and_determinism():
    from act_plus_plus_b200.engine import ActEngine
    eng = ActEngine(batch_max=1, num_cameras=1, img_h=64, img_w=64, train=True)
    eng.set_dropout(0.1, seed=1234)
    n = 1 << 20
    a = eng.dropout_mask(0x101, n, step=3)
    assert torch.equal(a, eng.dropout_mask(0x101, n, step=3))
    keep = a.float().mean().item()
    assert abs(keep - 0.9) < 2e-3, keep                                  # sigma = 3e-4
    for other in (eng.dropout_mask(0x102, n, step=3), eng.dropout_mask(0x101, n, step=4)):
        agree = (a == other).float().mean().item()
        assert abs(agree - 0.82) < 5e-3, agree                           # independent streams: 0.9^2 + 0.1^2
    eng.set_dropout(0.1, seed=1235)
    assert abs((a == eng.dropout_mask(0x101, n, step=3)).float().mean().item() - 0.82) < 5e-3
    # lag-1 autocorrelation of the keep bits (eight 16-bit lanes per Philox call must be independent)
    x = a.float() - keep
    assert abs((x[1:] * x[:-1]).mean().item()) < 1e-3


def test_training_call_with_dropout_matches_oracle_with_injected_masks():
    O, gold, cfg, sd, batch, eng = _engine("small_k1")
    B = batch["qpos"].shape[0]
    d, ff, h, nq = cfg["hidden_dim"], cfg["dim_feedforward"], cfg["nheads"], cfg["num_queries"]
    p = 0.1
    eng.set_dropout(p, seed=77)
    cb = {k: v.cuda() for k, v in batch.items()}
    eng.zero_grad()
    out = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])
    eng.backward()
    torch.cuda.synchronize()
    _, _, step = eng.dropout_state()
    assert step == 1
    K = cfg["num_cameras"]
    L = 2 + (batch["image"].shape[-2] // 32) * (batch["image"].shape[-1] // 32) * K
    Lc = nq + 2
    r8 = lambda n: (n + 7) // 8 * 8
    used = set()

    def hook(site, x):
        if site.startswith("transformer.encoder.layers."):
            base, Lx = eng.SITE_ENC + 8 * int(site.split(".")[3]), L
        elif site.startswith("encoder.layers."):
            base, Lx = eng.SITE_CVAE + 8 * int(site.split(".")[2]), Lc
        elif site.startswith("transformer.decoder.layers.0."):
            base, Lx = eng.SITE_DEC, nq
        else:
            raise AssertionError(site)
        used.add(site)
        dec = base == eng.SITE_DEC
        leaf = site.split(".")[-1] if not site.endswith(".attn") else ".".join(site.split(".")[-2:])
        if leaf in ("self_attn.attn", "multihead_attn.attn"):
            k = 4 if leaf.startswith("multihead") else 0
            Lq, Lk = x.shape[1], x.shape[2]
            m = eng.dropout_mask(base + k, B * h * Lq * r8(Lk)).view(B, h, Lq, r8(Lk))[..., :Lk]
            return m.reshape(B * h, Lq, Lk).cpu()
        k, width = {"drop1": (1, d), "ffn": (2, ff), "drop2": (5 if dec else 3, d), "drop3": (3, d)}[leaf]
        m = eng.dropout_mask(base + k, B * Lx * width).view(B, Lx, width)
        return m.permute(1, 0, 2).cpu()                                  # oracle tensors are sequence-first [L, B, *]

    O.DROP_HOOK = hook
    try:
        ref = O.loss_and_grads(sd, cfg, batch, dropout_p=p)
    finally:
        O.DROP_HOOK = None
    assert len(used) == 4 * cfg["enc_layers"] * 2 + 6, sorted(used)       # every dropout site of the live path was hit
    err = float((out["a_hat"].cpu() - ref["out"]["a_hat"]).abs().max())
    ls = out["losses"].cpu()
    lerr = [abs(float(ls[i]) - float(ref["out"][k])) / abs(float(ref["out"][k])) for i, k in enumerate(("l1", "kl", "loss"))]
    assert err <= 1e-2 and max(lerr) <= 1e-3, (err, lerr)
    # and it is NOT the deterministic result (the masks really act)
    assert float((out["a_hat"].cpu() - torch.from_numpy(gold["a_hat"])).abs().max()) > 5e-2
    worst = {}
    for n, g in ref["grads"].items():
        if g is None or float(g.norm()) < 1e-6 * max(float(v.norm()) for v in ref["grads"].values() if v is not None):
            continue
        ours = eng.grad_view(n).cpu()
        worst[n] = float((ours - g).norm() / g.norm())
    top = sorted(worst.items(), key=lambda kv: -kv[1])[:5]
    assert np.median(list(worst.values())) < 3e-2 and top[0][1] < 0.3, top


def test_eval_mode_and_inference_are_deterministic():
    O, gold, cfg, sd, batch, eng = _engine("small_k1")
    cb = {k: v.cuda() for k, v in batch.items()}
    eng.set_dropout(0.1, seed=5)
    a = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])["a_hat"].clone()
    b = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])["a_hat"].clone()
    assert float((a - b).abs().max()) > 1e-3                              # new step, new masks
    i1 = eng.forward(cb["qpos"], cb["image"])["a_hat"].clone()            # inference call: never dropped
    assert float((i1.cpu() - torch.from_numpy(gold["a_hat_infer"])).abs().max()) <= 1e-2
    eng.set_dropout(0.0)
    c = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])["a_hat"]
    assert float((c.cpu() - torch.from_numpy(gold["a_hat"])).abs().max()) <= 1e-2
