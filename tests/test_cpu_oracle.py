"""CPU suite (`-m "not gpu"`): the oracle against the golden vectors written by the UNMODIFIED reference
(oracle/make_golden.py), the literal temporal-ensemble lines, and the reference AdamW semantics the engine mirrors.
Nothing here touches the CUDA path; /root/reference is not needed (the goldens are committed)."""
import os

import numpy as np
import pytest
import torch

from oracle import act_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _load(case):
    gold = np.load(os.path.join(GOLDEN, f"act_{case}.npz"))
    B, H, W, K, seed = [int(x) for x in gold["meta"]]
    cfg = O.make_cfg(num_cameras=K)
    sd = O.init_state_dict(cfg, seed)
    batch = O.make_inputs(cfg, B, H, W, seed)
    return gold, cfg, sd, batch


@pytest.mark.parametrize("case", ["small_k1", "small_k3"])
def test_oracle_reproduces_reference_golden(case):
    """fp32 restatement == reference outputs (same machine class: <= 2e-5 abs on outputs, 1e-5 rel on losses)."""
    gold, cfg, sd, batch = _load(case)
    assert np.allclose(O.state_checksum(sd), gold["state_checksum"], rtol=1e-9)
    o = O.loss_and_grads(sd, cfg, batch)
    for k in ("a_hat", "mu", "logvar"):
        assert float(np.abs(o["out"][k].numpy() - gold[k]).max()) <= 2e-5 * max(1.0, float(np.abs(gold[k]).max())), k
    for k in ("l1", "kl", "loss"):
        assert abs(float(o["out"][k]) - float(gold[k])) <= 1e-5 * abs(float(gold[k])), k
    inf = O.policy_forward(sd, cfg, batch["qpos"], batch["image"])
    assert float(np.abs(inf["a_hat"].numpy() - gold["a_hat_infer"]).max()) <= 2e-5
    # gradients: norms of every parameter the reference reaches, None / exact-zero structure
    names = [str(n) for n in gold["grad_names"]]
    for i, n in enumerate(names):
        g = o["grads"][n]
        if gold["grad_none"][i]:
            assert g is None, n
        elif g is None:                       # dead decoder layers: reference gives exact zeros
            assert gold["grad_norm"][i] == 0.0, n
        else:
            assert abs(float(g.double().norm()) - gold["grad_norm"][i]) <= 1e-3 * gold["grad_norm"][i] + 1e-7, n


def test_dead_decoder_layers_and_is_pad_head():
    gold, cfg, sd, batch = _load("small_k1")
    names = [str(n) for n in gold["grad_names"]]
    dead = [n for i, n in enumerate(names) if n.startswith("transformer.decoder.layers.") and
            not n.startswith("transformer.decoder.layers.0.")]
    assert dead and all(gold["grad_norm"][names.index(n)] == 0.0 and not gold["grad_none"][names.index(n)] for n in dead)
    assert all(gold["grad_none"][names.index(n)] for n in names if n.startswith("is_pad_head."))


def test_temporal_agg_oracle_matches_golden_bit_exact():
    g = np.load(os.path.join(GOLDEN, "temporal_agg.npz"))
    T, nq, A = [int(x) for x in g["meta"]]
    agg = O.TemporalAggOracle(T, nq, A)
    chunks = torch.from_numpy(g["chunks"])
    for t in range(T):
        out = agg.step(t, chunks[t][None])
        assert out.dtype == torch.float64 and out.shape == (1, A)
        assert np.array_equal(out.numpy()[0], g["out"][t]), t


def test_temporal_agg_full_size_golden_bit_exact():
    """BASELINE shape (nq = 100, T = 400: a ring of nq chunks wraps 4x): the literal lines on the regenerated stream
    reproduce the committed float64 outputs bit for bit."""
    g = np.load(os.path.join(GOLDEN, "temporal_agg_full.npz"))
    T, nq, A, seed = [int(x) for x in g["meta"]]
    chunks = O.temporal_agg_stream(T, nq, A, seed)
    assert float(chunks.double().sum()) == g["chunk_checksum"][0] and float((chunks == 0).sum()) == g["chunk_checksum"][1]
    agg = O.TemporalAggOracle(T, nq, A)
    for t in range(T):
        out = agg.step(t, chunks[t])
        assert np.array_equal(out.numpy()[0], g["out"][t]), t


def test_temporal_agg_weights_oldest_first_and_zero_filter():
    """imitate_episodes.py:405-409: rows with any exact zero are dropped; index 0 (oldest) has the largest weight."""
    nq, A = 4, 3
    agg = O.TemporalAggOracle(6, nq, A)
    c0 = torch.ones((1, nq, A))
    c1 = 2 * torch.ones((1, nq, A))
    c1[0, 0, 1] = 0.0                       # the entry chunk 1 contributes to step 1 contains a zero -> dropped
    agg.step(0, c0)
    out = agg.step(1, c1)
    assert torch.equal(out, torch.ones((1, A), dtype=torch.float64))
    c2 = 3 * torch.ones((1, nq, A))
    out = agg.step(2, c2)                   # rows: chunk0 (1), chunk1 (2), chunk2 (3)
    w = np.exp(-0.01 * np.arange(3)); w /= w.sum()
    assert np.allclose(out.numpy()[0], (w * np.array([1.0, 2.0, 3.0])).sum())
    assert w[0] > w[1] > w[2]


def test_adamw_restatement_matches_torch_and_zero_grad_decays():
    torch.manual_seed(0)
    p0, g0 = torch.randn(1000), torch.randn(1000)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=1e-5, weight_decay=1e-4)
    p, m, v = p0.clone(), torch.zeros(1000), torch.zeros(1000)
    for step in (1, 2, 3):
        ref.grad = g0 * step
        opt.step()
        p, m, v = O.adamw_step(p, g0 * step, m, v, step, 1e-5)
        assert torch.allclose(p, ref.detach(), atol=1e-7, rtol=1e-6)
    # a parameter reached with exact zeros (decoder layers 1..) is still weight-decayed (detr/main.py:109)
    # (at the reference's lr 1e-5 * wd 1e-4 the factor 1 - 1e-9 rounds to 1 in fp32, so use lr 1e-2 to see it)
    z = torch.nn.Parameter(p0.clone()); z.grad = torch.zeros(1000)
    torch.optim.AdamW([z], lr=1e-2, weight_decay=1e-4).step()
    pz, _, _ = O.adamw_step(p0.clone(), torch.zeros(1000), torch.zeros(1000), torch.zeros(1000), 1, 1e-2)
    assert torch.allclose(pz, z.detach(), atol=0, rtol=1e-6) and not torch.equal(pz, p0)


def test_sinusoid_table_and_sine_embedding_shapes():
    t = O.get_sinusoid_encoding_table(102, 512)
    assert t.shape == (1, 102, 512) and t.dtype == torch.float32
    assert float(t[0, 0, 0]) == 0.0 and float(t[0, 0, 1]) == 1.0
    pos = O.position_embedding_sine(15, 20, 256)
    assert pos.shape == (1, 512, 15, 20)
    # y embedding constant along x, x embedding constant along y (position_encoding.py:36-52)
    assert torch.equal(pos[0, :256, :, 0], pos[0, :256, :, 5]) and torch.equal(pos[0, 256:, 0, :], pos[0, 256:, 7, :])


def test_bf16_operand_error_budget_of_the_backbone():
    """Where the engine's a_hat error budget goes (GPU-measured on full_k1_b8: relative error 0.61 % at the layer4
    features, 0.59 % after the encoder, 0.39 % after the decoder -- the transformer adds nothing net).  Emulating the
    engine's roundings in the fp32 oracle reproduces that figure and splits it: bf16 weights alone 0.35 %, bf16 stored
    activations alone 0.49 %, an fp32 identity trunk would only bring the total from 0.60 % to 0.50 %.  So the margin to
    the 1e-2 tolerance is what "bf16 operands, fp32 accumulate" (BASELINE.json north_star) costs on this network, not a
    defect of one kernel."""
    import torch.nn.functional as F
    cfg = O.make_cfg()
    sd = O.init_state_dict(cfg, 3)
    x0 = O.normalize_image(O.make_inputs(cfg, 1, 480, 640, 3)["image"])[:, 0]
    bf = lambda t: t.bfloat16().float()

    def run(rw, ra, ridt, rimg):
        Wt = (lambda n: bf(sd[n])) if rw else (lambda n: sd[n])
        A = bf if ra else (lambda t: t)
        p = "backbones.0.0.body."
        x = F.conv2d(bf(x0) if rimg else x0, Wt(p + "conv1.weight"), stride=2, padding=3)
        x = F.max_pool2d(A(F.relu(O.frozen_bn(sd, p + "bn1.", x))), 3, 2, 1)
        xi = x
        for name, cin, cout, stride in O._RESNET_STAGES:
            for blk in range(2):
                q, s = f"{p}{name}.{blk}.", (stride if blk == 0 else 1)
                out = A(F.relu(O.frozen_bn(sd, q + "bn1.", F.conv2d(x, Wt(q + "conv1.weight"), stride=s, padding=1))))
                out = O.frozen_bn(sd, q + "bn2.", F.conv2d(out, Wt(q + "conv2.weight"), stride=1, padding=1))
                idt = xi
                if (q + "downsample.0.weight") in sd:
                    idt = O.frozen_bn(sd, q + "downsample.1.", F.conv2d(x, Wt(q + "downsample.0.weight"), stride=s))
                    idt = bf(idt) if ridt else idt
                y = F.relu(out + idt)
                xi, x = (bf(y) if ridt else y), A(y)
        return xi

    ref = run(False, False, False, False)
    rel = lambda y: float((y - ref).norm() / ref.norm())
    all_bf16, fp32_trunk = rel(run(True, True, True, True)), rel(run(True, True, False, True))
    w_only, a_only = rel(run(True, False, False, False)), rel(run(False, True, True, False))
    assert 4.5e-3 < all_bf16 < 7.5e-3, all_bf16              # the engine's measured 6.1e-3
    assert 2.5e-3 < w_only < 4.5e-3 and 3.5e-3 < a_only < 6e-3, (w_only, a_only)
    assert 0.75 < fp32_trunk / all_bf16 < 0.95, (fp32_trunk, all_bf16)
