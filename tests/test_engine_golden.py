"""End-to-end parity of the CUDA engine against golden vectors produced by the UNMODIFIED reference
(oracle/make_golden.py): a_hat / mu / logvar max-abs <= 1e-2, losses relative <= 1e-3 (BASELINE.json tolerances)."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
# full_k3_b2: 3 cameras, 902 encoder tokens (BASELINE configs[3] per-sample shape); small_noenc_k1: encoder=None variant
# (detr_vae.py:109-112); small_s7a6_k4: the fork's single-arm shape -- state_dim 7, action_dim 6, 4 cameras
CASES = ["small_k1", "small_k3", "mid_k1", "full_k1_b8", "full_k3_b2", "small_noenc_k1", "small_s7a6_k4"]
TOL_ABS, TOL_LOSS = 1e-2, 1e-3


def _setup(case):
    from oracle import act_oracle as O
    from act_plus_plus_b200.engine import ActEngine
    gold = np.load(os.path.join(GOLDEN, f"act_{case}.npz"))
    B, H, W, K, seed = [int(x) for x in gold["meta"]]
    over = eval(str(gold["cfg_over"])) if "cfg_over" in gold.files else {}      # a dict literal written by make_golden
    cfg = O.make_cfg(num_cameras=K, **over)
    sd = O.init_state_dict(cfg, seed)
    batch = O.make_inputs(cfg, B, H, W, seed)
    assert np.allclose(O.state_checksum(sd), gold["state_checksum"], rtol=1e-9), "seeded weights differ from the golden run"
    eng = ActEngine(batch_max=B, num_cameras=K, img_h=H, img_w=W, train=True, state_dim=cfg["state_dim"],
                    action_dim=cfg["action_dim"], no_encoder=bool(cfg["no_encoder"]))
    eng.load_state_dict(sd)
    cb = {k: v.cuda() for k, v in batch.items()}
    return gold, cfg, sd, cb, eng


def _report(case, what, vals):
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/parity_report.jsonl", "a") as f:
        f.write(json.dumps(dict(case=case, what=what, **vals)) + "\n")


@pytest.mark.parametrize("case", CASES)
def test_forward_train_call_matches_reference(case):
    gold, cfg, sd, cb, eng = _setup(case)
    out = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])
    torch.cuda.synchronize()
    errs = {}
    for k in ("a_hat", "mu", "logvar"):
        errs[k] = float((out[k].cpu() - torch.from_numpy(gold[k])).abs().max())
    ls = out["losses"].cpu().numpy()
    for i, k in enumerate(("l1", "kl", "loss")):
        errs[k] = float(abs(ls[i] - gold[k]) / abs(gold[k])) if float(gold[k]) != 0.0 else float(abs(ls[i]))   # kl = 0: exact
    _report(case, "train_call", errs)
    for k in ("a_hat", "mu", "logvar"):
        assert errs[k] <= TOL_ABS, (case, k, errs)
    for k in ("l1", "kl", "loss"):
        assert errs[k] <= TOL_LOSS, (case, k, errs)


@pytest.mark.parametrize("case", CASES)
def test_forward_inference_matches_reference(case):
    gold, cfg, sd, cb, eng = _setup(case)
    out = eng.forward(cb["qpos"], cb["image"])
    torch.cuda.synchronize()
    err = float((out["a_hat"].cpu() - torch.from_numpy(gold["a_hat_infer"])).abs().max())
    _report(case, "inference", dict(a_hat=err))
    assert err <= TOL_ABS, (case, err)


def test_three_camera_forward_through_the_chunked_attention_kernel(monkeypatch):
    """Rows of 902 keys take the chunked fused kernel by default; ACT_ATTN_CHUNKED=0 routes them to tc_gemm + softmax
    (the round-1 path, kept as an A/B switch): same golden, same tolerances, and the two paths agree with each other far
    inside the tolerance, forward and backward."""
    gold, cfg, sd, cb, eng = _setup("full_k3_b2")
    monkeypatch.setenv("ACT_ATTN_CHUNKED", "0")
    ref = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])
    ref = {k: ref[k].clone() for k in ("a_hat", "mu", "logvar", "losses")}
    eng.zero_grad()
    eng.backward(0)
    g_ref = eng.flat_grads.clone()
    monkeypatch.delenv("ACT_ATTN_CHUNKED")
    out = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])
    torch.cuda.synchronize()
    for k in ("a_hat", "mu", "logvar"):
        err = float((out[k].cpu() - torch.from_numpy(gold[k])).abs().max())
        assert err <= TOL_ABS, (k, err)
    between = float((out["a_hat"] - ref["a_hat"]).abs().max())
    assert 0.0 < between <= 4e-3, between            # different kernels (not bit-equal), same arithmetic
    # backward through the chunked dS kernel -- one pass over the keys, rowsum(P o dP) taken as dO . O -- against the
    # unfused softmax backward: every region's gradient agrees to a fraction of the bf16 noise of the step
    eng.zero_grad()
    eng.backward(0)
    torch.cuda.synchronize()
    g = eng.flat_grads
    assert torch.isfinite(g).all()
    rel = {}
    for r in range(3):
        off, n = eng.region(r)
        if n > 0:
            rel[r] = float((g[off:off + n] - g_ref[off:off + n]).norm() / g_ref[off:off + n].norm().clamp_min(1e-20))
    _report("full_k3_b2", "chunked_vs_unfused", dict(a_hat=between, **{f"grad_rel_region{r}": v for r, v in rel.items()}))
    assert max(rel.values()) <= 2e-2, rel


def test_inference_batch16_matches_reference_and_is_batch_invariant():
    """Inference golden at batch 16 on the full image (per-GPU batch of BASELINE configs[2] at N = 4; the rollout leg's
    many-environment regime), and: every sample's chunk is independent of what else is in the batch -- 128 copies of one
    observation give 128 identical chunks equal to the batch-1 result (what bench.py's 128-env step relies on)."""
    from oracle import act_oracle as O
    from act_plus_plus_b200.engine import ActEngine
    gold = np.load(os.path.join(GOLDEN, "act_full_k1_b16_infer.npz"))
    B, H, W, K, seed = [int(x) for x in gold["meta"]]
    cfg = O.make_cfg(num_cameras=K)
    sd = O.init_state_dict(cfg, seed)
    assert np.allclose(O.state_checksum(sd), gold["state_checksum"], rtol=1e-9)
    batch = O.make_inputs(cfg, B, H, W, seed)
    eng = ActEngine(batch_max=B, num_cameras=K, img_h=H, img_w=W, train=False)
    eng.load_state_dict(sd)
    qpos, image = batch["qpos"].cuda(), batch["image"].cuda()
    out = eng.forward(qpos, image)["a_hat"]
    torch.cuda.synchronize()
    err = float((out.cpu() - torch.from_numpy(gold["a_hat_infer"])).abs().max())
    _report("full_k1_b16_infer", "inference", dict(a_hat=err))
    assert err <= TOL_ABS, err
    big = ActEngine.sharing(eng, 128, H, W, False)
    rep = big.forward(qpos[3:4].repeat(128, 1), image[3:4].repeat(128, 1, 1, 1, 1))["a_hat"]
    torch.cuda.synchronize()
    assert float((rep - rep[:1]).abs().max()) == 0.0, "a sample's chunk depends on its position in the batch"
    one = ActEngine.sharing(eng, 1, H, W, False).forward(qpos[3:4], image[3:4])["a_hat"]
    d1 = float((rep[0] - one[0]).abs().max())
    _report("full_k1_b16_infer", "batch128_vs_batch1", dict(a_hat=d1))
    # different batch sizes pick different tile shapes (same bf16 operands, other fp32 summation order): tiny, not zero
    assert d1 <= 2e-3, d1
    assert float((rep[0].cpu() - torch.from_numpy(gold["a_hat_infer"][3])).abs().max()) <= TOL_ABS


@pytest.mark.parametrize("case", ["full_k1_b8", "full_k3_b2"])
def test_stage_errors_report(case):
    """Where the bf16 engine spends its a_hat budget: relative error (on strided samples) of the backbone features, the
    encoder input / output and the decoder output against the oracle's fp32 activations stored with the golden."""
    gold, cfg, sd, cb, eng = _setup(case)
    eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])
    torch.cuda.synchronize()
    rows = {}
    for k in [f"feat{c}" for c in range(cfg["num_cameras"])] + ["src", "memory", "hs"]:
        ref = torch.from_numpy(gold[f"stage_{k}"]).double()
        numel, stride = int(gold[f"stage_{k}_meta"][0]), int(gold[f"stage_{k}_meta"][1])
        ours = eng.debug_buffer(k).double().cpu()
        assert ours.numel() == numel, (k, ours.numel(), numel)
        ours = ours[::stride][:ref.numel()]
        rel = float((ours - ref).norm() / (ref.norm() + 1e-30))
        rows[k] = dict(rel=rel, max_abs=float((ours - ref).abs().max()), ref_max=float(ref.abs().max()))
        assert rel < 5e-2, (case, k, rows[k])
    _report(case, "stages", {k: v["rel"] for k, v in rows.items()})
    _report(case, "stages_max_abs", {k: v["max_abs"] for k, v in rows.items()})


@pytest.mark.parametrize("group", ["1", "2"])
def test_stem_group_fallbacks_match_reference(group, monkeypatch):
    """ACT_STEM_GROUP=1|2: the one- and two-pixel-per-GEMM-row stem (the fallbacks of the default four-pixel form for
    widths whose conv1 output is not a multiple of 4) against the same goldens."""
    monkeypatch.setenv("ACT_STEM_GROUP", group)
    gold, cfg, sd, cb, eng = _setup("mid_k1")
    out = eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])
    torch.cuda.synchronize()
    err = float((out["a_hat"].cpu() - torch.from_numpy(gold["a_hat"])).abs().max())
    lerr = float(abs(out["losses"].cpu().numpy()[2] - gold["loss"]) / abs(gold["loss"]))
    _report("mid_k1", f"stem_group_{group}", dict(a_hat=err, loss=lerr))
    assert err <= TOL_ABS and lerr <= TOL_LOSS, (group, err, lerr)


def _sample(t, n=256):
    f = t.detach().flatten()
    stride = max(1, f.numel() // n)
    return f[::stride][:n].double().cpu().numpy()


@pytest.mark.parametrize("case", CASES)
def test_backward_matches_reference_autograd(case):
    """loss.backward() parity: per-parameter gradient norms and strided samples vs the reference's autograd (fp32 CPU).
    bf16 operands -> tolerance is on relative error, judged per tensor against that tensor's own norm."""
    gold, cfg, sd, cb, eng = _setup(case)
    eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], cb["eps"])
    eng.zero_grad()
    eng.backward()
    torch.cuda.synchronize()
    names = [str(n) for n in gold["grad_names"]]
    max_norm = float(gold["grad_norm"].max())
    rows = []
    for i, n in enumerate(names):
        if gold["grad_none"][i]:
            if not n.startswith("is_pad_head."):      # encoder=None variant: the unused CVAE projections take no gradient
                assert float(eng.grad_view(n).abs().max()) == 0.0, n
            continue
        ref_norm = float(gold["grad_norm"][i])
        g = eng.grad_view(n)
        ours_norm = float(g.double().norm())
        if ref_norm == 0.0:
            assert ours_norm == 0.0, (n, ours_norm)       # dead decoder layers: exact zeros
            continue
        if ref_norm < 1e-6 * max_norm:
            # analytically zero (decoder layer-0 self-attention q/k see tgt = 0, all values equal): rounding noise only
            assert ours_norm < 1e-4 * max_norm, (n, ref_norm, ours_norm)
            continue
        rs, os_ = gold["grad_sample"][i][:min(256, g.numel())], _sample(g)
        rs = rs[:len(os_)]
        rel = float(np.linalg.norm(os_ - rs) / (np.linalg.norm(rs) + 1e-30))
        rows.append((n, ref_norm, ours_norm, abs(ours_norm - ref_norm) / ref_norm, rel))
    rows.sort(key=lambda r: -r[4])
    with open(f"gpurun_out/grad_report_{case}.txt", "w") as f:
        for r in rows:
            f.write("%-70s ref=%.4e ours=%.4e norm_rel=%.3e sample_rel=%.3e\n" % r)
    norm_rel = np.array([r[3] for r in rows])
    samp_rel = np.array([r[4] for r in rows])
    _report(case, "backward", dict(norm_rel_max=float(norm_rel.max()), norm_rel_med=float(np.median(norm_rel)),
                                   sample_rel_max=float(samp_rel.max()), sample_rel_med=float(np.median(samp_rel))))
    # bf16 operands: per-tensor norms within 6 %, medians within 2-3 %.  Element samples of the deepest convs are noisier
    # on the tiny fixtures (3x4 feature maps: a handful of pixels per weight, ReLU masks flip on near-zero bf16 values).
    # (the no_encoder fixture has fewer, noisier gradients: its KL term is absent, so the tiny fixture's median is larger)
    assert np.median(samp_rel) < (3e-2 if not case.startswith("small") else 4e-2) and np.median(norm_rel) < 2e-2, \
        (case, "median", float(np.median(samp_rel)))
    assert norm_rel.max() < 0.06, (case, max(rows, key=lambda r: r[3]))
    # worst single tensor: 0.1 at batch 8 on the full image; the batch-2 three-camera case averages 4x fewer samples per
    # conv weight (measured 0.20 on one layer3 conv), the 3x4-feature-map fixtures are noisier still
    limit = 0.1 if case == "full_k1_b8" else 0.25 if case.startswith("full") else 0.45
    assert samp_rel.max() < limit, (case, rows[0])


def test_overlapped_step_equals_serial_step():
    """train_step (async zero_grad, CVAE || backbone branches, weight gradients and per-region AdamW on trailing
    streams) must give the same parameters as the plain call sequence zero_grad -> forward -> backward(1,2,3) -> adamw
    on one stream.  Differences come only from fp32 atomic accumulation order."""
    from act_plus_plus_b200.train import train_step
    _, _, sd, cb, eng_a = _setup("small_k1")
    _, _, _, _, eng_b = _setup("small_k1")
    for eng in (eng_a, eng_b):
        eng.set_lr(1e-3, 1e-3)          # large enough that a wrong / missing update is far outside the tolerance
    eps = torch.randn((cb["qpos"].shape[0], 32), device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
    for it in range(3):
        la = train_step(eng_a, cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], eps).clone()
        eng_b.zero_grad()
        lb = eng_b.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], eps)["losses"].clone()
        for stage in (1, 2, 3):
            eng_b.backward(stage, 1.0)
        eng_b.adamw_step()
        if it == 0:                     # same parameters, deterministic forward: the two losses agree tightly
            assert torch.allclose(la, lb, rtol=1e-5, atol=1e-6), (la, lb)
    torch.cuda.synchronize()
    # after two updates at lr 1e-3 the fp32 atomic-order noise of the backward (either path, run to run) has been
    # amplified by Adam's normalisation: measured spread of the KL term is ~0.3 %
    assert torch.allclose(la, lb, rtol=1.5e-2, atol=1e-5), (la, lb)
    from act_plus_plus_b200.engine import R_BACKBONE, R_CVAE, R_MAIN
    pa, pb, p0 = eng_a.flat_params, eng_b.flat_params, _setup("small_k1")[4].flat_params
    for name, r in (("main", R_MAIN), ("cvae", R_CVAE), ("backbone", R_BACKBONE)):
        off, n = eng_b.region(r)
        da = (pa[off:off + n] - pb[off:off + n]).abs()
        moved = (pb[off:off + n] - p0[off:off + n]).abs().mean().item()
        # Adam normalises every update to ~lr, so a skipped / doubled / stale-gradient region update moves the mean by
        # ~1e-3 per step; elements whose true gradient is zero (e.g. key biases) may flip sign on rounding noise.
        assert moved > 5e-4, f"{name}: parameters did not move ({moved})"
        assert da.mean().item() <= 0.02 * moved, f"{name}: mean |overlapped - serial| = {da.mean().item()} vs moved {moved}"
        # (split-K / atomic fp32 reductions make the two runs differ by bf16 roundings of activations, which moves
        # Adam's normalised update of small-gradient elements; a wrong region update diverges everywhere)
        frac = (da > 0.5 * moved).float().mean().item()
        assert frac < 3e-2, f"{name}: {frac:.3e} of the elements diverge"


def test_grads_ready_hook_ranges_and_identity_step():
    """act_engine_backward_and_step_dp with an averaging callback that does nothing (world size 1) must (a) announce
    ranges that tile the main, CVAE-encoder and backbone regions exactly once, each on a live CUDA stream, and (b) leave
    the same parameters as the single-process backward_and_step."""
    from act_plus_plus_b200.engine import R_BACKBONE, R_CVAE, R_MAIN
    _, _, sd, cb, eng_a = _setup("small_k1")
    _, _, _, _, eng_b = _setup("small_k1")
    p0 = eng_a.flat_params.clone()
    for eng in (eng_a, eng_b):
        eng.set_lr(1e-3, 1e-3)
    eps = torch.randn((cb["qpos"].shape[0], 32), device="cuda", generator=torch.Generator(device="cuda").manual_seed(5))
    seen = []
    for _ in range(2):
        seen.clear()
        for eng in (eng_a, eng_b):
            eng.zero_grad()
            eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], eps)
        eng_a.backward_and_step_dp(lambda off, n, stream: seen.append((off, n, stream)), 1.0)
        eng_b.backward_and_step(1.0)
    torch.cuda.synchronize()
    want = sorted(eng_a.region(r) for r in (R_MAIN, R_CVAE, R_BACKBONE))
    got, merged = sorted((o, n) for o, n, _ in seen), []
    for o, n in got:                                    # merge adjacent ranges
        if merged and merged[-1][0] + merged[-1][1] == o:
            merged[-1] = (merged[-1][0], merged[-1][1] + n)
        else:
            assert not merged or merged[-1][0] + merged[-1][1] < o, f"overlapping ranges {merged[-1]} / {(o, n)}"
            merged.append((o, n))
    want_m = []
    for o, n in want:
        if want_m and want_m[-1][0] + want_m[-1][1] == o:
            want_m[-1] = (want_m[-1][0], want_m[-1][1] + n)
        else:
            want_m.append((o, n))
    assert merged == want_m, (merged, want_m)
    assert len(seen) >= 4 and all(s != 0 for _, _, s in seen)   # per-layer ranges, each on an internal stream
    for name, r in (("main", R_MAIN), ("cvae", R_CVAE), ("backbone", R_BACKBONE)):
        off, n = eng_b.region(r)
        da = (eng_a.flat_params[off:off + n] - eng_b.flat_params[off:off + n]).abs()
        moved = (eng_b.flat_params[off:off + n] - p0[off:off + n]).abs().mean().item()
        assert moved > 5e-4, f"{name}: parameters did not move ({moved})"
        assert da.mean().item() <= 0.02 * moved, f"{name}: {da.mean().item()} vs moved {moved}"


def test_grads_ready_hook_error_propagates():
    """An exception inside the averaging callback surfaces as that exception, not as a crash through the C frames."""
    _, _, sd, cb, eng = _setup("small_k1")
    eng.zero_grad()
    eng.forward(cb["qpos"], cb["image"], cb["actions"], cb["is_pad"], None)

    def bad(off, n, stream):
        raise ValueError("boom")
    with pytest.raises(ValueError, match="boom"):
        eng.backward_and_step_dp(bad, 1.0)
    torch.cuda.synchronize()
