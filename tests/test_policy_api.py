"""The reference-facing surface (policy.ACTPolicy / build_ACT_model_and_optimizer) driven the way
imitate_episodes.py drives it: make_policy -> forward_pass -> loss.backward() -> optimizer.step() -> eval call."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _policy_config(K=1, lr=1e-5, dropout=0.0, **extra):
    # imitate_episodes.py:78-94 (+ dropout 0 where a test compares against the deterministic goldens; the reference's
    # default 0.1 is exercised by test_train_mode_dropout_and_eval_mode / tests/test_dropout.py)
    return {'dropout': dropout, **extra, 'lr': lr, 'num_queries': 100, 'kl_weight': 10, 'hidden_dim': 512, 'dim_feedforward': 3200,
            'lr_backbone': 1e-5, 'backbone': 'resnet18', 'enc_layers': 4, 'dec_layers': 7, 'nheads': 8,
            'camera_names': [f'cam{k}' for k in range(K)], 'vq': False, 'vq_class': None, 'vq_dim': None,
            'action_dim': 16, 'no_encoder': False}


def _load_case(case):
    from oracle import act_oracle as O
    gold = np.load(os.path.join(GOLDEN, f"act_{case}.npz"))
    B, H, W, K, seed = [int(x) for x in gold["meta"]]
    cfg = O.make_cfg(num_cameras=K)
    sd = O.init_state_dict(cfg, seed)
    batch = {k: v.cuda() for k, v in O.make_inputs(cfg, B, H, W, seed).items()}
    # episodes are longer than the chunk: the policy slices [:, :num_queries] (policy.py:289-290)
    batch["actions"] = torch.cat([batch["actions"], torch.randn(B, 7, 16, device="cuda")], 1)
    batch["is_pad"] = torch.cat([batch["is_pad"], torch.ones(B, 7, dtype=torch.bool, device="cuda")], 1)
    return gold, cfg, sd, batch, K


def test_train_step_like_imitate_episodes():
    from act_plus_plus_b200.policy import ACTPolicy
    gold, cfg, sd, batch, K = _load_case("small_k1")
    policy = ACTPolicy(_policy_config(K))
    status = policy.deserialize({"model." + k: v for k, v in sd.items()})
    assert not status.missing_keys and not status.unexpected_keys
    policy.cuda(); policy.train()
    optimizer = policy.configure_optimizers()
    assert policy.model.num_queries == 100 and policy.model.encoder is not None and policy.kl_weight == 10
    before = {k: v.clone() for k, v in policy.serialize().items()}
    optimizer.zero_grad()
    out = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])     # actions longer than nq: sliced
    assert set(out) == {"l1", "kl", "loss"}
    loss = out["loss"]
    assert loss.requires_grad and loss.dim() == 0
    # l1 is independent of the reparametrisation noise up to its effect through z; kl is exactly noise-free
    assert abs(out["kl"].item() - float(gold["kl"])) / float(gold["kl"]) < 1e-3
    loss.backward()
    eng = policy.model._master
    g = {k: eng.grad_view(k).clone() for k in eng.tensors}
    assert float(g["action_head.weight"].abs().sum()) > 0
    assert float(g["transformer.decoder.layers.3.linear1.weight"].abs().sum()) == 0
    optimizer.step()
    torch.cuda.synchronize()
    after = policy.serialize()
    lr, wd = 1e-5, 1e-4
    for name in ("model.action_head.weight", "model.encoder.layers.2.linear1.weight", "model.backbones.0.0.body.layer3.0.conv1.weight",
                 "model.additional_pos_embed.weight"):
        p0, gg = before[name], g[name[6:]]
        m = 0.1 * gg
        v = 0.001 * gg * gg
        ref = p0 * (1 - lr * wd) - (lr / 0.1) * m / (v.sqrt() / (0.001 ** 0.5) + 1e-8)
        assert (after[name] - ref).abs().max().item() < 2e-7, name
    n = "model.transformer.decoder.layers.5.linear2.weight"          # dead layer: decoupled decay only
    assert (after[n] - before[n] * (1 - lr * wd)).abs().max().item() < 1e-9
    for n in ("model.is_pad_head.weight", "model.pos_table", "model.backbones.0.0.body.bn1.running_var"):
        assert torch.equal(after[n], before[n]), n                   # no grad / buffers: untouched
    # next forward uses the updated weights (shadows refreshed by the fused optimizer)
    out2 = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])
    assert out2["loss"].item() != loss.item()


def test_train_mode_dropout_and_eval_mode():
    """policy.train() applies the reference's dropout 0.1 to training calls (detr/main.py:45), policy.eval() is the
    deterministic validation path (imitate_episodes.py:566-571)."""
    from act_plus_plus_b200.policy import ACTPolicy
    gold, cfg, sd, batch, K = _load_case("small_k1")
    cfgd = _policy_config(K)
    del cfgd["dropout"]                                   # default, as imitate_episodes.py builds the dict
    policy = ACTPolicy(dict(cfgd, seed=3))
    policy.deserialize({"model." + k: v for k, v in sd.items()})
    assert policy.model.dropout_p == 0.1
    policy.train()
    with torch.no_grad():
        a = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])
        b = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])
    assert abs(a["l1"].item() - b["l1"].item()) > 1e-5     # fresh masks every call
    policy.eval()
    with torch.inference_mode():
        v = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])
    assert abs(v["kl"].item() - float(gold["kl"])) / float(gold["kl"]) < 1e-3
    names = [n for n, _ in policy.named_parameters()]
    assert "model.action_head.weight" in names and len(names) > 100 and len(list(policy.parameters())) == len(names)


def test_fused_step_with_cuda_graphs_matches_separate_calls():
    """args_override['fused_step']: loss.backward() is deferred to optimizer.step() (one overlapped engine call, replayed
    from CUDA graphs from the third step on).  Six steps must land on the same parameters as the plain
    backward-then-step policy (dropout 0, same reparametrisation noise)."""
    from act_plus_plus_b200.policy import ACTPolicy
    gold, cfg, sd, batch, K = _load_case("small_k1")
    res = []
    for fused in (False, True):
        policy = ACTPolicy(_policy_config(K, lr=1e-3, fused_step=fused))
        policy.deserialize({"model." + k: v for k, v in sd.items()})
        policy.train()
        opt = policy.configure_optimizers()
        losses = []
        for it in range(6):
            torch.manual_seed(100 + it)                    # the policy draws eps from torch's generator (detr_vae.py:21)
            out = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])
            out["loss"].backward()
            opt.step()
            opt.zero_grad()
            losses.append(out["loss"].item())
        torch.cuda.synchronize()
        if fused:
            assert policy.model._graphs, "the fused policy did not capture its step graphs"
            assert all(g.fwd is not None and g.bwd is not None for g in policy.model._graphs.values())
        res.append((losses, policy.serialize()["model.action_head.weight"].clone(), policy.model._master.step_count))
    (la, wa, sa), (lb, wb, sb) = res
    assert sa == sb == 6
    assert abs(la[0] - lb[0]) <= 1e-5 * abs(la[0])
    assert all(abs(x - y) <= 2e-2 * abs(x) for x, y in zip(la, lb)), (la, lb)     # Adam amplifies fp32 atomic-order noise
    assert la[-1] < la[0]
    moved = (wa - torch.from_numpy(np.asarray(sd["action_head.weight"])).cuda()).abs().mean().item()
    assert moved > 1e-3 and (wa - wb).abs().mean().item() < 0.05 * moved


def test_no_encoder_variant_through_the_policy():
    """no_encoder=True (detr_vae.py:109-112, 373-377; policy.py:306-307): kl = 0, loss = l1, no `encoder.*` keys."""
    from oracle import act_oracle as O
    from act_plus_plus_b200.policy import ACTPolicy
    gold = np.load(os.path.join(GOLDEN, "act_small_noenc_k1.npz"))
    B, H, W, K, seed = [int(x) for x in gold["meta"]]
    cfg = O.make_cfg(num_cameras=K, no_encoder=True)
    sd = O.init_state_dict(cfg, seed)
    batch = {k: v.cuda() for k, v in O.make_inputs(cfg, B, H, W, seed).items()}
    policy = ACTPolicy(dict(_policy_config(K), no_encoder=True))
    st = policy.deserialize({"model." + k: v for k, v in sd.items()})
    assert not st.missing_keys and not st.unexpected_keys
    assert policy.model.encoder is None and not any(k.startswith("model.encoder.") for k in policy.serialize())
    policy.train()
    out = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])
    assert out["kl"].item() == 0.0
    assert abs(out["loss"].item() - float(gold["loss"])) / float(gold["loss"]) < 1e-3
    out["loss"].backward()
    policy.configure_optimizers().step()
    a2, _, (mu, logvar), _, _ = policy.model(batch["qpos"], batch["image"], None, batch["actions"], batch["is_pad"])
    assert mu is None and logvar is None
    policy.eval()
    with torch.inference_mode():
        a_hat = policy(batch["qpos"], batch["image"])
    assert a_hat.shape == (B, 100, 16)


def test_inference_and_validation_calls():
    from act_plus_plus_b200.policy import ACTPolicy
    gold, cfg, sd, batch, K = _load_case("small_k3")
    policy = ACTPolicy(_policy_config(K))
    policy.deserialize({"model." + k: v for k, v in sd.items()})
    policy.eval()
    with torch.inference_mode():
        a_hat = policy(batch["qpos"], batch["image"])
        val = policy(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])
    assert a_hat.shape == (batch["qpos"].shape[0], 100, 16)
    assert (a_hat.cpu() - torch.from_numpy(gold["a_hat_infer"])).abs().max().item() <= 1e-2
    assert not val["loss"].requires_grad
    assert abs(val["kl"].item() - float(gold["kl"])) / float(gold["kl"]) < 1e-3
    # model() keeps the reference's 5-tuple (detr_vae.py:254)
    a2, pad_hat, (mu, logvar), probs, binaries = policy.model(batch["qpos"], batch["image"], None)
    assert pad_hat.shape == (batch["qpos"].shape[0], 100, 1) and mu is None and probs is None and binaries is None


def test_unsupported_configs_raise():
    from act_plus_plus_b200.policy import ACTPolicy
    bad = _policy_config()
    bad["vq"] = True
    with pytest.raises(NotImplementedError):
        ACTPolicy(bad)
    bad = _policy_config()
    bad["backbone"] = "resnet50"
    with pytest.raises(ValueError):
        ACTPolicy(bad)


def test_batched_rollout_matches_reference_ensemble_lines():
    """BatchedRollout (policy query + ring-buffer ensemble kernel, 3 envs) == the literal imitate_episodes.py:403-411
    buffer fed with the same policy outputs; also env sharding and the device-side (de)normalisation."""
    from oracle import act_oracle as O
    from act_plus_plus_b200.rollout import BatchedRollout
    from act_plus_plus_b200.policy import ACTPolicy
    torch.manual_seed(0)
    pol = ACTPolicy(dict(_policy_config(1), seed=5))
    pol.eval()
    E, T, H, W = 3, 6, 96, 128
    nq, A = pol.model.num_queries, pol.model.action_dim
    stats = dict(qpos_mean=torch.randn(14), qpos_std=torch.rand(14) + 0.5, action_mean=torch.randn(A), action_std=torch.rand(A) + 0.5)
    ro = BatchedRollout(pol, num_envs_global=E, stats=stats)
    half = BatchedRollout(pol, num_envs_global=E, rank=1, world=2)
    assert (ro.E, half.lo, half.hi) == (3, 2, 3)
    oracles = [O.TemporalAggOracle(T, nq, A) for _ in range(E)]
    g = torch.Generator().manual_seed(1)
    for t in range(T):
        qpos = torch.randn((E, 14), generator=g).cuda()
        image = torch.rand((E, 1, 3, H, W), generator=g).cuda()
        out = ro.step(qpos, image)
        assert out.dtype == torch.float64 and out.shape == (E, A)
        with torch.inference_mode():
            chunks = pol((qpos - stats["qpos_mean"].cuda()) / stats["qpos_std"].cuda(), image).cpu()
        for e in range(E):
            ref = oracles[e].step(t, chunks[e][None])[0] * stats["action_std"].double() + stats["action_mean"].double()
            assert torch.allclose(out[e].cpu(), ref, rtol=1e-12, atol=1e-12), (t, e)


def test_open_loop_replay_with_base_delay_matches_reference_lines():
    """No temporal_agg, real-robot timing (imitate_episodes.py:307, 311-313, 401, 415): query every nq - BASE_DELAY steps,
    all_actions = cat([a[:, :-D, :-2], a[:, D:, -2:]], 2), raw_action = all_actions[:, t % query_frequency]; uint8 frames,
    raw qpos and de-normalisation handled on the device."""
    from act_plus_plus_b200.rollout import BatchedRollout
    from act_plus_plus_b200.policy import ACTPolicy
    torch.manual_seed(0)
    pol = ACTPolicy(dict(_policy_config(1), seed=6))
    pol.eval()
    E, H, W, D = 2, 96, 128, 13
    nq, A = pol.model.num_queries, pol.model.action_dim
    stats = dict(qpos_mean=torch.randn(14), qpos_std=torch.rand(14) + 0.5, action_mean=torch.randn(A), action_std=torch.rand(A) + 0.5)
    ro = BatchedRollout(pol, num_envs_global=E, stats=stats, temporal_agg=False, base_delay=D)
    qf = nq - D
    g = torch.Generator().manual_seed(2)
    all_actions = None
    for t in range(qf + 5):
        qpos = torch.randn((E, 14), generator=g).cuda()
        u8 = torch.randint(0, 256, (E, 1, H, W, 3), generator=g, dtype=torch.uint8).cuda()
        out = ro.step(qpos, u8)
        if t % qf == 0:
            img = (u8.permute(0, 1, 4, 2, 3).float() / 255.0).contiguous()
            with torch.inference_mode():
                a = pol((qpos - stats["qpos_mean"].cuda()) / stats["qpos_std"].cuda(), img).cpu()
            all_actions = torch.cat([a[:, :-D, :-2], a[:, D:, -2:]], dim=2)                      # :401
        raw = all_actions[:, t % qf].double()                                                   # :415
        ref = raw * stats["action_std"].double() + stats["action_mean"].double()                # :294
        assert torch.allclose(out.cpu(), ref, rtol=1e-12, atol=1e-12), t


def test_input_prefetcher_delivers_batches_in_order():
    from act_plus_plus_b200.train import InputPrefetcher
    host = [dict(x=torch.full((1 << 20,), float(i)).pin_memory(), y=torch.arange(8).add(i).pin_memory()) for i in range(5)]
    pf = InputPrefetcher(host[0], "cuda", depth=2)
    pf.submit(host[0])
    seen = []
    for i in range(5):
        if i + 1 < 5:
            pf.submit(host[i + 1])
        b = pf.next()
        seen.append((b["x"].sum().item() / (1 << 20), b["y"].clone()))     # consume on the current stream
        pf.consumed()
    torch.cuda.synchronize()
    for i, (x, y) in enumerate(seen):
        assert x == float(i)
        assert torch.equal(y.cpu(), torch.arange(8) + i)


def test_uint8_hwc_ingest_equals_float_path():
    """Dataset-native uint8 [B,K,H,W,3] frames (utils.py:133-158) through the fused ingest == the reference-side
    `einsum('k h w c -> k c h w') / 255.0` float path, bit for bit (same fp32 division, same bf16 packing)."""
    from act_plus_plus_b200.engine import ActEngine
    from act_plus_plus_b200.policy import _init_like_reference
    B, K, H, W = 2, 1, 96, 128
    g = torch.Generator(device="cpu").manual_seed(11)
    u8 = torch.randint(0, 256, (B, K, H, W, 3), generator=g, dtype=torch.uint8)
    f32 = (u8.permute(0, 1, 4, 2, 3).float() / 255.0).contiguous()
    qpos = torch.randn((B, 14), generator=g)
    eng = ActEngine(batch_max=B, num_cameras=K, img_h=H, img_w=W, train=False)
    _init_like_reference(eng, seed=2)
    a = eng.forward(qpos.cuda(), f32.cuda())["a_hat"].clone()
    b = eng.forward(qpos.cuda(), u8.cuda())["a_hat"].clone()
    c = eng.forward(qpos.cuda(), f32.cuda())["a_hat"].clone()       # switching back works too
    torch.cuda.synchronize()
    assert torch.equal(a, b) and torch.equal(a, c)
    assert a.abs().max().item() > 0


def test_checkpoint_file_round_trip(tmp_path):
    """imitate_episodes.py:614-619 saves `policy.serialize()` with torch.save and :168-177 restores it with
    `policy.deserialize(torch.load(...))`: a checkpoint written by one policy object must reproduce the same actions in
    a fresh one, keep the reference's key names, and the optimiser state (which the reference never saves) round-trips."""
    from act_plus_plus_b200.policy import ACTPolicy
    gold, cfg, sd, batch, K = _load_case("small_k1")
    pol = ACTPolicy(_policy_config(K))
    pol.deserialize({"model." + k: v for k, v in sd.items()})
    pol.cuda(); pol.train()
    opt = pol.configure_optimizers()
    for _ in range(2):
        opt.zero_grad()
        pol(batch["qpos"], batch["image"], batch["actions"], batch["is_pad"])["loss"].backward()
        opt.step()
    pol.eval()
    with torch.inference_mode():
        a_ref = pol(batch["qpos"], batch["image"]).clone()
    ck = tmp_path / "policy_last.ckpt"
    torch.save(pol.serialize(), ck)
    torch.save(opt.state_dict(), tmp_path / "optim.ckpt")
    loaded = torch.load(ck, map_location="cpu")
    assert set(loaded) == {"model." + k for k in sd}, "checkpoint keys must be the reference's state_dict names"
    pol2 = ACTPolicy(_policy_config(K))
    st = pol2.deserialize(loaded)
    assert not st.missing_keys and not st.unexpected_keys
    pol2.cuda(); pol2.eval()
    with torch.inference_mode():
        a2 = pol2(batch["qpos"], batch["image"]).clone()
    assert torch.equal(a_ref, a2)
    # optimiser state: continue training from the restored moments == continue in the original object
    opt2 = pol2.configure_optimizers()
    opt2.load_state_dict(torch.load(tmp_path / "optim.ckpt", map_location="cpu"))
    eps = torch.randn((batch["qpos"].shape[0], 32), device="cuda")
    outs = []
    for p, o in ((pol, opt), (pol2, opt2)):
        p.train()
        eng = p.model._engine_for(batch["qpos"], batch["image"], True)
        eng.zero_grad()
        eng.forward(batch["qpos"], batch["image"], batch["actions"][:, :100], batch["is_pad"][:, :100], eps)
        eng.backward(0, 1.0)
        o.step()
        outs.append(p.serialize()["model.action_head.weight"].clone())
    torch.cuda.synchronize()
    assert torch.allclose(outs[0], outs[1], rtol=0, atol=2e-6)
