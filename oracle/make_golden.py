"""Generate tests/golden/*.npz by running the UNMODIFIED reference on CPU (build container only).

TEST INFRASTRUCTURE ONLY.  Run here (``python -m oracle.make_golden``); `/root/reference` does not
exist on the GPU box, so nothing imports this module at test time.  For every case it
  1. regenerates the seeded state_dict / inputs from ``oracle.act_oracle``,
  2. loads the state_dict (strict) into the reference's own ``DETRVAE`` (SURVEY.md 8c recipe),
  3. runs the reference forward + the loss lines of policy.py:309-318 + backward,
  4. asserts the oracle restatement reproduces the reference (fp32, <= 2e-5), and
  5. stores outputs, per-parameter grad norms and strided grad samples as the golden vector.
"""
from __future__ import annotations

import argparse
import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

from oracle import act_oracle as O

REF = "/root/reference"
GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = {
    # name: (batch, H, W, num_cameras, seed)
    "small_k1": (2, 96, 128, 1, 0),
    "small_k3": (2, 96, 128, 3, 1),
    "mid_k1": (3, 224, 288, 1, 2),
    "full_k1_b8": (8, 480, 640, 1, 3),      # BASELINE config C1/C2 shape
    "full_k3_b2": (2, 480, 640, 3, 4),      # BASELINE configs[3] shape per sample: 3 cameras -> 902 encoder tokens
}
# model variants of SURVEY.md 8(f)-4: name -> (batch, H, W, num_cameras, seed, config overrides)
VARIANTS = {
    "small_noenc_k1": (2, 96, 128, 1, 6, dict(no_encoder=True)),                 # detr_vae.py:109-112, policy.py:306-307
    "small_s7a6_k4": (2, 96, 128, 4, 7, dict(state_dim=7, action_dim=6)),        # train_single_arm.py:72,93: 7-DoF arm, 4 cameras
}
# inference-only cases (no backward): per-GPU batch of BASELINE configs[2] at N=4 / the rollout leg's many-env regime
INFER_CASES = {
    "full_k1_b16_infer": (16, 480, 640, 1, 5),
}
STAGE_SAMPLES = 8192


def _import_reference():
    sys.modules.setdefault("IPython", types.SimpleNamespace(embed=lambda *a, **k: None))
    for p in (os.path.join(REF, "detr"), REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    import detr.models.backbone as bb
    bb.is_main_process = lambda: False           # no ImageNet download (backbone.py:123)
    import detr.models.detr_vae as dv
    return dv


def build_reference_model(cfg):
    dv = _import_reference()
    from detr.models.backbone import build_backbone
    from detr.models.transformer import build_transformer
    args = argparse.Namespace(
        hidden_dim=cfg["hidden_dim"], dropout=0.1, nheads=cfg["nheads"], dim_feedforward=cfg["dim_feedforward"],
        enc_layers=cfg["enc_layers"], dec_layers=cfg["dec_layers"], pre_norm=False, position_embedding="sine",
        lr_backbone=cfg["lr_backbone"], masks=False, backbone="resnet18", dilation=False,
        camera_names=[f"cam{k}" for k in range(cfg["num_cameras"])])
    backbones = [build_backbone(args) for _ in args.camera_names]
    encoder = None if cfg.get("no_encoder") else dv.build_encoder(args)            # detr_vae.py:373-377
    model = dv.DETRVAE(backbones, build_transformer(args), encoder, cfg["state_dim"],
                       cfg["num_queries"], args.camera_names, False, None, None, cfg["action_dim"], None, None)
    return dv, model


def reference_run(cfg, sd, batch, train_call=True):
    dv, model = build_reference_model(cfg)
    missing = model.load_state_dict(sd, strict=True)
    model.eval()                                   # validation path: dropout off (imitate_episodes.py:566-571)
    eps = batch["eps"]
    dv.reparametrize = lambda mu, logvar: mu + logvar.div(2).exp() * eps   # inject the noise of detr_vae.py:19-22
    image = O.normalize_image(batch["image"])
    nq = cfg["num_queries"]
    if train_call:
        actions, is_pad = batch["actions"][:, :nq], batch["is_pad"][:, :nq]
        a_hat, is_pad_hat, (mu, logvar), _, _ = model(batch["qpos"], image, None, actions, is_pad)
        if model.encoder is None:                                           # policy.py:306-307
            total_kld = torch.zeros(1)
            mu = logvar = torch.zeros((batch["qpos"].shape[0], 32))
        else:
            klds = -0.5 * (1 + logvar - mu.pow(2) - logvar.exp())           # policy.py:386-387
            total_kld = klds.sum(1).mean(0, True)
        all_l1 = F.l1_loss(actions, a_hat, reduction="none")                # policy.py:314
        l1 = (all_l1 * ~is_pad.unsqueeze(-1)).mean()
        loss = l1 + total_kld[0] * cfg["kl_weight"]
        loss.backward()
        grads = {n: p.grad for n, p in model.named_parameters()}
        return dict(a_hat=a_hat.detach(), mu=mu.detach(), logvar=logvar.detach(), l1=l1.detach(),
                    kl=total_kld[0].detach(), loss=loss.detach(), is_pad_hat=is_pad_hat.detach()), grads
    with torch.inference_mode():
        a_hat, _, _, _, _ = model(batch["qpos"], image, None)
    return dict(a_hat=a_hat), None


def _stage_record(taps: dict) -> dict:
    """Strided samples + norms of the intermediate activations (per-stage parity diagnostics: where does the bf16
    engine spend its a_hat error budget)."""
    rec = {}
    for k, t in taps.items():
        f = t.detach().flatten()
        stride = max(1, f.numel() // STAGE_SAMPLES)
        rec[f"stage_{k}"] = f[::stride][:STAGE_SAMPLES].float().numpy()
        rec[f"stage_{k}_meta"] = np.asarray([f.numel(), stride, float(f.double().norm()), float(f.abs().max())])
    return rec


def make_infer_case(name: str):
    B, H, W, K, seed = INFER_CASES[name]
    cfg = O.make_cfg(num_cameras=K)
    sd = O.init_state_dict(cfg, seed)
    batch = O.make_inputs(cfg, B, H, W, seed)
    torch.set_num_threads(os.cpu_count())
    ref_inf, _ = reference_run(cfg, sd, batch, False)
    oi = O.policy_forward(sd, cfg, batch["qpos"], batch["image"])
    assert float((oi["a_hat"] - ref_inf["a_hat"]).abs().max()) <= 2e-5
    rec = {"meta": np.asarray([B, H, W, K, seed], dtype=np.int64), "state_checksum": O.state_checksum(sd),
           "a_hat_infer": ref_inf["a_hat"].numpy()}
    np.savez_compressed(os.path.join(GOLDEN_DIR, f"act_{name}.npz"), **rec)
    print(f"[{name}] inference golden written, |a_hat|max={float(ref_inf['a_hat'].abs().max()):.4f}")


def _sample(t: torch.Tensor, n: int = 256) -> np.ndarray:
    f = t.detach().flatten()
    stride = max(1, f.numel() // n)
    return f[::stride][:n].double().numpy()


def make_case(name: str, check_oracle: bool = True):
    B, H, W, K, seed, over = (CASES[name] + ({},)) if name in CASES else VARIANTS[name]
    cfg = O.make_cfg(num_cameras=K, **over)
    sd = O.init_state_dict(cfg, seed)
    batch = O.make_inputs(cfg, B, H, W, seed)
    torch.set_num_threads(os.cpu_count())
    ref, grads = reference_run(cfg, sd, batch, True)
    ref_inf, _ = reference_run(cfg, sd, batch, False)
    rec = {
        "meta": np.asarray([B, H, W, K, seed], dtype=np.int64),
        "state_checksum": O.state_checksum(sd),
        "input_checksum": np.asarray([float(batch["image"].double().sum()), float(batch["qpos"].double().sum()),
                                      float(batch["actions"].double().sum()), float(batch["is_pad"].sum()),
                                      float(batch["eps"].double().sum())]),
        "a_hat": ref["a_hat"].numpy(), "mu": ref["mu"].numpy(), "logvar": ref["logvar"].numpy(),
        "l1": ref["l1"].numpy(), "kl": ref["kl"].numpy(), "loss": ref["loss"].numpy(),
        "a_hat_infer": ref_inf["a_hat"].numpy(),
        "cfg_over": np.asarray(repr(over)),
    }
    names = O.trainable_names(cfg)
    gnorm = np.zeros(len(names))
    gnone = np.zeros(len(names), dtype=np.int64)
    gsamp = np.zeros((len(names), 256))
    for i, n in enumerate(names):
        g = grads[n]
        if g is None:
            gnone[i] = 1
            continue
        gnorm[i] = float(g.double().norm())
        s = _sample(g)
        gsamp[i, :len(s)] = s
    rec["grad_names"] = np.asarray(names)
    rec["grad_norm"], rec["grad_none"], rec["grad_sample"] = gnorm, gnone, gsamp
    if name.startswith("full"):
        taps = {}
        with torch.no_grad():
            ot = O.policy_forward(sd, cfg, batch["qpos"], batch["image"], batch["actions"], batch["is_pad"], batch["eps"], taps=taps)
        assert float((ot["a_hat"] - ref["a_hat"]).abs().max()) <= 2e-5
        rec.update(_stage_record(taps))
    if check_oracle:
        o = O.loss_and_grads(sd, cfg, batch)
        for k in ("a_hat", "mu", "logvar", "l1", "kl", "loss"):
            if o["out"][k] is None:
                assert cfg.get("no_encoder") and k in ("mu", "logvar")
                continue
            err = float((o["out"][k] - ref[k]).abs().max())
            assert err <= 2e-5 * max(1.0, float(ref[k].abs().max())), (name, k, err)
        oi = O.policy_forward(sd, cfg, batch["qpos"], batch["image"])
        assert float((oi["a_hat"] - ref_inf["a_hat"]).abs().max()) <= 2e-5
        worst = 0.0
        for n in names:
            g, go = grads[n], o["grads"][n]
            if g is None:
                assert go is None, n
                continue
            if go is None:      # dead decoder layers: reference gives exact zeros, the restatement skips them
                assert float(g.abs().max()) == 0.0, n
                continue
            rel = float((g - go).norm() / (g.norm() + 1e-12))
            worst = max(worst, rel)
        assert worst < 1e-3, (name, "grad rel", worst)
        print(f"[{name}] oracle == reference: outputs <=2e-5, worst grad rel err {worst:.2e}")
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    np.savez_compressed(os.path.join(GOLDEN_DIR, f"act_{name}.npz"), **rec)
    print(f"[{name}] loss={float(ref['loss']):.6f} l1={float(ref['l1']):.6f} kl={float(ref['kl']):.6f} "
          f"|a_hat|max={float(ref['a_hat'].abs().max()):.4f}")


def make_temporal_agg():
    """Literal imitate_episodes.py:403-411 on a random chunk stream (with exact zeros planted to
    exercise the `!= 0` population filter).  Two fixtures: a small one (chunks committed) and the BASELINE shape
    nq=100 / T=400 (the ring wraps 4x; only the float64 outputs are committed, the stream is regenerated from its seed)."""
    T, nq, A = 60, 20, 16
    agg = O.TemporalAggOracle(T, nq, A)
    chunks = O.temporal_agg_stream(T, nq, A, 7)
    outs = np.stack([agg.step(t, chunks[t]).numpy()[0] for t in range(T)])
    np.savez_compressed(os.path.join(GOLDEN_DIR, "temporal_agg.npz"), chunks=chunks[:, 0].numpy(), out=outs,
                        meta=np.asarray([T, nq, A]))
    print("[temporal_agg] golden written", outs.shape, outs.dtype)
    T, nq, A = 400, 100, 16
    agg = O.TemporalAggOracle(T, nq, A)
    chunks = O.temporal_agg_stream(T, nq, A, 11)
    outs = np.stack([agg.step(t, chunks[t]).numpy()[0] for t in range(T)])
    np.savez_compressed(os.path.join(GOLDEN_DIR, "temporal_agg_full.npz"), out=outs, meta=np.asarray([T, nq, A, 11]),
                        chunk_checksum=np.asarray([float(chunks.double().sum()), float((chunks == 0).sum())]))
    print("[temporal_agg_full] golden written", outs.shape, outs.dtype)


if __name__ == "__main__":
    which = sys.argv[1:] or list(CASES) + list(VARIANTS) + list(INFER_CASES) + ["temporal_agg"]
    for w in which:
        if w == "temporal_agg":
            make_temporal_agg()
        elif w in INFER_CASES:
            make_infer_case(w)
        else:
            make_case(w)
