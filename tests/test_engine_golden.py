"""End-to-end parity of the CUDA engine against golden vectors produced by the UNMODIFIED reference
(oracle/make_golden.py): a_hat / mu / logvar max-abs <= 1e-2, losses relative <= 1e-3 (BASELINE.json tolerances)."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["small_k1", "small_k3", "mid_k1", "full_k1_b8"]
TOL_ABS, TOL_LOSS = 1e-2, 1e-3


def _setup(case):
    from oracle import act_oracle as O
    from act_plus_plus_b200.engine import ActEngine
    gold = np.load(os.path.join(GOLDEN, f"act_{case}.npz"))
    B, H, W, K, seed = [int(x) for x in gold["meta"]]
    cfg = O.make_cfg(num_cameras=K)
    sd = O.init_state_dict(cfg, seed)
    batch = O.make_inputs(cfg, B, H, W, seed)
    assert np.allclose(O.state_checksum(sd), gold["state_checksum"], rtol=1e-9), "seeded weights differ from the golden run"
    eng = ActEngine(batch_max=B, num_cameras=K, img_h=H, img_w=W, train=True)
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
        errs[k] = float(abs(ls[i] - gold[k]) / abs(gold[k]))
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
