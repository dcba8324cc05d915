"""Drop-in mirror of the reference's ``policy.ACTPolicy`` / ``detr.main.build_ACT_model_and_optimizer``.

Same constructor dict, call signatures, return types, attribute names and error behaviour as
``/root/reference/policy.py:243-348`` and ``/root/reference/detr/main.py:92-112``; the work underneath is done by
the CUDA engine (``act_engine_*`` C ABI).  There is no PyTorch/CPU fallback: without the built library or a CUDA
device construction raises.

Deviations from the reference, all deliberate and documented in DESIGN.md:
  * dropout (0.1 in the reference's train mode) is not applied -- parity is defined on the reference's
    deterministic path (``.eval()`` / dropout 0), SURVEY.md 8c;
  * ``state_dim`` is a config key (default 14, the ALOHA qpos size) instead of the fork's hard-coded 7
    (detr_vae.py:345); ``depth_camera_names`` may be absent (policy.py:249 raises KeyError upstream);
  * ``sys.argv`` is never parsed (detr/main.py:93-94 does).
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, Optional

import torch
import torch.nn as nn

from .engine import (KIND_BACKBONE, KIND_BUFFER, KIND_NOGRAD, R_BACKBONE, R_CVAE, R_MAIN, ActEngine)

_DEFAULTS = dict(  # detr/main.py:12-60 argparse defaults that ACT never overrides, plus the overridden ones
    lr=1e-4, lr_backbone=1e-5, weight_decay=1e-4, backbone="resnet18", dilation=False, position_embedding="sine",
    camera_names=[], enc_layers=4, dec_layers=6, dim_feedforward=2048, hidden_dim=256, dropout=0.1, nheads=8,
    num_queries=400, pre_norm=False, masks=False, use_pcd=False, use_depth=False, vq=False, vq_class=None, vq_dim=None,
    action_dim=16, no_encoder=False, state_dim=14, kl_weight=10,
)


def _check_supported(a: dict):
    if a.get("vq"):
        raise NotImplementedError("vq=True (VQ-CVAE, detr_vae.py:137-145) is outside the B200 hot path")
    if a.get("use_depth") or a.get("use_pcd"):
        raise NotImplementedError("use_depth / use_pcd branches (detr_vae.py:188-210) are outside the B200 hot path")
    if a.get("backbone", "resnet18") != "resnet18":
        raise ValueError(f"backbone {a.get('backbone')!r} not supported (resnet18 only)")
    if a.get("pre_norm"):
        raise ValueError("pre_norm=True not supported (the reference's ACT configs are post-norm)")
    if a.get("position_embedding", "sine") not in ("sine", "v2"):
        raise ValueError("only the sine position embedding is supported")
    if a.get("no_encoder"):
        raise NotImplementedError("no_encoder=True (detr_vae.py:109-112) is not built yet")
    if a.get("masks") or a.get("dilation"):
        raise ValueError("masks / dilation are not supported")
    if a["hidden_dim"] != 512 or a["nheads"] != 8:
        raise ValueError("the sm_100a kernels are specialised for hidden_dim=512, nheads=8 (the ACT configuration)")


def _init_like_reference(eng: ActEngine, seed: Optional[int] = None):
    """Random init with the reference's families: xavier_uniform on transformer.* matrices (transformer.py:44-47),
    nn.Linear / nn.Embedding / nn.MultiheadAttention / LayerNorm defaults elsewhere, torchvision's kaiming-normal
    (fan_out) for the ResNet convs, identity FrozenBatchNorm buffers (backbone.py:29-34; no ImageNet download here)."""
    g = torch.Generator(device="cpu")
    if seed is not None:
        g.manual_seed(seed)
    else:
        g.manual_seed(torch.initial_seed() & 0x7FFFFFFF)
    from .policy_tables import sinusoid_table
    with torch.no_grad():
        for name, t in eng.tensors.items():
            shape = t["shape"]
            v = eng.view(name)
            leaf = name.split(".")[-1]
            if name == "pos_table":
                val = sinusoid_table(shape[1], shape[2])
            elif "running_var" in name or (t["kind"] == KIND_BUFFER and leaf == "weight"):
                val = torch.ones(shape)
            elif t["kind"] == KIND_BUFFER:
                val = torch.zeros(shape)
            elif leaf in ("in_proj_bias",) or (leaf == "bias" and "out_proj" in name):
                val = torch.zeros(shape)
            elif "norm" in name and leaf == "weight":
                val = torch.ones(shape)
            elif "norm" in name and leaf == "bias":
                val = torch.zeros(shape)
            elif name.startswith("backbones.") and len(shape) == 4:
                fan_out = shape[0] * shape[2] * shape[3]
                val = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_out)
            elif leaf in ("query_embed.weight",) or name in ("query_embed.weight", "cls_embed.weight", "additional_pos_embed.weight"):
                val = torch.randn(shape, generator=g)
            elif len(shape) >= 2 and (name.startswith("transformer.") or leaf == "in_proj_weight"):
                fan_out, fan_in = shape[0], int(torch.tensor(shape[1:]).prod())
                a = math.sqrt(6.0 / (fan_in + fan_out))
                val = (torch.rand(shape, generator=g) * 2 - 1) * a
            elif len(shape) >= 2:
                fan_in = int(torch.tensor(shape[1:]).prod())
                val = (torch.rand(shape, generator=g) * 2 - 1) / math.sqrt(fan_in)
            else:   # biases of nn.Linear / nn.Conv2d: U(-1/sqrt(fan_in), 1/sqrt(fan_in)) of the matching weight
                wname = name[:-len("bias")] + "weight"
                fan_in = int(torch.tensor(eng.tensors[wname]["shape"][1:]).prod()) if wname in eng.tensors else shape[0]
                val = (torch.rand(shape, generator=g) * 2 - 1) / math.sqrt(fan_in)
            v.copy_(val.to(v.device))
    eng.refresh()


class _ActStep(torch.autograd.Function):
    """Bridges ``loss.backward()`` (imitate_episodes.py:605-606) to ``act_engine_backward``."""

    @staticmethod
    def forward(ctx, anchor, model, qpos, image, actions, is_pad, eps):
        out = model._engine_for(qpos, image, train=True).forward(qpos, image, actions, is_pad, eps)
        ctx.model = model
        model._last = out
        return out["losses"].clone()

    @staticmethod
    def backward(ctx, g):
        model = ctx.model
        g = g.detach().float().cpu()
        if g[0] != 0 or g[1] != 0:
            raise NotImplementedError("backward is supported through loss_dict['loss'] only (l1 + kl_weight * kl)")
        model._run_backward(float(g[2]))
        return torch.zeros(1, device=model.device), None, None, None, None, None, None


class DETRVAE(nn.Module):
    """Stand-in for ``detr.models.detr_vae.DETRVAE`` (detr_vae.py:36-254): same call signature and return tuple."""

    def __init__(self, args: dict):
        super().__init__()
        if not torch.cuda.is_available():
            raise RuntimeError("act_plus_plus_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.args = dict(args)
        self.num_queries = args["num_queries"]
        self.camera_names = list(args["camera_names"])
        self.state_dim, self.action_dim = args["state_dim"], args["action_dim"]
        self.latent_dim = 32
        self.vq, self.vq_class, self.vq_dim = False, args.get("vq_class"), args.get("vq_dim")
        self.device = torch.device(f"cuda:{torch.cuda.current_device()}")
        self._anchor = nn.Parameter(torch.zeros(1, device=self.device))
        self._engines: Dict[tuple, ActEngine] = {}
        self._dp_group = None
        self._last = None
        self._step = 0
        # a minimal engine fixes the parameter layout and owns the master tensors; compute engines share them
        self._master = self._make_engine(1, 64, 64, train=True)
        self.encoder = object()          # policy.py:306 only tests `self.model.encoder is None`
        _init_like_reference(self._master, args.get("seed"))

    # -------------------------------------------------------------- engines
    def _make_engine(self, batch, h, w, train, share: Optional[ActEngine] = None) -> ActEngine:
        a = self.args
        eng = ActEngine(batch_max=batch, num_cameras=len(self.camera_names), img_h=h, img_w=w,
                        hidden_dim=a["hidden_dim"], nheads=a["nheads"], dim_feedforward=a["dim_feedforward"],
                        enc_layers=a["enc_layers"], dec_layers=a["dec_layers"], num_queries=a["num_queries"],
                        state_dim=a["state_dim"], action_dim=a["action_dim"], latent_dim=32, kl_weight=float(a["kl_weight"]),
                        lr=a["lr"], lr_backbone=a["lr_backbone"], weight_decay=a["weight_decay"], train=train,
                        device=self.device) if share is None else ActEngine.sharing(share, batch, h, w, train)
        return eng

    def _engine_for(self, qpos, image, train: bool) -> ActEngine:
        B = int(qpos.shape[0])
        if image.dtype == torch.uint8:                 # dataset-native frames [B, K, H, W, 3]
            H, W = int(image.shape[2]), int(image.shape[3])
        else:
            H, W = int(image.shape[-2]), int(image.shape[-1])
        key = (H, W, bool(train))
        eng = self._engines.get(key)
        if eng is None or eng.cfg.batch_max < B:
            eng = self._make_engine(B, H, W, train, share=self._master)
            self._engines[key] = eng
        return eng

    def _run_backward(self, loss_grad: float):
        eng = next(e for e in self._engines.values() if e.train_enabled and e.last_was_train)
        if self._dp_group is None:
            eng.backward(0, loss_grad)
            return
        from .train import backward_data_parallel
        backward_data_parallel(eng, self._dp_group, loss_grad)

    # -------------------------------------------------------------- nn.Module surface
    def forward(self, qpos, image, env_state, actions=None, is_pad=None, vq_sample=None, pointcloud=None, depth_img=None):
        """image must already be ImageNet-normalised?  No: like everything in this package the engine fuses the
        normalisation of policy.py:268-272 into conv1, so DETRVAE.forward takes the *un-normalised* [0,1] image and
        ACTPolicy does not normalise separately."""
        if pointcloud is not None or depth_img is not None or vq_sample is not None:
            raise NotImplementedError("pointcloud / depth / vq inputs are outside the B200 hot path")
        if actions is not None:
            eps = torch.randn((qpos.shape[0], self.latent_dim), device=qpos.device, dtype=torch.float32)
            out = self._engine_for(qpos, image, True).forward(qpos, image, actions, is_pad, eps, want_is_pad_hat=True)
            return out["a_hat"], out["is_pad_hat"], [out["mu"], out["logvar"]], None, None
        out = self._engine_for(qpos, image, False).forward(qpos, image, want_is_pad_hat=True)
        return out["a_hat"], out["is_pad_hat"], [None, None], None, None

    def state_dict(self, *args, destination=None, prefix="", keep_vars=False, **kw):
        sd = OrderedDict() if destination is None else destination
        for k, v in self._master.state_dict().items():
            sd[prefix + k] = v if keep_vars else v.detach()
        return sd

    def load_state_dict(self, state_dict, strict: bool = True, prefix: str = ""):
        sd = {k[len(prefix):]: v for k, v in state_dict.items() if k.startswith(prefix)}
        missing, unexpected = self._master.load_state_dict(sd, strict=strict)
        for e in self._engines.values():
            e.refresh()
        return torch.nn.modules.module._IncompatibleKeys(missing, unexpected)

    def named_parameters(self, prefix: str = "", recurse: bool = True, remove_duplicate: bool = True):
        for k, t in self._master.tensors.items():
            if t["kind"] != KIND_BUFFER:
                yield prefix + ("." if prefix else "") + k, self._master.param(k)

    def parameters(self, recurse: bool = True):
        for _, p in self.named_parameters():
            yield p

    def cuda(self, device=None):
        return self

    def to(self, *a, **k):
        return self


class ActAdamW:
    """Stand-in for the ``torch.optim.AdamW`` built at detr/main.py:102-110: two groups (non-backbone @ lr,
    backbone @ lr_backbone), weight_decay 1e-4, betas (0.9, 0.999).  ``step()`` is one fused multi-tensor kernel per
    optimizer region; parameters whose gradient is None in the reference (is_pad_head) are skipped, parameters the
    reference reaches with exact zeros (decoder layers 1..) are only weight-decayed -- SURVEY.md 3.2."""

    def __init__(self, model: DETRVAE, lr: float, lr_backbone: float, weight_decay: float):
        self.model = model
        named = list(model.named_parameters())
        self.param_groups = [
            dict(params=[p for n, p in named if "backbone" not in n], lr=lr, weight_decay=weight_decay, betas=(0.9, 0.999), eps=1e-8),
            dict(params=[p for n, p in named if "backbone" in n], lr=lr_backbone, weight_decay=weight_decay, betas=(0.9, 0.999), eps=1e-8),
        ]

    def _engine(self) -> ActEngine:
        for e in self.model._engines.values():
            if e.train_enabled and e.last_was_train:
                return e
        return self.model._master

    def zero_grad(self, set_to_none: bool = True):
        self.model._master.zero_grad()

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        eng = self._engine()
        eng.set_lr(float(self.param_groups[0]["lr"]), float(self.param_groups[1]["lr"]))
        eng.adamw_step()
        return loss

    def state_dict(self):
        m = self.model._master
        return dict(step=m.step_count, exp_avg=m.exp_avg.clone(), exp_avg_sq=m.exp_avg_sq.clone(),
                    lr=[g["lr"] for g in self.param_groups])

    def load_state_dict(self, sd):
        m = self.model._master
        m._shared["step"] = int(sd["step"])
        m.exp_avg.copy_(sd["exp_avg"]); m.exp_avg_sq.copy_(sd["exp_avg_sq"])
        for g, lr in zip(self.param_groups, sd.get("lr", [])):
            g["lr"] = lr


def build_ACT_model_and_optimizer(args_override: dict):
    """detr/main.py:92-112 without the ``sys.argv`` parse and without ``.cuda()`` side effects beyond device memory."""
    args = dict(_DEFAULTS)
    args.update(args_override)
    _check_supported(args)
    model = DETRVAE(args)
    optimizer = ActAdamW(model, args["lr"], args["lr_backbone"], args["weight_decay"])
    return model, optimizer


def kl_divergence(mu, logvar):
    """policy.py:378-391 (kept for callers that import it; the engine computes the same numbers in its loss kernel)."""
    klds = -0.5 * (1 + logvar - mu.pow(2) - logvar.exp())
    return klds.sum(1).mean(0, True), klds.mean(0), klds.mean(1).mean(0, True)


class ACTPolicy(nn.Module):
    """policy.py:243-348.  ``policy(qpos, image, actions, is_pad)`` -> {'l1','kl','loss'} (loss supports .backward());
    ``policy(qpos, image)`` -> a_hat [B, num_queries, action_dim]."""

    def __init__(self, args_override: dict):
        super().__init__()
        self.use_depth = args_override.get("use_depth", False)
        self.use_pcd = args_override.get("use_pcd", False)
        self.depth_camera_names = args_override.get("depth_camera_names", [])
        model, optimizer = build_ACT_model_and_optimizer(args_override)
        self.model = model
        self.optimizer = optimizer
        self.kl_weight = args_override["kl_weight"]
        self.vq = args_override.get("vq", False)
        print(f"KL Weight {self.kl_weight}")

    def __call__(self, qpos, image, actions=None, is_pad=None, vq_sample=None, depth_img=None, pointcloud=None):
        if vq_sample is not None or depth_img is not None or pointcloud is not None:
            raise NotImplementedError("vq / depth / pointcloud inputs are outside the B200 hot path")
        m = self.model
        if actions is not None:                       # training / validation call (policy.py:288-320)
            nq = m.num_queries
            actions, is_pad = actions[:, :nq], is_pad[:, :nq]
            eps = torch.randn((qpos.shape[0], m.latent_dim), device=qpos.device, dtype=torch.float32)
            if torch.is_grad_enabled() and not torch.is_inference_mode_enabled():
                losses = _ActStep.apply(m._anchor, m, qpos, image, actions, is_pad, eps)
            else:
                losses = m._engine_for(qpos, image, True).forward(qpos, image, actions, is_pad, eps)["losses"]
            return dict(l1=losses[0], kl=losses[1], loss=losses[2])
        return m._engine_for(qpos, image, False).forward(qpos, image)["a_hat"]

    def configure_optimizers(self):
        return self.optimizer

    def set_data_parallel(self, process_group=None):
        """Gradient all-reduce (NCCL, averaged) overlapped with the staged backward.  Not a reference feature
        (the reference is single-GPU); see DESIGN.md."""
        import torch.distributed as dist
        self.model._dp_group = process_group if process_group is not None else dist.group.WORLD

    def serialize(self):
        return self.state_dict()

    def deserialize(self, model_dict):
        return self.load_state_dict(model_dict)

    def state_dict(self, *args, **kw):
        return self.model.state_dict(prefix="model.")

    def load_state_dict(self, state_dict, strict: bool = True):
        return self.model.load_state_dict(state_dict, strict=strict, prefix="model.")

    def cuda(self, device=None):
        return self

    def train(self, mode: bool = True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)
