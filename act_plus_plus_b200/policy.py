"""Drop-in mirror of the reference's ``policy.ACTPolicy`` / ``detr.main.build_ACT_model_and_optimizer``.

Same constructor dict, call signatures, return types, attribute names and error behaviour as
``/root/reference/policy.py:243-348`` and ``/root/reference/detr/main.py:92-112``; the work underneath is done by
the CUDA engine (``act_engine_*`` C ABI).  There is no PyTorch/CPU fallback: without the built library or a CUDA
device construction raises.

Behaviour notes (all in DESIGN.md):
  * ``policy.train()`` enables the reference's dropout (``args_override['dropout']``, default 0.1 as detr/main.py:45)
    on training calls, ``policy.eval()`` disables it; masks are counter-based Philox draws (csrc/dropout.cuh), so the
    random stream differs from torch's -- parity is defined on the deterministic path (dropout 0 / ``.eval()``) and on
    injected masks (tests/test_dropout.py);
  * ``state_dim`` is a config key (default 14, the ALOHA qpos size) instead of the fork's hard-coded 7
    (detr_vae.py:345); ``depth_camera_names`` may be absent (policy.py:249 raises KeyError upstream);
  * ``sys.argv`` is never parsed (detr/main.py:93-94 does);
  * backbone weights: the reference starts from torchvision's ImageNet ResNet-18 (backbone.py:121-124, downloaded);
    there is no network here, so ``args_override['backbone_ckpt']`` (a torchvision ``resnet18`` state_dict file) or a
    cached ``resnet18-*.pth`` in the torch hub cache is loaded when present, otherwise construction WARNS and the trunk
    keeps its random Kaiming init with identity FrozenBatchNorm statistics;
  * ``args_override['fused_step']=True`` (or ACT_FUSED_STEP=1) runs ``loss.backward()`` + ``optimizer.step()`` as the
    engine's single overlapped call at ``optimizer.step()`` time and replays the step from two CUDA graphs; gradients
    are then only materialised by ``optimizer.step()`` / ``model.flush_backward()``.
"""
from __future__ import annotations

import math
from collections import OrderedDict
import glob
import os
import warnings
from typing import Dict, Optional

import torch
import torch.nn as nn

from .engine import (KIND_BACKBONE, KIND_BUFFER, KIND_NOGRAD, R_BACKBONE, R_CVAE, R_MAIN, ActEngine)

_DEFAULTS = dict(  # detr/main.py:12-60 argparse defaults that ACT never overrides, plus the overridden ones
    lr=1e-4, lr_backbone=1e-5, weight_decay=1e-4, backbone="resnet18", dilation=False, position_embedding="sine",
    camera_names=[], enc_layers=4, dec_layers=6, dim_feedforward=2048, hidden_dim=256, dropout=0.1, nheads=8,
    num_queries=400, pre_norm=False, masks=False, use_pcd=False, use_depth=False, vq=False, vq_class=None, vq_dim=None,
    action_dim=16, no_encoder=False, state_dim=14, kl_weight=10, backbone_ckpt=None, fused_step=None, cuda_graph=True,
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


def _find_resnet18_ckpt(args: dict) -> Optional[str]:
    if args.get("backbone_ckpt"):
        return args["backbone_ckpt"]
    hub = os.path.join(os.environ.get("TORCH_HOME", os.path.join(os.path.expanduser("~"), ".cache", "torch")), "hub", "checkpoints")
    hits = sorted(glob.glob(os.path.join(hub, "resnet18-*.pth")))
    return hits[0] if hits else None


def _load_pretrained_backbones(eng: ActEngine, args: dict) -> bool:
    """backbone.py:121-124 builds torchvision resnet18(pretrained=is_main_process()) and freezes the BatchNorm statistics:
    copy conv / bn weight, bias, running_mean, running_var of a torchvision resnet18 state_dict into every camera's
    ``backbones.<k>.0.body.*``.  Returns False (after a loud warning) when no checkpoint is available offline."""
    path = _find_resnet18_ckpt(args)
    if path is None:
        warnings.warn("act_plus_plus_b200: no ImageNet ResNet-18 checkpoint found (args_override['backbone_ckpt'] or the torch "
                      "hub cache); the camera backbones start from RANDOM weights with identity FrozenBatchNorm statistics, "
                      "unlike the reference (backbone.py:121-124). Load a checkpoint before comparing task results.",
                      RuntimeWarning, stacklevel=3)
        return False
    sd = torch.load(path, map_location="cpu")
    n = 0
    with torch.no_grad():
        for name in eng.tensors:
            if not name.startswith("backbones."):
                continue
            key = name.split(".0.body.", 1)[1]
            if key not in sd:
                raise RuntimeError(f"{path}: missing resnet18 entry {key}")
            eng.view(name).copy_(sd[key].to(eng.device, torch.float32))
            n += 1
    eng.mark_params_changed()
    eng.refresh()
    return n > 0


class _ActStep(torch.autograd.Function):
    """Bridges ``loss.backward()`` (imitate_episodes.py:605-606) to ``act_engine_backward``.  The incoming gradient of
    the three returned scalars stays on the device (``act_engine_set_loss_grad``): no host synchronisation."""

    @staticmethod
    def forward(ctx, anchor, model, eng, qpos, image, actions, is_pad, eps):
        losses = model._train_forward(eng, qpos, image, actions, is_pad, eps)
        ctx.model, ctx.eng = model, eng
        return losses

    @staticmethod
    def backward(ctx, g):
        model, eng = ctx.model, ctx.eng
        model._pending = (eng, g.detach().to(torch.float32).contiguous())
        if not model.fused_step:
            model.flush_backward()
        return torch.zeros(1, device=model.device), None, None, None, None, None, None, None


class _StepGraphs:
    """fused_step mode: the training step of one engine at fixed shapes as two CUDA graphs -- F = zero_grad + forward +
    loss (replayed by ``policy(...)``), B = backward + AdamW with every region's optimiser overlapped
    (replayed by ``optimizer.step()``).  Inputs are copied into static buffers, the loss gradient is read from a static
    device vector, the optimiser step and the dropout step advance through device counters."""

    def __init__(self, model, eng, qpos, image, actions, is_pad):
        self.eng = eng
        self.key = (tuple(qpos.shape), tuple(image.shape), image.dtype, tuple(actions.shape), tuple(is_pad.shape))
        dev = eng.device
        self.qpos = torch.empty_like(qpos, dtype=torch.float32).contiguous()
        self.image = torch.empty_like(image).contiguous()
        self.actions = torch.empty(actions.shape, dtype=torch.float32, device=dev)
        self.is_pad = torch.empty(is_pad.shape, dtype=torch.bool, device=dev)
        self.eps = torch.empty((qpos.shape[0], model.latent_dim), dtype=torch.float32, device=dev)
        self.gvec = torch.zeros(3, dtype=torch.float32, device=dev)
        self.fwd = self.bwd = None
        self.losses = None

    def load(self, qpos, image, actions, is_pad, eps):
        self.qpos.copy_(qpos, non_blocking=True); self.image.copy_(image, non_blocking=True)
        self.actions.copy_(actions, non_blocking=True); self.is_pad.copy_(is_pad, non_blocking=True)
        self.eps.copy_(eps, non_blocking=True)

    def _fwd_body(self):
        self.eng.zero_grad(overlap_forward=True)
        out = self.eng.forward(self.qpos, self.image, self.actions, self.is_pad, self.eps)
        self.eng.join_lanes()
        return out["losses"]

    def capture_forward(self):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._fwd_body()                              # warm-up on a side stream (lazy one-time host set-up)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.losses = self._fwd_body()
        self.fwd = g

    def capture_backward(self, step_fn):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            step_fn(self.eng, advance_on_device=True)
        self.bwd = g


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
        self._dp_seed_offset = 0
        self._pending = None             # (engine, device gradient of {l1, kl, loss}) of a loss.backward() not yet run
        self._graphs: Dict[int, _StepGraphs] = {}
        self._seen: Dict[tuple, int] = {}
        fs = args.get("fused_step")
        self.fused_step = bool(fs) if fs is not None else os.environ.get("ACT_FUSED_STEP", "0") == "1"
        self.use_graphs = bool(args.get("cuda_graph", True)) and os.environ.get("ACT_POLICY_GRAPH", "1") == "1"
        self.dropout_p = float(args.get("dropout", 0.1))
        self._drop_seed = int(args["seed"]) if args.get("seed") is not None else (torch.initial_seed() & 0x7FFFFFFF)
        self.training_mode = True        # nn.Module default: constructed in train mode
        # a minimal engine fixes the parameter layout and owns the master tensors; compute engines share them
        self._master = self._make_engine(1, 64, 64, train=True)
        # policy.py:306 only tests `self.model.encoder is None` (the no_encoder variant, detr_vae.py:373-377)
        self.encoder = None if args.get("no_encoder") else object()
        _init_like_reference(self._master, args.get("seed"))
        self.pretrained_backbone = _load_pretrained_backbones(self._master, args)

    # -------------------------------------------------------------- engines
    def _make_engine(self, batch, h, w, train, share: Optional[ActEngine] = None) -> ActEngine:
        a = self.args
        eng = ActEngine(batch_max=batch, num_cameras=len(self.camera_names), img_h=h, img_w=w,
                        hidden_dim=a["hidden_dim"], nheads=a["nheads"], dim_feedforward=a["dim_feedforward"],
                        enc_layers=a["enc_layers"], dec_layers=a["dec_layers"], num_queries=a["num_queries"],
                        state_dim=a["state_dim"], action_dim=a["action_dim"], latent_dim=32, kl_weight=float(a["kl_weight"]),
                        lr=a["lr"], lr_backbone=a["lr_backbone"], weight_decay=a["weight_decay"], train=train,
                        device=self.device, no_encoder=bool(a.get("no_encoder"))) if share is None else \
            ActEngine.sharing(share, batch, h, w, train)
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
            if eng is not None:
                # a larger batch needs a larger workspace: say so (each engine holds its own activation buffers) and
                # drop the smaller one instead of keeping both alive
                warnings.warn(f"act_plus_plus_b200: re-allocating the {H}x{W} {'train' if train else 'inference'} workspace "
                              f"for batch {B} (was {eng.cfg.batch_max})", RuntimeWarning, stacklevel=3)
                if self._pending is not None and self._pending[0] is eng:
                    self.flush_backward()
                self._graphs.pop(id(eng), None)
            eng = self._make_engine(B, H, W, train, share=self._master)
            if self._dp_group is not None and train:
                self._connect_dp(eng)
            self._engines[key] = eng
        return eng

    def _apply_mode(self, eng: ActEngine):
        p = self.dropout_p if (self.training_mode and eng.train_enabled) else 0.0
        if getattr(eng, "dropout_p", 0.0) != p or getattr(eng, "_drop_seed", None) != self._drop_seed + self._dp_seed_offset:
            eng.set_dropout(p, self._drop_seed + self._dp_seed_offset)
            eng._drop_seed = self._drop_seed + self._dp_seed_offset
            self._graphs.pop(id(eng), None)              # masks on/off is baked into a captured graph

    # -------------------------------------------------------------- training step plumbing
    def _train_forward(self, eng: ActEngine, qpos, image, actions, is_pad, eps) -> torch.Tensor:
        """zero_grad (fused mode) + forward + loss of one training call; replays graph F when the shapes repeat."""
        if self._pending is not None:
            self.flush_backward()                         # a backward that never reached optimizer.step(): run it first
        self._apply_mode(eng)
        if not (self.fused_step and self.use_graphs):
            return eng.forward(qpos, image, actions, is_pad, eps)["losses"].clone()
        key = (tuple(qpos.shape), tuple(image.shape), image.dtype, tuple(actions.shape), tuple(is_pad.shape))
        sg = self._graphs.get(id(eng))
        if sg is not None and sg.key != key:
            sg = None
            self._graphs.pop(id(eng), None)
        if sg is None:
            n = self._seen[(id(eng),) + key] = self._seen.get((id(eng),) + key, 0) + 1
            if n < 3:                                     # the first two steps at a shape run eagerly
                eng.zero_grad(overlap_forward=True)
                return eng.forward(qpos, image, actions, is_pad, eps)["losses"].clone()
            sg = _StepGraphs(self, eng, qpos, image, actions, is_pad)
            sg.load(qpos, image, actions, is_pad, eps)
            try:
                sg.capture_forward()
            except Exception as ex:                       # capture is an optimisation, never a requirement
                warnings.warn(f"act_plus_plus_b200: CUDA-graph capture of the forward failed ({ex}); launching directly")
                self.use_graphs = False
                eng.zero_grad(overlap_forward=True)
                return eng.forward(qpos, image, actions, is_pad, eps)["losses"].clone()
            self._graphs[id(eng)] = sg
        sg.load(qpos, image, actions, is_pad, eps)
        eng.last_was_train = True
        sg.fwd.replay()
        return sg.losses.clone()

    def _step_fn(self, eng: ActEngine, advance_on_device: bool = False):
        """backward + AdamW as one overlapped engine call (single GPU, peer-memory or NCCL data parallel)."""
        if self._dp_group is None:
            eng.backward_and_step(1.0, advance_on_device=advance_on_device)
        elif eng.dp_connected:
            eng.backward_and_step_p2p(1.0, advance_on_device=advance_on_device)
        else:
            from .train import backward_and_step_hooked
            backward_and_step_hooked(eng, self._dp_group, 1.0)

    def flush_backward(self):
        """Run a pending ``loss.backward()`` now (gradients into the flat buffer; no optimiser step)."""
        if self._pending is None:
            return
        eng, g = self._pending
        self._pending = None
        eng.set_loss_grad(g)
        try:
            if self._dp_group is None:
                eng.backward(0, 1.0)
            else:
                from .train import backward_data_parallel
                backward_data_parallel(eng, self._dp_group, 1.0)
        finally:
            eng.set_loss_grad(None)

    def optimizer_step(self, lr: float, lr_backbone: float):
        """``optimizer.step()``: with a pending backward (fused_step) the engine's overlapped backward_and_step, otherwise
        the fused AdamW over gradients that are already in the flat buffer."""
        if self._pending is None:
            eng = next((e for e in self._engines.values() if e.train_enabled and e.last_was_train), self._master)
            eng.set_lr(lr, lr_backbone)
            eng.adamw_step()
            return
        eng, g = self._pending
        self._pending = None
        eng.set_lr(lr, lr_backbone)
        sg = self._graphs.get(id(eng))
        # (the NCCL-hook exchange counts its optimiser step on the host: it is launched directly, not replayed)
        if sg is not None and self.use_graphs and (self._dp_group is None or eng.dp_connected):
            sg.gvec.copy_(g, non_blocking=True)
            if sg.bwd is None or getattr(sg, "lr", None) != (lr, lr_backbone):
                eng.set_loss_grad(sg.gvec)
                try:
                    sg.capture_backward(self._step_fn)    # learning rates are baked into the graph: re-capture on change
                    sg.lr = (lr, lr_backbone)
                except Exception as ex:
                    warnings.warn(f"act_plus_plus_b200: CUDA-graph capture of backward+step failed ({ex}); launching directly")
                    self.use_graphs = False
                    sg.bwd = None
                finally:
                    eng.set_loss_grad(None)               # the graph keeps reading sg.gvec; eager calls use their own
            if sg.bwd is not None:
                sg.bwd.replay()
                eng.note_graph_step()
                return
        eng.set_loss_grad(g)
        try:
            self._step_fn(eng)
        finally:
            eng.set_loss_grad(None)
        if eng.dp_connected and eng.step_count % 50 == 1 and eng.dp_status() != 0:
            raise RuntimeError("peer-memory data-parallel step: a flag barrier timed out (a peer never arrived); "
                               "parameters were left untouched from that barrier on")

    def _connect_dp(self, eng: ActEngine):
        import torch.distributed as dist
        if os.environ.get("ACT_DP_P2P", "1") != "1" or dist.get_backend(self._dp_group) != "nccl":
            return
        try:
            eng.dp_connect(self._dp_group)                # collective; raises on every rank if any rank failed
        except RuntimeError as ex:
            warnings.warn(f"act_plus_plus_b200: peer-memory gradient exchange unavailable ({str(ex)[:120]}); using NCCL")

    # -------------------------------------------------------------- nn.Module surface
    def forward(self, qpos, image, env_state, actions=None, is_pad=None, vq_sample=None, pointcloud=None, depth_img=None):
        """Like everything in this package the engine fuses the ImageNet normalisation of policy.py:268-272 into the
        stem, so DETRVAE.forward takes the *un-normalised* [0,1] image and ACTPolicy does not normalise separately."""
        if pointcloud is not None or depth_img is not None or vq_sample is not None:
            raise NotImplementedError("pointcloud / depth / vq inputs are outside the B200 hot path")
        if actions is not None:
            eng = self._engine_for(qpos, image, True)
            if self._pending is not None:
                self.flush_backward()
            self._apply_mode(eng)
            eps = torch.randn((qpos.shape[0], self.latent_dim), device=qpos.device, dtype=torch.float32)
            out = eng.forward(qpos, image, actions, is_pad, eps, want_is_pad_hat=True)
            if self.encoder is None:                      # detr_vae.py:109-112: mu = logvar = None without the CVAE encoder
                return out["a_hat"], out["is_pad_hat"], [None, None], None, None
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

    def train(self, mode: bool = True):
        self.training_mode = bool(mode)
        return super().train(mode)

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

    def zero_grad(self, set_to_none: bool = True):
        m = self.model
        m._pending = None                                 # a backward that was never stepped is discarded with its gradients
        if m.fused_step:
            return                                        # the clear is part of the next training forward (overlapped)
        m._master.zero_grad()

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.model.optimizer_step(float(self.param_groups[0]["lr"]), float(self.param_groups[1]["lr"]))
        return loss

    def state_dict(self):
        m = self.model._master
        self.model.flush_backward()
        if any(e.dp_connected for e in self.model._engines.values()):
            for e in self.model._engines.values():
                if e.dp_connected:
                    if e.dp_status() != 0:
                        raise RuntimeError("peer-memory data-parallel step: a flag barrier timed out; refusing to checkpoint")
                    e.dp_gather_moments(self.model._dp_group)      # moments are sharded by the peer-memory step
        return dict(step=m.step_count, exp_avg=m.exp_avg.clone(), exp_avg_sq=m.exp_avg_sq.clone(),
                    lr=[g["lr"] for g in self.param_groups])

    def load_state_dict(self, sd):
        m = self.model._master
        m._shared["step"] = int(sd["step"])
        m.exp_avg.copy_(sd["exp_avg"]); m.exp_avg_sq.copy_(sd["exp_avg_sq"])
        for g, lr in zip(self.param_groups, sd.get("lr", [])):
            g["lr"] = lr
        self.model._graphs.clear()                        # the device step counter is re-seeded by the next eager steps
        self.model._seen.clear()


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
            eng = m._engine_for(qpos, image, True)
            if torch.is_grad_enabled() and not torch.is_inference_mode_enabled():
                losses = _ActStep.apply(m._anchor, m, eng, qpos, image, actions, is_pad, eps)
            else:
                if m._pending is not None:
                    m.flush_backward()
                m._apply_mode(eng)
                losses = eng.forward(qpos, image, actions, is_pad, eps)["losses"]
            return dict(l1=losses[0], kl=losses[1], loss=losses[2])
        eng = m._engine_for(qpos, image, False)
        if m.use_graphs and not torch.is_grad_enabled():
            return eng.infer_graphed(qpos, image, tag="policy").clone()    # the graph's output buffer is reused by the next call
        return eng.forward(qpos, image)["a_hat"]

    def configure_optimizers(self):
        return self.optimizer

    def set_data_parallel(self, process_group=None, rank_seed_offset: Optional[int] = None):
        """Data-parallel training (new work: the reference is single-GPU): every rank holds a replica and
        feeds its own batch shard; ``optimizer.step()`` averages the gradients -- fused into the optimiser kernel over
        NVLink peer memory when the ranks can map each other's buffers (csrc/dp_p2p.cu), through NCCL all-reduces
        overlapped with the backward otherwise.  Dropout seeds are offset by rank."""
        import torch.distributed as dist
        m = self.model
        m._dp_group = process_group if process_group is not None else dist.group.WORLD
        m._dp_seed_offset = (dist.get_rank(m._dp_group) if rank_seed_offset is None else rank_seed_offset) * 1000003
        for eng in m._engines.values():
            if eng.train_enabled:
                m._connect_dp(eng)
        m._graphs.clear()

    def serialize(self):
        return self.state_dict()

    def deserialize(self, model_dict):
        return self.load_state_dict(model_dict)

    def state_dict(self, *args, **kw):
        return self.model.state_dict(prefix="model.")

    def load_state_dict(self, state_dict, strict: bool = True):
        return self.model.load_state_dict(state_dict, strict=strict, prefix="model.")

    def named_parameters(self, prefix: str = "", recurse: bool = True, remove_duplicate: bool = True):
        # nn.Module would walk `_parameters` and find only the autograd anchor: delegate to the engine-backed views
        return self.model.named_parameters(prefix=(prefix + "." if prefix else "") + "model")

    def parameters(self, recurse: bool = True):
        return self.model.parameters()

    def cuda(self, device=None):
        return self

    def train(self, mode: bool = True):
        """train(): the reference's dropout acts on training calls; eval(): deterministic (imitate_episodes.py:566)."""
        self.training = mode
        self.model.train(mode)
        return self

    def eval(self):
        return self.train(False)
