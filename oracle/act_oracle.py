"""CPU restatement (torch fp32, CPU tensors only) of the reference ACT hot path.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  Every function cites the
reference file:line (relative to /root/reference) whose arithmetic it restates.
The restatement is *functional*: it takes a ``state_dict`` with the reference's
own key names (344 tensors for one camera) and never
builds ``nn.Module`` objects, so that the same dictionary can be loaded into the
unmodified reference (``oracle/make_golden.py``) and into the CUDA engine.

Parity: pinned against the reference run on CPU in the build container
(``oracle/make_golden.py``); the reference itself has no golden vectors.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

IMAGENET_MEAN = (0.485, 0.456, 0.406)   # policy.py:268-271
IMAGENET_STD = (0.229, 0.224, 0.225)

DEFAULT_CFG = dict(
    hidden_dim=512, nheads=8, dim_feedforward=3200, enc_layers=4, dec_layers=7,
    num_queries=100, state_dim=14, action_dim=16, num_cameras=1, latent_dim=32,
    kl_weight=10, lr=1e-5, lr_backbone=1e-5, weight_decay=1e-4,
    no_encoder=False,       # detr_vae.py:373-377: encoder = None -> latent 0, kl 0, no `encoder.*` tensors
)

# torchvision resnet18 topology (backbone.py:121-124 builds it; layer4 is returned, backbone.py:71)
_RESNET_STAGES = [("layer1", 64, 64, 1), ("layer2", 64, 128, 2), ("layer3", 128, 256, 2), ("layer4", 256, 512, 2)]


def make_cfg(**kw) -> dict:
    cfg = dict(DEFAULT_CFG)
    cfg.update(kw)
    return cfg


# --------------------------------------------------------------------------------------
# state_dict specification (names / shapes / init kind), in the reference's own order
# --------------------------------------------------------------------------------------
def _enc_layer_spec(prefix: str, d: int, ff: int, xavier: bool) -> List[tuple]:
    k = "xavier" if xavier else "linear"
    return [
        (prefix + "self_attn.in_proj_weight", (3 * d, d), "xavier"),     # nn.MultiheadAttention always xavier-inits in_proj
        (prefix + "self_attn.in_proj_bias", (3 * d,), "zeros"),
        (prefix + "self_attn.out_proj.weight", (d, d), k),
        (prefix + "self_attn.out_proj.bias", (d,), "zeros"),
        (prefix + "linear1.weight", (ff, d), k),
        (prefix + "linear1.bias", (ff,), "bias:%d" % d),
        (prefix + "linear2.weight", (d, ff), k),
        (prefix + "linear2.bias", (d,), "bias:%d" % ff),
        (prefix + "norm1.weight", (d,), "ln_w"),
        (prefix + "norm1.bias", (d,), "ln_b"),
        (prefix + "norm2.weight", (d,), "ln_w"),
        (prefix + "norm2.bias", (d,), "ln_b"),
    ]


def _dec_layer_spec(prefix: str, d: int, ff: int) -> List[tuple]:
    return [
        (prefix + "self_attn.in_proj_weight", (3 * d, d), "xavier"),
        (prefix + "self_attn.in_proj_bias", (3 * d,), "zeros"),
        (prefix + "self_attn.out_proj.weight", (d, d), "xavier"),
        (prefix + "self_attn.out_proj.bias", (d,), "zeros"),
        (prefix + "multihead_attn.in_proj_weight", (3 * d, d), "xavier"),
        (prefix + "multihead_attn.in_proj_bias", (3 * d,), "zeros"),
        (prefix + "multihead_attn.out_proj.weight", (d, d), "xavier"),
        (prefix + "multihead_attn.out_proj.bias", (d,), "zeros"),
        (prefix + "linear1.weight", (ff, d), "xavier"),
        (prefix + "linear1.bias", (ff,), "bias:%d" % d),
        (prefix + "linear2.weight", (d, ff), "xavier"),
        (prefix + "linear2.bias", (d,), "bias:%d" % ff),
        (prefix + "norm1.weight", (d,), "ln_w"), (prefix + "norm1.bias", (d,), "ln_b"),
        (prefix + "norm2.weight", (d,), "ln_w"), (prefix + "norm2.bias", (d,), "ln_b"),
        (prefix + "norm3.weight", (d,), "ln_w"), (prefix + "norm3.bias", (d,), "ln_b"),
    ]


def _bn_spec(prefix: str, c: int) -> List[tuple]:
    # FrozenBatchNorm2d buffers (backbone.py:29-34); not parameters, never trained
    return [(prefix + "weight", (c,), "bn_w"), (prefix + "bias", (c,), "bn_b"),
            (prefix + "running_mean", (c,), "bn_m"), (prefix + "running_var", (c,), "bn_v")]


def _backbone_spec(prefix: str) -> List[tuple]:
    s = [(prefix + "conv1.weight", (64, 3, 7, 7), "conv")] + _bn_spec(prefix + "bn1.", 64)
    for name, cin, cout, stride in _RESNET_STAGES:
        for blk in range(2):
            p = f"{prefix}{name}.{blk}."
            c_in = cin if blk == 0 else cout
            s.append((p + "conv1.weight", (cout, c_in, 3, 3), "conv"))
            s += _bn_spec(p + "bn1.", cout)
            s.append((p + "conv2.weight", (cout, cout, 3, 3), "conv"))
            s += _bn_spec(p + "bn2.", cout)
            if blk == 0 and (stride != 1 or c_in != cout):
                s.append((p + "downsample.0.weight", (cout, c_in, 1, 1), "conv"))
                s += _bn_spec(p + "downsample.1.", cout)
    return s


def state_spec(cfg: dict) -> List[tuple]:
    """(name, shape, init-kind) for every state_dict entry, in the reference's order
    (detr_vae.py:38-105 registration order; verified against the imported reference)."""
    d, ff, nq = cfg["hidden_dim"], cfg["dim_feedforward"], cfg["num_queries"]
    S, A, Z = cfg["state_dim"], cfg["action_dim"], cfg["latent_dim"]
    spec: List[tuple] = [("pos_table", (1, nq + 2, d), "pos_table")]
    for i in range(cfg["enc_layers"]):
        spec += _enc_layer_spec(f"transformer.encoder.layers.{i}.", d, ff, True)
    for i in range(cfg["dec_layers"]):
        spec += _dec_layer_spec(f"transformer.decoder.layers.{i}.", d, ff)
    spec += [("transformer.decoder.norm.weight", (d,), "ln_w"), ("transformer.decoder.norm.bias", (d,), "ln_b")]
    if not cfg.get("no_encoder"):
        for i in range(cfg["enc_layers"]):
            spec += _enc_layer_spec(f"encoder.layers.{i}.", d, ff, False)
    spec += [
        ("action_head.weight", (A, d), "linear"), ("action_head.bias", (A,), "bias:%d" % d),
        ("is_pad_head.weight", (1, d), "linear"), ("is_pad_head.bias", (1,), "bias:%d" % d),
        ("query_embed.weight", (nq, d), "normal"),
        ("input_proj.weight", (d, 512, 1, 1), "conv1x1"), ("input_proj.bias", (d,), "bias:512"),
    ]
    for k in range(cfg["num_cameras"]):
        spec += _backbone_spec(f"backbones.{k}.0.body.")
    spec += [
        ("input_proj_robot_state.weight", (d, S), "linear"), ("input_proj_robot_state.bias", (d,), "bias:%d" % S),
        ("cls_embed.weight", (1, d), "normal"),
        ("encoder_action_proj.weight", (d, A), "linear"), ("encoder_action_proj.bias", (d,), "bias:%d" % A),
        ("encoder_joint_proj.weight", (d, S), "linear"), ("encoder_joint_proj.bias", (d,), "bias:%d" % S),
        ("latent_proj.weight", (2 * Z, d), "linear"), ("latent_proj.bias", (2 * Z,), "bias:%d" % d),
        ("latent_out_proj.weight", (d, Z), "linear"), ("latent_out_proj.bias", (d,), "bias:%d" % Z),
        ("additional_pos_embed.weight", (2, d), "normal"),
    ]
    return spec


_BUFFER_KINDS = ("pos_table", "bn_w", "bn_b", "bn_m", "bn_v")


def is_buffer(kind: str) -> bool:
    return kind in _BUFFER_KINDS


def get_sinusoid_encoding_table(n_position: int, d_hid: int) -> torch.Tensor:
    """detr_vae.py:25-33 -- float64 numpy table, sin on even / cos on odd columns, cast to float32."""
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    j = np.arange(d_hid)[None, :]
    table = pos / np.power(10000, 2 * (j // 2) / d_hid)
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.FloatTensor(table).unsqueeze(0)


def init_state_dict(cfg: dict, seed: int = 0, perturb: bool = True) -> Dict[str, torch.Tensor]:
    """Seeded random-init state_dict with the reference's names, shapes and init families
    (xavier_uniform on transformer.* matrices, transformer.py:44-47; torch defaults elsewhere;
    kaiming-normal fan_out for torchvision resnet convs).  With ``perturb`` the values torch
    would initialise to constants (biases of MHA, LayerNorm affine, FrozenBN buffers) get
    small random values so that parity tests exercise every term.  Deterministic for a given
    torch build: a single CPU ``torch.Generator`` walked in ``state_spec`` order."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    for name, shape, kind in state_spec(cfg):
        if kind == "pos_table":
            t = get_sinusoid_encoding_table(shape[1], shape[2])
        elif kind == "xavier":
            fan_out, fan_in = shape[0], shape[1]
            a = math.sqrt(6.0 / (fan_in + fan_out))
            t = (torch.rand(shape, generator=g) * 2 - 1) * a
        elif kind in ("linear", "conv1x1"):
            fan_in = shape[1]
            a = 1.0 / math.sqrt(fan_in)          # kaiming_uniform(a=sqrt(5)) bound
            t = (torch.rand(shape, generator=g) * 2 - 1) * a
        elif kind.startswith("bias:"):
            a = 1.0 / math.sqrt(int(kind.split(":")[1]))
            t = (torch.rand(shape, generator=g) * 2 - 1) * a
        elif kind == "conv":
            fan_out = shape[0] * shape[2] * shape[3]
            t = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_out)
        elif kind == "normal":
            t = torch.randn(shape, generator=g)
        elif kind == "zeros":
            t = torch.randn(shape, generator=g) * 0.02 if perturb else torch.zeros(shape)
        elif kind == "ln_w":
            t = 1.0 + (torch.randn(shape, generator=g) * 0.05 if perturb else 0.0) * torch.ones(shape)
        elif kind == "ln_b":
            t = torch.randn(shape, generator=g) * 0.05 if perturb else torch.zeros(shape)
        elif kind == "bn_w":
            t = 0.6 + 0.3 * torch.rand(shape, generator=g) if perturb else torch.ones(shape)
        elif kind == "bn_b":
            t = torch.randn(shape, generator=g) * 0.05 if perturb else torch.zeros(shape)
        elif kind == "bn_m":
            t = torch.randn(shape, generator=g) * 0.05 if perturb else torch.zeros(shape)
        elif kind == "bn_v":
            t = 0.8 + 0.4 * torch.rand(shape, generator=g) if perturb else torch.ones(shape)
        else:
            raise ValueError(kind)
        sd[name] = t.to(torch.float32).contiguous()
    return sd


def state_checksum(sd: Dict[str, torch.Tensor]) -> np.ndarray:
    """Small fingerprint used by the golden fixtures to prove that both sides regenerated
    the same seeded weights."""
    names = ["transformer.encoder.layers.0.linear1.weight", "encoder.layers.3.self_attn.in_proj_weight",
             "backbones.0.0.body.layer4.1.conv2.weight", "action_head.weight", "query_embed.weight"]
    out = []
    for n in names:
        if n not in sd:                          # no_encoder variant
            n = "latent_proj.weight"
        t = sd[n].double().flatten()
        out += [float(t.sum()), float(t.abs().sum()), float(t[::997].sum())]
    return np.asarray(out, dtype=np.float64)


def make_inputs(cfg: dict, batch: int, height: int = 480, width: int = 640, seed: int = 0,
                extra_steps: int = 0) -> Dict[str, torch.Tensor]:
    """Synthetic ALOHA-shaped batch: image U[0,1) f32 [B,K,3,H,W], qpos N(0,1),
    actions N(0,1) [B,nq(+extra),A], is_pad with a random tail (utils.py:118-124), eps N(0,1) [B,32]."""
    g = torch.Generator(device="cpu")
    g.manual_seed(1000 + seed)
    K, nq = cfg["num_cameras"], cfg["num_queries"] + extra_steps
    image = torch.rand((batch, K, 3, height, width), generator=g)
    qpos = torch.randn((batch, cfg["state_dim"]), generator=g)
    actions = torch.randn((batch, nq, cfg["action_dim"]), generator=g)
    n_valid = torch.randint(1, cfg["num_queries"] + 1, (batch,), generator=g)
    is_pad = torch.arange(nq)[None, :] >= n_valid[:, None]
    eps = torch.randn((batch, cfg["latent_dim"]), generator=g)
    return dict(image=image, qpos=qpos, actions=actions, is_pad=is_pad, eps=eps)


# --------------------------------------------------------------------------------------
# model pieces
# --------------------------------------------------------------------------------------
def frozen_bn(sd, p: str, x: torch.Tensor) -> torch.Tensor:
    """backbone.py:47-57 (eps = 1e-5 inside the rsqrt)."""
    w, b = sd[p + "weight"].reshape(1, -1, 1, 1), sd[p + "bias"].reshape(1, -1, 1, 1)
    rv, rm = sd[p + "running_var"].reshape(1, -1, 1, 1), sd[p + "running_mean"].reshape(1, -1, 1, 1)
    scale = w * (rv + 1e-5).rsqrt()
    bias = b - rm * scale
    return x * scale + bias


def resnet18_layer4(sd, p: str, x: torch.Tensor) -> torch.Tensor:
    """torchvision resnet18 up to layer4 with FrozenBatchNorm2d (backbone.py:71,121-124)."""
    x = F.conv2d(x, sd[p + "conv1.weight"], stride=2, padding=3)
    x = F.relu(frozen_bn(sd, p + "bn1.", x))
    x = F.max_pool2d(x, kernel_size=3, stride=2, padding=1)
    for name, cin, cout, stride in _RESNET_STAGES:
        for blk in range(2):
            q = f"{p}{name}.{blk}."
            s = stride if blk == 0 else 1
            idt = x
            out = F.conv2d(x, sd[q + "conv1.weight"], stride=s, padding=1)
            out = F.relu(frozen_bn(sd, q + "bn1.", out))
            out = F.conv2d(out, sd[q + "conv2.weight"], stride=1, padding=1)
            out = frozen_bn(sd, q + "bn2.", out)
            if (q + "downsample.0.weight") in sd:
                idt = frozen_bn(sd, q + "downsample.1.", F.conv2d(x, sd[q + "downsample.0.weight"], stride=s))
            x = F.relu(out + idt)
    return x


def position_embedding_sine(h: int, w: int, num_pos_feats: int, temperature: float = 10000.0) -> torch.Tensor:
    """position_encoding.py:30-52 with normalize=True, scale=2*pi -> [1, 2*num_pos_feats, h, w]."""
    not_mask = torch.ones((1, h, w), dtype=torch.float32)
    y_embed = not_mask.cumsum(1, dtype=torch.float32)
    x_embed = not_mask.cumsum(2, dtype=torch.float32)
    eps, scale = 1e-6, 2 * math.pi
    y_embed = y_embed / (y_embed[:, -1:, :] + eps) * scale
    x_embed = x_embed / (x_embed[:, :, -1:] + eps) * scale
    dim_t = torch.arange(num_pos_feats, dtype=torch.float32)
    dim_t = temperature ** (2 * (dim_t // 2) / num_pos_feats)
    pos_x = x_embed[:, :, :, None] / dim_t
    pos_y = y_embed[:, :, :, None] / dim_t
    pos_x = torch.stack((pos_x[:, :, :, 0::2].sin(), pos_x[:, :, :, 1::2].cos()), dim=4).flatten(3)
    pos_y = torch.stack((pos_y[:, :, :, 0::2].sin(), pos_y[:, :, :, 1::2].cos()), dim=4).flatten(3)
    return torch.cat((pos_y, pos_x), dim=3).permute(0, 3, 1, 2)


def mha(sd, p: str, query, key, value, nheads: int, key_padding_mask=None, dropout_p: float = 0.0):
    """nn.MultiheadAttention forward as called at transformer.py:214-215,283-291.
    Inputs are [L, B, d] (sequence first).  q is scaled by 1/sqrt(head_dim) after projection,
    masked keys get -inf before the softmax, dropout acts on the attention weights."""
    Lq, B, d = query.shape
    Lk = key.shape[0]
    hd = d // nheads
    W, bvec = sd[p + "in_proj_weight"], sd[p + "in_proj_bias"]
    q = F.linear(query, W[:d], bvec[:d])
    k = F.linear(key, W[d:2 * d], bvec[d:2 * d])
    v = F.linear(value, W[2 * d:], bvec[2 * d:])
    q = q.reshape(Lq, B * nheads, hd).transpose(0, 1) * (1.0 / math.sqrt(hd))
    k = k.reshape(Lk, B * nheads, hd).transpose(0, 1)
    v = v.reshape(Lk, B * nheads, hd).transpose(0, 1)
    logits = torch.bmm(q, k.transpose(1, 2))                        # [B*h, Lq, Lk]
    if key_padding_mask is not None:
        m = key_padding_mask[:, None, None, :].expand(B, nheads, Lq, Lk).reshape(B * nheads, Lq, Lk)
        logits = logits.masked_fill(m, float("-inf"))
    attn = torch.softmax(logits, dim=-1)
    attn = _drop(attn, dropout_p, p + "attn")                       # nn.MultiheadAttention(dropout=p): on the weights
    out = torch.bmm(attn, v).transpose(0, 1).reshape(Lq, B, d)
    return F.linear(out, sd[p + "out_proj.weight"], sd[p + "out_proj.bias"])


def _ln(sd, p, x):
    return F.layer_norm(x, (x.shape[-1],), sd[p + "weight"], sd[p + "bias"], 1e-5)


# Mask injection for the dropout parity test: DROP_HOOK(site, x) -> keep mask (0/1, broadcastable to x) or None.
# Sites are named by the parameter prefix of the layer + {"self_attn.attn", "multihead_attn.attn", "drop1", "drop2",
# "drop3", "ffn"}.  Without a hook the masks come from torch's global generator like in the reference.
DROP_HOOK = None


def _drop(x, p, site=None):
    if p <= 0.0:
        return x
    if DROP_HOOK is not None:
        m = DROP_HOOK(site, x)
        if m is not None:
            return x * m.to(x.dtype) / (1.0 - p)
    return F.dropout(x, p=p)


def encoder_layer(sd, p: str, src, pos, nheads: int, key_padding_mask=None, dropout_p: float = 0.0):
    """TransformerEncoderLayer.forward_post, transformer.py:211-224."""
    q = k = src + pos
    src2 = mha(sd, p + "self_attn.", q, k, src, nheads, key_padding_mask, dropout_p)
    src = _ln(sd, p + "norm1.", src + _drop(src2, dropout_p, p + "drop1"))
    src2 = F.linear(_drop(F.relu(F.linear(src, sd[p + "linear1.weight"], sd[p + "linear1.bias"])), dropout_p, p + "ffn"),
                    sd[p + "linear2.weight"], sd[p + "linear2.bias"])
    return _ln(sd, p + "norm2.", src + _drop(src2, dropout_p, p + "drop2"))


def decoder_layer(sd, p: str, tgt, memory, pos, query_pos, nheads: int, dropout_p: float = 0.0):
    """TransformerDecoderLayer.forward_post, transformer.py:274-295."""
    q = k = tgt + query_pos
    tgt2 = mha(sd, p + "self_attn.", q, k, tgt, nheads, None, dropout_p)
    tgt = _ln(sd, p + "norm1.", tgt + _drop(tgt2, dropout_p, p + "drop1"))
    tgt2 = mha(sd, p + "multihead_attn.", tgt + query_pos, memory + pos, memory, nheads, None, dropout_p)
    tgt = _ln(sd, p + "norm2.", tgt + _drop(tgt2, dropout_p, p + "drop2"))
    tgt2 = F.linear(_drop(F.relu(F.linear(tgt, sd[p + "linear1.weight"], sd[p + "linear1.bias"])), dropout_p, p + "ffn"),
                    sd[p + "linear2.weight"], sd[p + "linear2.bias"])
    return _ln(sd, p + "norm3.", tgt + _drop(tgt2, dropout_p, p + "drop3"))


def cvae_encode(sd, cfg, qpos, actions, is_pad, eps, dropout_p: float = 0.0):
    """DETRVAE.encode training branch, detr_vae.py:115-151, with the reparametrisation noise
    injected (detr_vae.py:19-22 draws it from torch's global generator)."""
    B = qpos.shape[0]
    Z = cfg["latent_dim"]
    action_embed = F.linear(actions, sd["encoder_action_proj.weight"], sd["encoder_action_proj.bias"])
    qpos_embed = F.linear(qpos, sd["encoder_joint_proj.weight"], sd["encoder_joint_proj.bias"]).unsqueeze(1)
    cls_embed = sd["cls_embed.weight"].unsqueeze(0).repeat(B, 1, 1)
    x = torch.cat([cls_embed, qpos_embed, action_embed], dim=1).permute(1, 0, 2)         # [nq+2, B, d]
    mask = torch.cat([torch.zeros((B, 2), dtype=torch.bool), is_pad], dim=1)
    pos = sd["pos_table"].permute(1, 0, 2)                                                # [nq+2, 1, d]
    for i in range(cfg["enc_layers"]):
        x = encoder_layer(sd, f"encoder.layers.{i}.", x, pos, cfg["nheads"], mask, dropout_p)
    latent_info = F.linear(x[0], sd["latent_proj.weight"], sd["latent_proj.bias"])
    mu, logvar = latent_info[:, :Z], latent_info[:, Z:]
    latent_sample = mu + logvar.div(2).exp() * eps
    return latent_sample, mu, logvar


def detrvae_forward(sd, cfg, qpos, image_normalized, actions=None, is_pad=None, eps=None,
                    dropout_p: float = 0.0, faithful_decoder: bool = False, taps: dict = None):
    """DETRVAE.forward, detr_vae.py:163-254 (RGB path, no depth / pointcloud / vq).

    ``image_normalized`` is [B,K,3,H,W] *after* the ImageNet normalisation of policy.py:268-272.
    Returns dict(a_hat [B,nq,A], is_pad_hat [B,nq,1], mu, logvar).  ``taps`` (optional dict) receives
    intermediate activations in the engine's layouts (per-stage parity diagnostics): feat{k} = camera k's
    layer4 output as NHWC, src / memory = encoder input / output as [B, L, d], hs = decoder output [B, nq, d]."""
    B = qpos.shape[0]
    d, nh = cfg["hidden_dim"], cfg["nheads"]
    if actions is not None and not cfg.get("no_encoder"):
        latent_sample, mu, logvar = cvae_encode(sd, cfg, qpos, actions, is_pad, eps, dropout_p)
    else:                                                                                  # detr_vae.py:109-112 / :155-158
        mu = logvar = None
        latent_sample = torch.zeros((B, cfg["latent_dim"]), dtype=torch.float32)          # detr_vae.py:157
    latent_input = F.linear(latent_sample, sd["latent_out_proj.weight"], sd["latent_out_proj.bias"])

    feats, poss = [], []
    for k in range(cfg["num_cameras"]):
        f = resnet18_layer4(sd, f"backbones.{k}.0.body.", image_normalized[:, k])
        if taps is not None:
            taps[f"feat{k}"] = f.detach().permute(0, 2, 3, 1).contiguous()
        poss.append(position_embedding_sine(f.shape[2], f.shape[3], d // 2))
        feats.append(F.conv2d(f, sd["input_proj.weight"], sd["input_proj.bias"]))
    proprio_input = F.linear(qpos, sd["input_proj_robot_state.weight"], sd["input_proj_robot_state.bias"])
    src = torch.cat(feats, dim=3)                                                          # detr_vae.py:216 (W axis)
    pos = torch.cat(poss, dim=3)

    # Transformer.forward, transformer.py:58-122
    src = src.flatten(2).permute(2, 0, 1)                                                  # [hw, B, d]
    pos = pos.flatten(2).permute(2, 0, 1)                                                  # [hw, 1, d]
    pos = torch.cat([sd["additional_pos_embed.weight"].unsqueeze(1), pos], dim=0)          # [2+hw, 1, d]
    src = torch.cat([torch.stack([latent_input, proprio_input], dim=0), src], dim=0)       # [2+hw, B, d]
    query_pos = sd["query_embed.weight"].unsqueeze(1)                                       # [nq, 1, d]
    memory = src
    for i in range(cfg["enc_layers"]):
        memory = encoder_layer(sd, f"transformer.encoder.layers.{i}.", memory, pos, nh, None, dropout_p)
    if taps is not None:
        taps["src"] = src.detach().permute(1, 0, 2).contiguous()
        taps["memory"] = memory.detach().permute(1, 0, 2).contiguous()
    tgt = torch.zeros((cfg["num_queries"], B, d), dtype=torch.float32)
    # transformer.py:157-186 with return_intermediate=True: every layer output goes through the shared
    # final norm; detr_vae.py:245 keeps index [0] only -> decoder layers 1.. are dead work.
    n_run = cfg["dec_layers"] if faithful_decoder else 1
    out, first = tgt, None
    for i in range(n_run):
        out = decoder_layer(sd, f"transformer.decoder.layers.{i}.", out, memory, pos, query_pos, nh, dropout_p)
        if i == 0:
            first = _ln(sd, "transformer.decoder.norm.", out)
    hs = first.transpose(0, 1)                                                              # [B, nq, d]
    if taps is not None:
        taps["hs"] = hs.detach().contiguous()
    a_hat = F.linear(hs, sd["action_head.weight"], sd["action_head.bias"])
    is_pad_hat = F.linear(hs, sd["is_pad_head.weight"], sd["is_pad_head.bias"])
    return dict(a_hat=a_hat, is_pad_hat=is_pad_hat, mu=mu, logvar=logvar)


def normalize_image(image: torch.Tensor) -> torch.Tensor:
    """policy.py:268-272 -- torchvision Normalize broadcast over [B,K,3,H,W]."""
    mean = torch.tensor(IMAGENET_MEAN, dtype=torch.float32).view(1, 1, 3, 1, 1)
    std = torch.tensor(IMAGENET_STD, dtype=torch.float32).view(1, 1, 3, 1, 1)
    return (image - mean) / std


def kl_divergence(mu, logvar):
    """policy.py:378-391 -> total_kld ([1] tensor): sum over latent dims, mean over batch."""
    klds = -0.5 * (1 + logvar - mu.pow(2) - logvar.exp())
    return klds.sum(1).mean(0, True)


def policy_forward(sd, cfg, qpos, image, actions=None, is_pad=None, eps=None, dropout_p: float = 0.0,
                   faithful_decoder: bool = False, taps: dict = None):
    """ACTPolicy.__call__, policy.py:264-332.  Training/validation call returns
    dict(l1, kl, loss, a_hat, mu, logvar); inference call returns dict(a_hat)."""
    image = normalize_image(image)
    nq = cfg["num_queries"]
    if actions is not None:
        actions = actions[:, :nq]
        is_pad = is_pad[:, :nq]
        o = detrvae_forward(sd, cfg, qpos, image, actions, is_pad, eps, dropout_p, faithful_decoder, taps)
        # policy.py:306-309: kl = 0 when the model has no CVAE encoder
        total_kld = torch.zeros(1) if cfg.get("no_encoder") else kl_divergence(o["mu"], o["logvar"])
        all_l1 = F.l1_loss(actions, o["a_hat"], reduction="none")
        l1 = (all_l1 * ~is_pad.unsqueeze(-1)).mean()                   # mean over B*nq*A incl. padded (policy.py:315)
        loss = l1 + total_kld[0] * cfg["kl_weight"]
        return dict(l1=l1, kl=total_kld[0], loss=loss, a_hat=o["a_hat"], mu=o["mu"], logvar=o["logvar"])
    o = detrvae_forward(sd, cfg, qpos, image, dropout_p=0.0, faithful_decoder=faithful_decoder, taps=taps)
    return dict(a_hat=o["a_hat"])


def trainable_names(cfg: dict) -> List[str]:
    return [n for n, _, kind in state_spec(cfg) if not is_buffer(kind)]


def loss_and_grads(sd, cfg, batch, dropout_p: float = 0.0) -> Dict[str, object]:
    """loss.backward() of imitate_episodes.py:603-606 through the restatement (autograd on CPU).
    Parameters that autograd never reaches (is_pad_head, decoder layers >= 1) get grad None,
    except decoder layers 1.. which the reference reaches with exact zeros;
    callers treat None as 'skipped by AdamW' and zeros as 'decayed only'."""
    names = trainable_names(cfg)
    leaf = {n: (sd[n].clone().requires_grad_(True) if n in names else sd[n]) for n in sd}
    out = policy_forward(leaf, cfg, batch["qpos"], batch["image"], batch["actions"], batch["is_pad"], batch["eps"],
                         dropout_p=dropout_p)
    out["loss"].backward()
    grads = {n: leaf[n].grad for n in names}
    return dict(out={k: (v.detach() if v is not None else None) for k, v in out.items()}, grads=grads)


def adamw_step(p, g, m, v, step: int, lr: float, wd: float = 1e-4, b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8):
    """torch.optim.AdamW single-tensor update (detr/main.py:109 defaults), step counted from 1."""
    p = p * (1.0 - lr * wd)
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    bc1 = 1.0 - b1 ** step
    bc2 = 1.0 - b2 ** step
    denom = v.sqrt() / math.sqrt(bc2) + eps
    p = p - (lr / bc1) * (m / denom)
    return p, m, v


# --------------------------------------------------------------------------------------
# temporal ensemble (imitate_episodes.py:339, 403-411), literal
# --------------------------------------------------------------------------------------
class TemporalAggOracle:
    """One environment's ``all_time_actions`` buffer, literal layout [T, T+nq, A] float32."""

    def __init__(self, max_timesteps: int, num_queries: int, action_dim: int = 16):
        self.nq = num_queries
        self.buf = torch.zeros([max_timesteps, max_timesteps + num_queries, action_dim])   # :339

    def step(self, t: int, all_actions: torch.Tensor) -> torch.Tensor:
        """all_actions [1,nq,A] f32 -> raw_action [1,A] float64 (f32 * f64 promotes)."""
        self.buf[[t], t:t + self.nq] = all_actions                                         # :403
        actions_for_curr_step = self.buf[:, t]                                             # :404
        actions_populated = torch.all(actions_for_curr_step != 0, axis=1)                  # :405
        actions_for_curr_step = actions_for_curr_step[actions_populated]                   # :406
        k = 0.01
        exp_weights = np.exp(-k * np.arange(len(actions_for_curr_step)))                   # :408
        exp_weights = exp_weights / exp_weights.sum()                                      # :409
        exp_weights = torch.from_numpy(exp_weights).unsqueeze(dim=1)                       # :410
        return (actions_for_curr_step * exp_weights).sum(dim=0, keepdim=True)              # :411


def temporal_agg_stream(T: int, nq: int, A: int, seed: int):
    """Chunk stream of the temporal-ensemble fixtures, regenerated identically by the tests (only the outputs are
    committed for the full-size case): N(0,1) chunks with single exact zeros planted in ~5 % of the rows."""
    g = torch.Generator().manual_seed(seed)
    chunks = torch.randn((T, 1, nq, A), generator=g)
    zero_at = torch.rand((T, nq), generator=g) < 0.05
    chunks[:, 0][zero_at] = chunks[:, 0][zero_at] * torch.tensor([0.0] + [1.0] * (A - 1))   # plant a single exact zero
    return chunks


# --------------------------------------------------------------------------------------
# data ingest contract (utils.py:47-52, 60-175, 239-248), literal per-item restatement without h5py
# --------------------------------------------------------------------------------------
def batch_sampler_reference(batch_size, episode_len_l, sample_weights):
    """utils.py:239-248 verbatim (numpy global RandomState)."""
    sample_probs = np.array(sample_weights) / np.sum(sample_weights) if sample_weights is not None else None
    sum_dataset_len_l = np.cumsum([0] + [np.sum(episode_len) for episode_len in episode_len_l])
    while True:
        batch = []
        for _ in range(batch_size):
            episode_idx = np.random.choice(len(episode_len_l), p=sample_probs)
            step_idx = np.random.randint(sum_dataset_len_l[episode_idx], sum_dataset_len_l[episode_idx + 1])
            batch.append(step_idx)
        yield batch


def episodic_getitem_reference(episodes, norm_stats, chunk_size, index):
    """EpisodicDataset.__getitem__ (utils.py:60-175, ACT branch, no augmentation) over in-memory episodes
    [dict(frames uint8 [T,K,H,W,3], qpos [T,S], action [T,A], sim bool)] -> (image f32 [K,3,H,W], qpos, action, is_pad)."""
    episode_len = [e["action"].shape[0] for e in episodes]
    cumulative_len = np.cumsum(episode_len)
    episode_index = int(np.argmax(cumulative_len > index))                                  # :49
    start_ts = int(index - (cumulative_len[episode_index] - episode_len[episode_index]))   # :50
    ep = episodes[episode_index]
    action = ep["action"]
    max_episode_len = max(episode_len)
    if ep["sim"]:
        action, action_len = action[start_ts:], episode_len[episode_index] - start_ts      # :109-110
    else:
        action = action[max(0, start_ts - 1):]                                              # :112
        action_len = episode_len[episode_index] - max(0, start_ts - 1)
    padded_action = np.zeros((max_episode_len, action.shape[1]), dtype=np.float32)          # :116
    padded_action[:action_len] = action
    is_pad = np.zeros(max_episode_len)
    is_pad[action_len:] = 1
    padded_action, is_pad = padded_action[:chunk_size], is_pad[:chunk_size]                 # :121-122
    image_data = torch.einsum("k h w c -> k c h w", torch.from_numpy(ep["frames"][start_ts])) / 255.0   # :139, 157
    action_data = (torch.from_numpy(padded_action).float() - norm_stats["action_mean"]) / norm_stats["action_std"]   # :164
    qpos_data = (torch.from_numpy(ep["qpos"][start_ts]).float() - norm_stats["qpos_mean"]) / norm_stats["qpos_std"]  # :166
    return image_data, qpos_data, action_data, torch.from_numpy(is_pad).bool()
