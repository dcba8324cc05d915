"""Python handle over the C-ABI policy engine (``act_engine_*`` in include/act_b200.h).

Owns the flat fp32 parameter / gradient / Adam-moment tensors (torch CUDA memory) and exposes each reference
``state_dict`` entry as a view into them, so checkpoints interchange with the reference by name and shape."""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from typing import Dict, Optional

import torch

from . import _lib
from ._lib import check

KIND_PARAM, KIND_DEAD, KIND_NOGRAD, KIND_BACKBONE, KIND_BUFFER = range(5)
R_MAIN, R_CVAE, R_DEAD, R_NOGRAD, R_BACKBONE, R_BUFFER = range(6)


class EngineConfig(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("batch_max", "num_cameras", "img_h", "img_w", "hidden_dim", "nheads",
                                      "dim_feedforward", "enc_layers", "dec_layers", "num_queries", "state_dim",
                                      "action_dim", "latent_dim")] + \
               [(n, C.c_float) for n in ("kl_weight", "lr", "lr_backbone", "weight_decay", "beta1", "beta2", "adam_eps")] + \
               [("no_encoder", C.c_int), ("train", C.c_int)]


def _stream():
    return torch.cuda.current_stream().cuda_stream


class ActEngine:
    def __init__(self, batch_max=8, num_cameras=1, img_h=480, img_w=640, hidden_dim=512, nheads=8,
                 dim_feedforward=3200, enc_layers=4, dec_layers=7, num_queries=100, state_dim=14, action_dim=16,
                 latent_dim=32, kl_weight=10.0, lr=1e-5, lr_backbone=1e-5, weight_decay=1e-4, betas=(0.9, 0.999),
                 eps=1e-8, train=True, device: Optional[torch.device] = None, share: "Optional[ActEngine]" = None,
                 no_encoder: bool = False):
        if not torch.cuda.is_available():
            raise _lib.ActB200Error("act_plus_plus_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.lib = _lib.lib()
        cfg = EngineConfig(batch_max, num_cameras, img_h, img_w, hidden_dim, nheads, dim_feedforward, enc_layers,
                           dec_layers, num_queries, state_dim, action_dim, latent_dim, kl_weight, lr, lr_backbone,
                           weight_decay, betas[0], betas[1], eps, 1 if no_encoder else 0, 1 if train else 0)
        self.cfg = cfg
        self.train_enabled = bool(train)
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(self.lib.act_engine_create(C.byref(cfg), C.byref(self._h)), "act_engine_create")
            n = self.lib.act_engine_total_floats(self._h)
            if share is not None:
                # several engines (different batch / image geometry, train vs inference workspace) over ONE set of
                # master tensors; `_shared` carries the optimizer step count and the parameter version
                assert share.flat_params.numel() == n, "engines sharing parameters must have the same model config"
                self.flat_params, self._shared = share.flat_params, share._shared
                self.flat_grads, self.exp_avg, self.exp_avg_sq = share.flat_grads, share.exp_avg, share.exp_avg_sq
                train = train and self.flat_grads is not None
            else:
                self.flat_params = torch.zeros(n, dtype=torch.float32, device=self.device)
                self.flat_grads = torch.zeros(n, dtype=torch.float32, device=self.device) if train else None
                self.exp_avg = torch.zeros(n, dtype=torch.float32, device=self.device) if train else None
                self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=self.device) if train else None
                self._shared = dict(step=0, version=0, params={})
            check(self.lib.act_engine_bind(self._h, self.flat_params.data_ptr(),
                                           self.flat_grads.data_ptr() if train else None,
                                           self.exp_avg.data_ptr() if train else None,
                                           self.exp_avg_sq.data_ptr() if train else None), "act_engine_bind")
        self.tensors: "OrderedDict[str, dict]" = OrderedDict()
        name = C.create_string_buffer(256)
        shape = (C.c_int64 * 4)()
        ndim, kind, region = C.c_int(), C.c_int(), C.c_int()
        off, numel = C.c_int64(), C.c_int64()
        for i in range(self.lib.act_engine_num_tensors(self._h)):
            check(self.lib.act_engine_tensor_info(self._h, i, name, 256, shape, C.byref(ndim), C.byref(off),
                                                  C.byref(numel), C.byref(kind), C.byref(region)), "tensor_info")
            shp = tuple(int(shape[k]) for k in range(ndim.value))
            self.tensors[name.value.decode()] = dict(shape=shp, offset=off.value, numel=numel.value, kind=kind.value,
                                                     region=region.value)
        self._synced_version = -1 if share is not None else 0
        self.last_was_train = False
        self._train_bound = bool(train)
        import weakref
        self._shared.setdefault("engines", weakref.WeakSet()).add(self)      # every engine over these master tensors

    # ------------------------------------------------------------------ parameters
    def view(self, name: str, flat: Optional[torch.Tensor] = None) -> torch.Tensor:
        t = self.tensors[name]
        flat = self.flat_params if flat is None else flat
        return flat[t["offset"]:t["offset"] + t["numel"]].view(t["shape"])

    def grad_view(self, name: str) -> torch.Tensor:
        return self.view(name, self.flat_grads)

    def region(self, r: int):
        off, n = C.c_int64(), C.c_int64()
        check(self.lib.act_engine_region(self._h, r, C.byref(off), C.byref(n)), "act_engine_region")
        return off.value, n.value

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((k, self.view(k)) for k in self.tensors)

    def load_state_dict(self, sd: Dict[str, torch.Tensor], strict: bool = True):
        missing = [k for k in self.tensors if k not in sd]
        unexpected = [k for k in sd if k not in self.tensors]
        if strict and (missing or unexpected):
            raise RuntimeError(f"state_dict mismatch: missing {missing[:5]}... unexpected {unexpected[:5]}...")
        with torch.no_grad():
            for k, v in sd.items():
                if k in self.tensors:
                    dst = self.view(k)
                    if tuple(v.shape) != tuple(dst.shape):
                        raise RuntimeError(f"size mismatch for {k}: {tuple(v.shape)} vs {tuple(dst.shape)}")
                    dst.copy_(v.to(device=self.device, dtype=torch.float32))
        self.mark_params_changed()
        self.refresh()
        return missing, unexpected

    @classmethod
    def sharing(cls, master: "ActEngine", batch_max: int, img_h: int, img_w: int, train: bool) -> "ActEngine":
        c = master.cfg
        return cls(batch_max=batch_max, num_cameras=c.num_cameras, img_h=img_h, img_w=img_w, hidden_dim=c.hidden_dim,
                   nheads=c.nheads, dim_feedforward=c.dim_feedforward, enc_layers=c.enc_layers, dec_layers=c.dec_layers,
                   num_queries=c.num_queries, state_dim=c.state_dim, action_dim=c.action_dim, latent_dim=c.latent_dim,
                   kl_weight=c.kl_weight, lr=c.lr, lr_backbone=c.lr_backbone, weight_decay=c.weight_decay,
                   betas=(c.beta1, c.beta2), eps=c.adam_eps, train=train, device=master.device, share=master,
                   no_encoder=bool(c.no_encoder))

    @property
    def step_count(self) -> int:
        return self._shared["step"]

    def param(self, name: str) -> torch.nn.Parameter:
        """nn.Parameter aliasing the flat storage (its .grad aliases the flat gradient buffer)."""
        cache = self._shared["params"]
        if name not in cache:
            p = torch.nn.Parameter(self.view(name), requires_grad=self.tensors[name]["kind"] != KIND_BUFFER)
            if self.flat_grads is not None and self.tensors[name]["kind"] not in (KIND_BUFFER, KIND_NOGRAD):
                p.grad = self.grad_view(name)
            cache[name] = p
        return cache[name]

    def mark_params_changed(self):
        self._shared["version"] += 1

    def refresh(self):
        check(self.lib.act_engine_refresh(self._h, _stream()), "act_engine_refresh")
        self._synced_version = self._shared["version"]

    # ------------------------------------------------------------------ steps
    def forward(self, qpos, image, actions=None, is_pad=None, eps=None, want_is_pad_hat=False):
        """Returns dict(a_hat[, is_pad_hat, mu, logvar, losses]) -- all fp32 CUDA tensors."""
        c = self.cfg
        B = qpos.shape[0]
        assert qpos.is_cuda and image.is_cuda, "inputs must be CUDA tensors"
        qpos = qpos.contiguous().float()
        if image.dtype == torch.uint8:             # dataset-native frames [B, K, H, W, 3]: /255 + normalise fused on device
            image = image.contiguous()
            assert image.shape == (B, c.num_cameras, c.img_h, c.img_w, 3), f"uint8 image shape {tuple(image.shape)}"
            fmt = 1
        else:
            image = image.contiguous().float()
            assert image.shape == (B, c.num_cameras, 3, c.img_h, c.img_w), f"image shape {tuple(image.shape)}"
            fmt = 0
        if fmt != getattr(self, "_image_fmt", 0):
            check(self.lib.act_engine_set_image_format(self._h, fmt), "act_engine_set_image_format")
            self._image_fmt = fmt
        assert qpos.shape[1] == c.state_dim
        out = dict(a_hat=torch.empty((B, c.num_queries, c.action_dim), dtype=torch.float32, device=self.device))
        iph = torch.empty((B, c.num_queries, 1), dtype=torch.float32, device=self.device) if want_is_pad_hat else None
        a_ptr = p_ptr = e_ptr = mu_ptr = lv_ptr = ls_ptr = None
        abs_ = pbs = 0
        keep = []
        if actions is not None:
            actions = actions.float()
            if actions.stride(2) != 1 or actions.stride(1) != actions.shape[2]:
                actions = actions.contiguous()
            pad8 = is_pad.to(torch.uint8) if is_pad.dtype != torch.uint8 else is_pad
            if is_pad.dtype == torch.bool:
                pad8 = is_pad.view(torch.uint8)
            if pad8.stride(1) != 1:
                pad8 = pad8.contiguous()
            assert actions.shape[1] >= c.num_queries and pad8.shape[1] >= c.num_queries
            if eps is None:
                eps = torch.randn((B, c.latent_dim), dtype=torch.float32, device=self.device)
            eps = eps.contiguous().float()
            out["mu"] = torch.empty((B, c.latent_dim), dtype=torch.float32, device=self.device)
            out["logvar"] = torch.empty_like(out["mu"])
            out["losses"] = torch.empty(3, dtype=torch.float32, device=self.device)
            a_ptr, p_ptr, e_ptr = actions.data_ptr(), pad8.data_ptr(), eps.data_ptr()
            abs_, pbs = actions.stride(0), pad8.stride(0)
            mu_ptr, lv_ptr, ls_ptr = out["mu"].data_ptr(), out["logvar"].data_ptr(), out["losses"].data_ptr()
            keep += [actions, pad8, eps]
        self._keep = keep + [qpos, image]          # inputs must stay alive until the backward has run
        if self._synced_version != self._shared["version"]:
            self.refresh()                         # another engine stepped the shared parameters
        self.last_was_train = actions is not None
        check(self.lib.act_engine_forward(self._h, B, qpos.data_ptr(), image.data_ptr(), a_ptr, abs_, p_ptr, pbs, e_ptr,
                                          out["a_hat"].data_ptr(), iph.data_ptr() if iph is not None else None,
                                          mu_ptr, lv_ptr, ls_ptr, _stream()), "act_engine_forward")
        if iph is not None:
            out["is_pad_hat"] = iph
        return out

    def infer_graphed(self, qpos, image, tag=0):
        """The inference call policy(qpos, image) (policy.py:320-322) replayed from a CUDA graph once the same shape / dtype
        has been seen 3 times: the ~100 launches of an inference pass become one graph launch (a batch-1 call is launch-bound).
        Returns a STATIC [B, num_queries, action_dim] tensor, overwritten by the next call with the same key -- clone it to
        keep it.  Parameters are read through the engine's buffers, so optimiser steps / load_state_dict are seen by replays."""
        key = (tag, tuple(qpos.shape), tuple(image.shape), image.dtype, qpos.dtype)
        gs = self.__dict__.setdefault("_infer_graphs", {})
        ent = gs.get(key)
        if ent is None:
            ent = gs[key] = dict(seen=0, graph=None)
        if ent["graph"] is None:
            ent["seen"] += 1
            if ent["seen"] < 3 or torch.cuda.is_current_stream_capturing():
                return self.forward(qpos, image)["a_hat"]
            ent["qpos"], ent["image"] = qpos.clone(), image.clone()
            if self._synced_version != self._shared["version"]:
                self.refresh()
            cur = torch.cuda.current_stream()
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):                      # warm-up on the capture stream (lazy inits happen here)
                self.forward(ent["qpos"], ent["image"])
            cur.wait_stream(side)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                ent["out"] = self.forward(ent["qpos"], ent["image"])["a_hat"]
            ent["graph"] = g
        if self._synced_version != self._shared["version"]:
            self.refresh()
        ent["qpos"].copy_(qpos, non_blocking=True)
        ent["image"].copy_(image, non_blocking=True)
        ent["graph"].replay()
        return ent["out"]

    def backward(self, stage: int = 0, loss_grad: float = 1.0):
        check(self.lib.act_engine_backward(self._h, stage, float(loss_grad), _stream()), "act_engine_backward")

    def zero_grad(self, overlap_forward: bool = False):
        """overlap_forward: clear the gradients beside the next forward (the next backward / optimiser call waits)."""
        if overlap_forward:
            check(self.lib.act_engine_zero_grad_async(self._h, _stream()), "act_engine_zero_grad_async")
        else:
            check(self.lib.act_engine_zero_grad(self._h, _stream()), "act_engine_zero_grad")

    def set_loss_grad(self, gvec: Optional[torch.Tensor]):
        """Device vector d/d{l1, kl, loss} read by the following backward calls (None: the scalar loss_grad argument)."""
        if gvec is not None:
            assert gvec.is_cuda and gvec.dtype == torch.float32 and gvec.numel() == 3 and gvec.is_contiguous()
        self._gvec = gvec                                  # keep alive while the engine points at it
        check(self.lib.act_engine_set_loss_grad(self._h, gvec.data_ptr() if gvec is not None else None), "act_engine_set_loss_grad")

    def join_lanes(self):
        """The current stream waits for the engine's internal lanes (closes a graph capture after a forward)."""
        check(self.lib.act_engine_join_lanes(self._h, _stream()), "act_engine_join_lanes")

    def note_graph_step(self):
        """Book-keeping after a CUDA-graph replay of backward+step (the device counters advanced inside the graph)."""
        self._shared["step"] += 1
        self._shared["version"] += 1
        self._synced_version = self._shared["version"]

    def backward_and_step(self, loss_grad: float = 1.0, advance_on_device: bool = False):
        """loss.backward(); optimizer.step() as one call: each region's AdamW overlaps the remaining backward.
        advance_on_device: the optimiser step number is taken from (and advanced in) the engine's device counter, so a
        CUDA graph of this call stays correct on replay; the caller then does the host book-keeping (note_graph_step)."""
        if advance_on_device:
            check(self.lib.act_engine_backward_and_step(self._h, 0, float(loss_grad), _stream()), "act_engine_backward_and_step")
            return
        self._shared["step"] += 1
        check(self.lib.act_engine_backward_and_step(self._h, self._shared["step"], float(loss_grad), _stream()),
              "act_engine_backward_and_step")
        self._shared["version"] += 1
        self._synced_version = self._shared["version"]

    def backward_and_step_dp(self, grads_ready, loss_grad: float = 1.0):
        """Data-parallel loss.backward(); optimizer.step(): `grads_ready(offset, count, stream_ptr)` is called while the
        step is enqueued, once per contiguous range of `flat_grads` whose producers are enqueued on CUDA stream
        `stream_ptr`; it must enqueue the cross-rank average of that range and order the stream after it.  The range's
        AdamW follows behind it, so every collective overlaps the backward work still running on other streams."""
        import ctypes as C
        err = []

        def _cb(_user, off, n, stream):
            try:
                grads_ready(int(off), int(n), int(stream or 0))
                return 0
            except Exception as ex:            # an exception must not unwind through the C frames
                err.append(ex)
                return -1

        cb = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_longlong, C.c_longlong, C.c_void_p)(_cb)
        self._shared["step"] += 1
        rc = self.lib.act_engine_backward_and_step_dp(self._h, self._shared["step"], float(loss_grad), _stream(), cb, None)
        if err:
            raise err[0]
        check(rc, "act_engine_backward_and_step_dp")
        self._shared["version"] += 1
        self._synced_version = self._shared["version"]

    # ---- data-parallel step over NVLink peer memory (csrc/dp_p2p.cu; include/act_b200.h "act_engine_dp_*")
    DP_BLOB_BYTES = 296

    def dp_local_ptrs(self):
        """(grads, params, shadow, signal) device addresses of this engine's exchange buffers."""
        import ctypes as C
        out = [C.c_void_p() for _ in range(4)]
        check(self.lib.act_engine_dp_local(self._h, *[C.byref(o) for o in out]), "act_engine_dp_local")
        return tuple(int(o.value) for o in out)

    def dp_set_peers(self, rank: int, world: int, table):
        """table: four sequences (grads, params, shadow, signal) of `world` device addresses valid on this device."""
        import ctypes as C
        arrs = [(C.c_void_p * world)(*[int(x) for x in col]) for col in table]
        check(self.lib.act_engine_dp_set_peers(self._h, rank, world, *arrs), "act_engine_dp_set_peers")
        self._dp_world, self._dp_rank = world, rank

    def _dp_connect_symmetric(self, group) -> bool:
        """NVLS route of dp_connect: move the flat parameter / gradient buffers, the bf16 shadow and the barrier flags into
        symmetric memory (torch.distributed._symmetric_memory: allocation + rendezvous only), hand the engine every rank's
        mapping of them plus the NVSwitch multicast addresses.  Returns False -- on every rank alike -- when the box has no
        multicast support (no NVSwitch / fabric manager) so that the caller falls back to CUDA-IPC peer mappings."""
        import torch.distributed as dist
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        n = self.flat_params.numel()
        err, sy, mine = None, self._shared.get("symm"), None
        try:
            import torch.distributed._symmetric_memory as symm
            grp = group if group is not None else dist.group.WORLD
            sig_words = int(self.lib.act_engine_dp_signal_bytes()) // 4
            fresh = None
            if sy is None:                                       # the master tensors move once; every engine over them follows
                fresh = [symm.empty(n, dtype=torch.float32, device=self.device) for _ in range(2)]
                fresh = dict(g=fresh[0], p=fresh[1], hg=symm.rendezvous(fresh[0], grp), hp=symm.rendezvous(fresh[1], grp))
            s = symm.empty(n, dtype=torch.bfloat16, device=self.device)          # bf16 shadow and barrier flags: per engine
            sig = symm.empty(sig_words, dtype=torch.int32, device=self.device)
            mine = dict(s=s, sig=sig, hs=symm.rendezvous(s, grp), hsig=symm.rendezvous(sig, grp))
            probe = [mine["hs"]] + ([fresh["hg"], fresh["hp"]] if fresh else [])
            if any(int(h.multicast_ptr) == 0 for h in probe):
                err = "no multicast mapping (NVLS unsupported on this box)"
        except Exception as ex:                                  # old torch, no symmetric-memory backend, no fabric ...
            err = f"{type(ex).__name__}: {ex}"
        ok = torch.tensor([0.0 if err else 1.0], device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if float(ok) == 0.0:
            self._dp_nvls_error = err or "unavailable on another rank"
            return False
        if sy is None:
            sy = fresh
            g, p = sy["g"], sy["p"]
            p.copy_(self.flat_params)
            g.zero_()
            torch.cuda.synchronize(self.device)
            for e in list(self._shared["engines"]):
                e.flat_params = p
                if e.flat_grads is not None:
                    e.flat_grads = g
                e.__dict__.pop("_infer_graphs", None)            # captured graphs hold the old addresses
                with torch.cuda.device(e.device):
                    check(e.lib.act_engine_bind(e._h, p.data_ptr(), g.data_ptr() if e._train_bound else None,
                                                e.exp_avg.data_ptr() if e._train_bound else None,
                                                e.exp_avg_sq.data_ptr() if e._train_bound else None), "act_engine_bind")
            for name, prm in self._shared["params"].items():     # nn.Parameter objects keep their identity (optimisers, hooks)
                prm.data = self.view(name)
                if prm.grad is not None:
                    prm.grad = self.grad_view(name)
            self._shared["symm"] = sy
        mine["sig"].zero_()
        torch.cuda.synchronize(self.device)
        check(self.lib.act_engine_dp_use_buffers(self._h, mine["s"].data_ptr(), mine["sig"].data_ptr()), "act_engine_dp_use_buffers")
        table = []
        for b, h in ((sy["g"], sy["hg"]), (sy["p"], sy["hp"]), (mine["s"], mine["hs"]), (mine["sig"], mine["hsig"])):
            ptrs = [int(x) for x in h.buffer_ptrs]
            ptrs[rank] = b.data_ptr()
            table.append(ptrs)
        self.dp_set_peers(rank, world, table)
        check(self.lib.act_engine_dp_set_multicast(self._h, int(sy["hg"].multicast_ptr), int(sy["hp"].multicast_ptr),
                                                   int(mine["hs"].multicast_ptr)), "act_engine_dp_set_multicast")
        self._dp_symm = mine                                     # keep the allocations and their mappings alive
        self._dp_nvls = True                                     # (parameter VALUES did not change: no refresh is due)
        dist.barrier(group=group)                                # nobody starts a step before every rank is mapped
        return True

    def dp_connect(self, group=None, multicast: Optional[bool] = None):
        """Map every rank's flat gradient / parameter / shadow buffers and barrier flags (one process per GPU on one node).
        With NVSwitch multicast (default above 4 ranks, where one switch-reduced load replaces world-1 peer loads and one
        store reaches every rank: 3.68 -> 3.57 ms/step at N=8; measured 0.7 % slower at N=4 and 7 % slower at N=2, where it
        saves little or no link traffic; `multicast=` / ACT_DP_NVLS=0|1 override) the buffers move into
        symmetric memory and the exchange kernel uses multimem.ld_reduce / multimem.st; otherwise CUDA-IPC handles are
        exchanged and the kernel does per-peer loads and stores.  Collective: every rank must call it; if any rank fails,
        every rank raises RuntimeError (so callers can fall back to NCCL consistently)."""
        import ctypes as C
        import os
        import torch.distributed as dist
        from ._lib import last_error
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        torch.cuda.synchronize(self.device)
        if getattr(self, "_dp_mapped", None) is not None:        # re-connect after dp_disconnect(): the mappings are still open
            self._dp_world, self._dp_rank = self._dp_mapped
            return
        self._dp_nvls = False
        if multicast is None:
            env = os.environ.get("ACT_DP_NVLS", "auto")
            multicast = world > 4 if env == "auto" else env == "1"
        if multicast and getattr(self, "_dp_world", 1) <= 1 and self._dp_connect_symmetric(group):
            return
        blob = C.create_string_buffer(self.DP_BLOB_BYTES)
        err = None
        if self.lib.act_engine_dp_export(self._h, blob, self.DP_BLOB_BYTES) != self.DP_BLOB_BYTES:
            err = "act_engine_dp_export: " + last_error()
        blobs = [None] * world
        dist.all_gather_object(blobs, None if err else bytes(blob.raw), group=group)
        if err is None and any(b is None for b in blobs):
            err = "a peer could not export its buffers"
        if err is None and self.lib.act_engine_dp_connect(self._h, rank, world, b"".join(blobs), self.DP_BLOB_BYTES) != 0:
            err = "act_engine_dp_connect: " + last_error()
        ok = torch.tensor([0.0 if err else 1.0], device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)   # also: nobody starts a step before every rank is mapped
        if float(ok) == 0.0:
            raise RuntimeError(f"peer-memory data-parallel set-up failed: {err or 'on another rank'}")
        self._dp_world, self._dp_rank = world, rank

    def dp_disconnect(self):
        """Go back to the NCCL / single-process step (the mappings stay open until the engine is destroyed)."""
        if getattr(self, "_dp_world", 1) > 1:
            self._dp_mapped = (self._dp_world, self._dp_rank)
        self._dp_world = 1

    @property
    def dp_connected(self) -> bool:
        return getattr(self, "_dp_world", 1) > 1

    def backward_and_step_p2p(self, loss_grad: float = 1.0, advance_on_device: bool = False):
        """Data-parallel loss.backward(); optimizer.step() with the gradient exchange fused into the optimiser kernel
        over peer memory (act_engine_backward_and_step_p2p).  Every rank must call it for the same step."""
        if advance_on_device:
            check(self.lib.act_engine_backward_and_step_p2p(self._h, 0, float(loss_grad), _stream()),
                  "act_engine_backward_and_step_p2p")
            return
        self._shared["step"] += 1
        check(self.lib.act_engine_backward_and_step_p2p(self._h, self._shared["step"], float(loss_grad), _stream()),
              "act_engine_backward_and_step_p2p")
        self._shared["version"] += 1
        self._synced_version = self._shared["version"]

    def dp_status(self) -> int:
        return int(self.lib.act_engine_dp_status(self._h))

    def dp_slices(self):
        """[(offset, count)] ranges of the flat buffers exchanged by the last p2p step."""
        import ctypes as C
        cap = 64
        off, cnt = (C.c_longlong * cap)(), (C.c_longlong * cap)()
        k = self.lib.act_engine_dp_slices(self._h, off, cnt, cap)
        return [(int(off[i]), int(cnt[i])) for i in range(min(k, cap))]

    def dp_gather_moments(self, group=None):
        """Moments are sharded by the p2p step (rank r holds the current exp_avg / exp_avg_sq of its slices only): make
        every rank's flat moment buffers complete again (before saving the optimiser state)."""
        import torch.distributed as dist
        from .dp import owner_slices
        world = dist.get_world_size(group)
        for off, n in self.dp_slices():
            for r, lo, hi in owner_slices(n, world):
                src = dist.get_global_rank(group, r) if group is not None else r
                for buf in (self.exp_avg, self.exp_avg_sq):
                    dist.broadcast(buf[off + lo:off + hi], src=src, group=group)

    def adamw_step(self):
        self._shared["step"] += 1
        check(self.lib.act_engine_adamw_step(self._h, self._shared["step"], _stream()), "act_engine_adamw_step")
        self._shared["version"] += 1
        self._synced_version = self._shared["version"]     # the fused step refreshed this engine's own shadows

    def begin_optimizer_step(self) -> int:
        """Advance the shared step counter; the per-region updates of this optimiser step all use the returned value."""
        self._shared["step"] += 1
        return self._shared["step"]

    def adamw_region(self, step: int, region: int, blocks_per_sm: int = 0):
        check(self.lib.act_engine_adamw_region(self._h, step, region, blocks_per_sm, _stream()), "act_engine_adamw_region")

    def end_optimizer_step(self):
        self._shared["version"] += 1
        self._synced_version = self._shared["version"]

    def set_concurrent(self, on: bool):
        """False: all launches in program order on the caller's stream (for per-kernel timing); call between steps."""
        check(self.lib.act_engine_set_concurrent(self._h, 1 if on else 0), "act_engine_set_concurrent")

    def set_lr(self, lr: float, lr_backbone: float):
        check(self.lib.act_engine_set_lr(self._h, lr, lr_backbone), "act_engine_set_lr")

    # ---- train-mode dropout (act_engine_set_dropout; csrc/dropout.cuh)
    SITE_ENC, SITE_CVAE, SITE_DEC = 0x100, 0x200, 0x300

    def set_dropout(self, p: float, seed: int = 0):
        """p = 0 switches dropout off (the reference's .eval()); masks are Philox(seed; element, site, step)."""
        check(self.lib.act_engine_set_dropout(self._h, float(p), int(seed) & 0xFFFFFFFFFFFFFFFF), "act_engine_set_dropout")
        self.dropout_p = float(p)

    def dropout_state(self):
        """(p, seed, step): step = value of the device counter, i.e. the step of the LAST training forward."""
        p, seed, step = C.c_float(), C.c_uint64(), C.c_uint32()
        check(self.lib.act_engine_dropout_state(self._h, C.byref(p), C.byref(seed), C.byref(step)), "act_engine_dropout_state")
        return p.value, seed.value, step.value

    def dropout_mask(self, site: int, n: int, step: Optional[int] = None) -> torch.Tensor:
        """uint8 keep mask of the first n elements of dropout site `site` (tests: the masks the kernels regenerate)."""
        p, seed, cur = self.dropout_state()
        out = torch.empty(n, dtype=torch.uint8, device=self.device)
        check(self.lib.act_dropout_mask(seed, site, cur if step is None else step, p, n, out.data_ptr(), _stream()),
              "act_dropout_mask")
        return out

    def debug_buffer(self, name: str) -> torch.Tensor:
        """Copy of an internal activation of the last forward ("src", "memory", "hs", "feat<k>"): parity diagnostics."""
        numel, dtype = C.c_int64(), C.c_int()
        check(self.lib.act_engine_debug_buffer(self._h, name.encode(), None, 0, C.byref(numel), C.byref(dtype), _stream()),
              "act_engine_debug_buffer")
        out = torch.empty(numel.value, dtype=torch.float32 if dtype.value == 0 else torch.bfloat16, device=self.device)
        check(self.lib.act_engine_debug_buffer(self._h, name.encode(), out.data_ptr(), out.numel() * out.element_size(),
                                               C.byref(numel), C.byref(dtype), _stream()), "act_engine_debug_buffer")
        return out

    def launches(self, reset: bool = False) -> int:
        return self.lib.act_engine_launches(self._h, 1 if reset else 0)

    def workspace_bytes(self) -> int:
        return self.lib.act_engine_workspace_bytes(self._h)

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self._h.value:
                self.lib.act_engine_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass
