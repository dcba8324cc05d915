#!/usr/bin/env python
"""Benchmark of the ACT DETRVAE hot path (BASELINE.json metric: ACT train samples/s; batched-inference chunks/s; p50).

  python bench.py --gpus N --steps K --warmup W            our CUDA engine (one process per GPU under torchrun for N>1)
  python bench.py --impl reference ...                     the reference algorithm's CPU path (oracle port) on host cores

A step = one optimisation step (zero_grad, forward+loss, backward, AdamW) of ACT sim_transfer_cube_scripted at per-GPU
batch 8, one 480x640 camera, qpos 14, chunk 100, hidden 512, ff 3200, enc 4 / dec 7, kl_weight 10 (BASELINE configs[1];
at N GPUs the global batch is 8*N, i.e. configs[2]'s batch 64 at N=8).  `--global-batch G` holds the global batch fixed
instead (configs[2] as worded: 64 over 1/2/4/8 GPUs, "scaling": "strong").  Synthetic inputs, random-init weights.
Legs of the default run, all in ONE JSON line: device-resident training step (`value`), the same step from pinned host
buffers (`e2e`), the same step through the drop-in ACTPolicy API (`e2e_policy_api`), tensor-kernel roofline (`roofline`),
AdamW HBM roofline (`roofline_hbm`), batched rollout inference on every rank (`inference`), CPU baselines.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU, H, W, K_CAM, NQ, A_DIM, S_DIM, Z_DIM = 8, 480, 640, 1, 100, 16, 14, 32
FLOP_TRAIN = {1: 115.6e9, 3: 332.0e9}        # live path (decoder layer 0 only), 3x forward, per sample
FLOP_INFER = {1: 34.9e9, 3: 107.1e9}         # per chunk


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="launch kernels directly instead of replaying a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-infer", action="store_true", help="skip the batched-inference / latency legs")
    ap.add_argument("--no-policy-api", action="store_true", help="skip the ACTPolicy drop-in leg")
    ap.add_argument("--infer-envs", type=int, default=128, help="environments per GPU in the rollout leg (1024/8)")
    ap.add_argument("--batch", type=int, default=8, help="per-GPU batch (default 8 = BASELINE configs[1]; 16 for configs[3])")
    ap.add_argument("--global-batch", type=int, default=0, help="fixed GLOBAL batch split over the ranks (configs[2]: 64); "
                    "reports scaling 'strong'")
    ap.add_argument("--cameras", type=int, default=1, help="cameras per sample (3 = BASELINE configs[3], Mobile-ALOHA)")
    ap.add_argument("--dropout", type=float, default=0.0, help="train-mode dropout probability of the timed step (reference "
                    "default 0.1; 0 = the deterministic path parity is defined on)")
    ap.add_argument("--eager-gpu", action="store_true", help="also time the reference algorithm (oracle port) as eager PyTorch "
                    "on this GPU, fp32 and bf16 autocast: the practical bar of SURVEY.md 8(d), not a reference-published number")
    return ap.parse_args()


def workload_name(b, k):
    if b == 8 and k == 1:
        tag = "ACT sim_transfer_cube_scripted training step (fwd+loss+bwd+AdamW), per-GPU batch 8, 1 camera 480x640"
        cfgs = "BASELINE configs[1]"
    elif k == 1:
        tag = f"ACT sim_insertion_scripted training step (fwd+loss+bwd+AdamW), per-GPU batch {b}, 1 camera 480x640"
        cfgs = "BASELINE configs[2] shape"
    else:
        tag = f"Mobile-ALOHA {k}-camera ACT training step (fwd+loss+bwd+AdamW), per-GPU batch {b}, {k} cameras 480x640"
        cfgs = "BASELINE configs[3] shape"
    return f"{tag}, qpos 14, chunk 100, hidden 512, ff 3200, enc 4/dec 7, kl_weight 10 ({cfgs})"


# ------------------------------------------------------------------------------------------------ CPU baselines
class CpuTrainer:
    """Reference algorithm on host cores (oracle port = torch CPU fp32 restatement of the reference modules, validated
    against the unmodified reference by oracle/make_golden.py): zero_grad + fwd + loss + backward + AdamW steps of the
    same 480x640 workload, built once and stepped repeatedly.  The live path of the reference is timed: decoder layer 0
    only (layers 1-6 are dead work upstream, detr_vae.py:245) and dropout 0 -- i.e. LESS work than the unmodified
    reference does per step, so every ratio against it is conservative."""

    def __init__(self, threads=None):
        import torch
        from oracle import act_oracle as O
        self.torch, self.O = torch, O
        self.threads = threads or os.cpu_count()
        torch.set_num_threads(self.threads)
        self.cfg = O.make_cfg(num_cameras=K_CAM)
        self.sd = O.init_state_dict(self.cfg, 0)
        self.names = O.trainable_names(self.cfg)
        self.params = [torch.nn.Parameter(self.sd[n]) for n in self.names]
        self.opt = torch.optim.AdamW(self.params, lr=1e-5, weight_decay=1e-4)
        self.leaf = dict(self.sd)
        self.leaf.update({n: p for n, p in zip(self.names, self.params)})
        self.inputs = {}

    def batch(self, b):
        if b not in self.inputs:
            self.inputs[b] = self.O.make_inputs(self.cfg, b, H, W, 0)
        return self.inputs[b]

    def step(self, batch):
        """one optimisation step at `batch` samples; returns seconds"""
        b = self.batch(batch)
        t0 = time.perf_counter()
        self.opt.zero_grad()
        out = self.O.policy_forward(self.leaf, self.cfg, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
        out["loss"].backward()
        self.opt.step()
        return time.perf_counter() - t0

    def forward_loss(self, batch):
        """BASELINE configs[0]: single forward + loss (no backward); returns seconds"""
        b = self.batch(batch)
        t0 = time.perf_counter()
        with self.torch.no_grad():
            self.O.policy_forward(self.sd, self.cfg, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
        return time.perf_counter() - t0

    def infer(self, batch):
        """the inference call policy(qpos, image) -> chunks; returns seconds"""
        b = self.batch(batch)
        t0 = time.perf_counter()
        with self.torch.inference_mode():
            self.O.policy_forward(self.sd, self.cfg, b["qpos"], b["image"])
        return time.perf_counter() - t0

    def pick_batch(self, steps, budget_s):
        """two warm-up steps at batch 2 (the first one pays allocator / thread-pool start-up and must not be the one
        that sizes the run), then the largest per-step batch (<= the workload's 8) whose `steps` steps fit the budget"""
        self.step(2)
        dt2 = min(self.step(2), self.step(2))
        for b in (8, 4, 2):
            if dt2 * (b / 2.0) * steps <= budget_s:
                return b
        return 1


def cpu_train_sample(target_s=12.0, max_steps=40):
    """cpu_baseline leg: a bounded sample (about 10-30 s of CPU work) of the bench workload on all host threads."""
    tr = CpuTrainer()
    b = tr.pick_batch(2, target_s)
    tr.step(b)                                           # untimed warm-up at the chosen batch
    n, tot = 0, 0.0
    while n < max_steps and (n < 2 or tot < target_s):
        tot += tr.step(b)
        n += 1
    base = dict(value=b * n / tot, unit="samples/s", cores=tr.threads, kind="port",
                sample=(f"{n} optimisation steps (zero_grad+fwd+loss+bwd+AdamW) at batch {b} of the 480x640 {K_CAM}-camera "
                        f"workload, {tot:.1f} s, torch {tr.torch.__version__} CPU fp32; live path only (decoder layer 0, "
                        f"dropout 0): less work than the unmodified reference"))
    # BASELINE configs[0]: single forward + loss on CPU, batch 8
    tr.forward_loss(8)
    ts = sorted(tr.forward_loss(8) for _ in range(3))
    base["config0_forward_loss"] = dict(value=8 / ts[1], unit="samples/s", ms_per_call=ts[1] * 1e3,
                                        sample="median of 3 forward+loss calls at batch 8 (BASELINE configs[0])")
    return base, tr


def cpu_infer_sample(tr, envs_sample=8, ens_envs=32, ens_steps=120):
    """CPU baselines of the rollout leg: the inference call on a bounded batch, and the temporal ensemble as the literal
    reference lines (imitate_episodes.py:403-411) looped per environment."""
    import torch
    from oracle import act_oracle as O
    tr.infer(envs_sample)
    ts = sorted(tr.infer(envs_sample) for _ in range(3))
    out = dict(value=envs_sample / ts[1], unit="chunks/s", cores=tr.threads, kind="port",
               sample=f"median of 3 policy(qpos, image) calls at batch {envs_sample} (480x640, {K_CAM} camera), torch CPU fp32")
    g = torch.Generator().manual_seed(3)
    aggs = [O.TemporalAggOracle(400, NQ, A_DIM) for _ in range(ens_envs)]
    chunks = torch.randn((ens_steps, ens_envs, 1, NQ, A_DIM), generator=g)
    t0 = time.perf_counter()
    for t in range(ens_steps):
        for e in range(ens_envs):
            aggs[e].step(t, chunks[t, e])
    dt = time.perf_counter() - t0
    out["temporal_agg"] = dict(value=ens_envs * ens_steps / dt, unit="env-steps/s", cores=1,
                               sample=f"literal imitate_episodes.py:403-411 per environment, {ens_envs} envs x {ens_steps} steps, "
                                      f"buffer [400, 500, 16] f32 each, {dt:.1f} s")
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port), all host threads.  Each timed
    step is one optimisation step at the largest batch <= 8 for which `--steps` steps end within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    global K_CAM
    K_CAM = args.cameras
    tr = CpuTrainer()
    b = tr.pick_batch(args.steps + args.warmup, 240.0)
    for _ in range(args.warmup):
        tr.step(b)
    tot = 0.0
    for _ in range(args.steps):
        tot += tr.step(b)
    value = b * args.steps / tot
    sample = (f"{args.steps} optimisation steps at batch {b} of the 480x640 {K_CAM}-camera workload, {tot:.1f} s, "
              f"torch {torch.__version__} CPU fp32; live path only (decoder layer 0, dropout 0)")
    line = dict(impl="reference", metric="ACT train samples/s", value=value, unit="samples/s", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=tot / args.steps * 1e3, higher_is_better=True,
                scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=workload_name(8, K_CAM), global_batch=8 * args.gpus, parallelism=f"dp{args.gpus}"),
                notes=dict(impl=f"reference algorithm on host CPU (oracle port), bounded sample: batch {b} per step",
                           torch=torch.__version__),
                cpu_baseline=dict(value=value, unit="samples/s", cores=tr.threads, kind="port", sample=sample),
                e2e=dict(value=value, unit="samples/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line))


def eager_gpu_leg(dev, batch, steps=5):
    """--eager-gpu: the reference algorithm (oracle port, same torch ops as the reference modules) as eager PyTorch on this
    GPU with torch defaults, fp32 and bf16 autocast.  A baseline leg only ("the practical bar"): nothing
    of the product path runs here and the product never calls this."""
    import torch
    from oracle import act_oracle as O
    cfg = O.make_cfg(num_cameras=K_CAM)
    sd = {k: v.to(dev) for k, v in O.init_state_dict(cfg, 0).items()}
    names = O.trainable_names(cfg)
    params = [torch.nn.Parameter(sd[n]) for n in names]
    opt = torch.optim.AdamW(params, lr=1e-5, weight_decay=1e-4)
    leaf = dict(sd)
    leaf.update({n: p for n, p in zip(names, params)})
    b = {k: v.to(dev) for k, v in O.make_inputs(cfg, batch, H, W, 0).items()}
    res = {}
    for tag, autocast in (("fp32", False), ("bf16_autocast", True)):
        def one():
            opt.zero_grad()
            with torch.device(dev), torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                out = O.policy_forward(leaf, cfg, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
            out["loss"].backward()
            opt.step()
        for _ in range(3):
            one()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            one()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        res[tag] = dict(ms_per_step=ms, samples_per_s=batch / ms * 1e3)
    res["note"] = (f"eager PyTorch {torch.__version__} (cuDNN/cuBLAS, torch default TF32 settings) running the reference's live op "
                   f"sequence (decoder layer 0 only, dropout 0) at batch {batch}; not a reference-published number")
    return res


def infer_legs(eng, dev, envs, steps, rank, world, peaks, lib):
    """Batched rollout inference (BASELINE configs[4]: 1024 envs over 8 GPUs = 128 per GPU; one policy query per env per
    control step + temporal_agg ensemble), run on EVERY rank over its own shard of the environments:
      device-resident : observations already in HBM (float frames), chunks/s of this rank
      e2e             : pinned uint8 frames + raw qpos -> H2D -> fused ingest -> query -> ensemble -> de-normalised
                        [E, A] float64 actions D2H, every step (the rollout loop's real traffic)
      roofline        : tensor-kernel time of one query step by differential graph timing
      latency         : batch-1 query incl. the H2D of one frame and the D2H of the chunk."""
    import torch
    from act_plus_plus_b200.engine import ActEngine
    from act_plus_plus_b200.rollout import DeviceRollout
    from act_plus_plus_b200.measure import tensor_time_by_class
    out = {}
    g = torch.Generator(device="cpu").manual_seed(7 + rank)
    big = ActEngine.sharing(eng, envs, H, W, False)
    stats = dict(qpos_mean=torch.randn(S_DIM, generator=g), qpos_std=torch.rand(S_DIM, generator=g) + 0.5,
                 action_mean=torch.randn(A_DIM, generator=g), action_std=torch.rand(A_DIM, generator=g) + 0.5)
    ro = DeviceRollout(big, envs, stats=stats)
    # ---- device-resident (float frames, two alternating sets: 2 x 472 MB at 128 envs > L2)
    imgs = [torch.rand((envs, K_CAM, 3, H, W), generator=g).to(dev) for _ in range(2)]
    qpos = torch.randn((envs, S_DIM), generator=g).to(dev)
    for i in range(3):
        ro.step(qpos, imgs[i % 2])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        act = ro.step(qpos, imgs[i % 2])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    fl = FLOP_INFER.get(K_CAM, 34.9e9 * K_CAM)
    out.update(envs_per_gpu=envs, ms_per_env_step_batch=ms, chunks_per_s=envs / ms * 1e3, tflops=fl * envs / ms / 1e9,
               action_checksum=float(act.sum()))
    del imgs
    # ---- roofline of the query step (tensor kernels only), differential graph timing, engine concurrency off
    try:
        frames = torch.rand((envs, K_CAM, 3, H, W), generator=g).to(dev)
        big.set_concurrent(False)
        tt = tensor_time_by_class(lib, lambda: big.forward(qpos, frames))
        big.set_concurrent(True)
        out["roofline"] = dict(bound="tensor", achieved=tt["tensor_tflops"], peak=peaks["tf_sust"], unit="TFLOP/s",
                               frac=tt["tensor_tflops"] / peaks["tf_sust"], traffic=None, peak_source=peaks["src"] + ", sustained bf16",
                               kernel="all tcgen05 launches of one batched query (GEMM / conv / conv3 / attention scores)",
                               timing="serial CUDA-graph replay with each kernel class skipped in turn (act_tc_set_skip_mask)",
                               tensor_ms=tt["tensor_ms"], non_tensor_ms=tt["non_tensor_ms"], serial_step_ms=tt["serial_step_ms"],
                               by_kernel_class=tt["by_class"])
        del frames
    except Exception as ex:
        out["roofline"] = dict(error=str(ex)[:200])
    # ---- end to end: pinned uint8 frames [E, K, H, W, 3] + raw qpos from the host, actions back to the host
    nb = 2
    h_img = [torch.randint(0, 256, (envs, K_CAM, H, W, 3), generator=g, dtype=torch.uint8).pin_memory() for _ in range(nb)]
    h_qpos = [torch.randn((envs, S_DIM), generator=g).pin_memory() for _ in range(nb)]
    h_act = torch.empty((envs, A_DIM), dtype=torch.float64).pin_memory()
    ro.reset()
    def e2e(n):
        ro.submit(h_qpos[0], h_img[0])
        for i in range(n):
            if i + 1 < n:
                ro.submit(h_qpos[(i + 1) % nb], h_img[(i + 1) % nb])      # next observation's H2D under this step
            a = ro.step_submitted()
            h_act.copy_(a, non_blocking=True)
            torch.cuda.current_stream().synchronize()                     # the environment needs the action to step
        return float(h_act.sum())
    e2e(3)
    torch.cuda.synchronize()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(); chk = e2e(steps); f1.record(); torch.cuda.synchronize()
    ms2 = f0.elapsed_time(f1) / steps
    out["e2e"] = dict(value=envs / ms2 * 1e3, unit="chunks/s", ms_per_env_step_batch=ms2,
                      h2d_bytes_per_step=h_img[0].numel() + h_qpos[0].numel() * 4, d2h_bytes_per_step=h_act.numel() * 8,
                      action_checksum=chk, input="pinned uint8 HWC frames + raw qpos (z-score, /255, ImageNet normalise on device)")
    del ro, big, h_img
    # ---- batch-1 query latency: H2D of one uint8 frame + query + D2H of the chunk, host wall clock, synchronised
    one = ActEngine.sharing(eng, 1, H, W, False)
    h1 = torch.randint(0, 256, (1, K_CAM, H, W, 3), generator=g, dtype=torch.uint8).pin_memory()
    hq = torch.randn((1, S_DIM), generator=g).pin_memory()
    d1, dq = torch.empty_like(h1, device=dev), torch.empty_like(hq, device=dev)
    hc = torch.empty((1, NQ, A_DIM)).pin_memory()
    lat, lat_dev, lat_direct = [], [], []
    for i in range(48):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        d1.copy_(h1, non_blocking=True); dq.copy_(hq, non_blocking=True)
        hc.copy_(one.infer_graphed(dq, d1), non_blocking=True)          # CUDA-graph replay from the 3rd call on
        torch.cuda.synchronize()
        lat.append((time.perf_counter() - t0) * 1e3)
        t0 = time.perf_counter()
        one.infer_graphed(dq, d1)
        torch.cuda.synchronize()
        lat_dev.append((time.perf_counter() - t0) * 1e3)
        t0 = time.perf_counter()
        d1.copy_(h1, non_blocking=True); dq.copy_(hq, non_blocking=True)
        hc.copy_(one.forward(dq, d1)["a_hat"], non_blocking=True)       # ~100 direct launches
        torch.cuda.synchronize()
        lat_direct.append((time.perf_counter() - t0) * 1e3)
    lat, lat_dev, lat_direct = sorted(lat[8:]), sorted(lat_dev[8:]), sorted(lat_direct[8:])
    out.update(p50_latency_ms_b1=lat[len(lat) // 2], p90_latency_ms_b1=lat[int(len(lat) * 0.9)],
               p50_latency_ms_b1_device_resident=lat_dev[len(lat_dev) // 2],
               p50_latency_ms_b1_direct_launch=lat_direct[len(lat_direct) // 2],
               latency_note="host wall clock of H2D(uint8 frame + qpos) + query (CUDA-graph replay) + D2H(chunk), synchronised "
                            "per call; direct_launch = the same without the graph")
    return out


def policy_api_leg(dev, B, steps, warmup, sd_src, host, fused):
    """The step as a user of the reference would write it (imitate_episodes.py:596-607): ACTPolicy(...) ->
    policy(qpos, image, actions, is_pad)['loss'].backward() -> optimizer.step() -> optimizer.zero_grad(), fed from pinned
    host batches (H2D every step) with a loss read-back every step (an asynchronous copy into pinned memory that is read one
    step later -- the reference loop itself only detaches the loss dict per step, imitate_episodes.py:608, so the host is
    never made to wait for the step it has just enqueued).  fused = args_override['fused_step']."""
    import torch
    from act_plus_plus_b200.policy import ACTPolicy
    cfg = {'lr': 1e-5, 'num_queries': NQ, 'kl_weight': 10, 'hidden_dim': 512, 'dim_feedforward': 3200, 'lr_backbone': 1e-5,
           'backbone': 'resnet18', 'enc_layers': 4, 'dec_layers': 7, 'nheads': 8, 'camera_names': [f'cam{k}' for k in range(K_CAM)],
           'vq': False, 'vq_class': None, 'vq_dim': None, 'action_dim': A_DIM, 'no_encoder': False, 'state_dim': S_DIM,
           'dropout': 0.0, 'fused_step': fused, 'seed': 0}
    import contextlib
    import warnings
    with warnings.catch_warnings(), contextlib.redirect_stdout(sys.stderr):   # (the reference prints "KL Weight ..." on construction)
        warnings.simplefilter("ignore")                  # (no ImageNet checkpoint on the box: random-init backbone, as everywhere here)
        policy = ACTPolicy(cfg)
    policy.model._master.flat_params.copy_(sd_src)       # the same random-init weights as the engine-level legs
    policy.model._master.mark_params_changed(); policy.model._master.refresh()
    policy.train()
    opt = policy.configure_optimizers()
    from act_plus_plus_b200.train import InputPrefetcher
    NB = len(host)
    feed = [{k: v for k, v in hb.items() if k != "eps"} for hb in host]      # the policy draws its own reparametrisation noise
    pf = InputPrefetcher(feed[0], dev, depth=2)          # the package's loader-side helper: H2D of batch i+1 under step i

    loss_pin = [torch.zeros(1).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]

    def run(n):
        last = 0.0
        pf.submit(feed[0])
        for i in range(n):
            if i + 1 < n:
                pf.submit(feed[(i + 1) % NB])
            d = pf.next()
            out = policy(d["qpos"], d["image"], d["actions"], d["is_pad"])
            pf.consumed()
            out["loss"].backward()
            opt.step()
            opt.zero_grad()
            loss_pin[i & 1].copy_(out["loss"].detach().reshape(1), non_blocking=True)     # D2H of this step's loss ...
            loss_ev[i & 1].record()
            if i > 0:                                                                    # ... read once the NEXT step is enqueued
                loss_ev[(i - 1) & 1].synchronize()
                last = float(loss_pin[(i - 1) & 1][0])
        loss_ev[(n - 1) & 1].synchronize()
        return float(loss_pin[(n - 1) & 1][0])

    run(max(4, warmup))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); last = run(steps); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    graphs = bool(policy.model._graphs) and all(g.fwd is not None and g.bwd is not None for g in policy.model._graphs.values())
    del policy, opt
    torch.cuda.empty_cache()
    return dict(value=B / ms * 1e3, unit="samples/s", ms_per_step=ms, last_loss=last, cuda_graphs=graphs)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from act_plus_plus_b200.engine import ActEngine, R_BACKBONE, R_CVAE, R_DEAD, R_MAIN
    from act_plus_plus_b200 import _lib
    from act_plus_plus_b200.measure import ClockSampler, measured_peaks, tensor_time_by_class
    from act_plus_plus_b200.train import GraphedTrainStep, InputPrefetcher, train_step
    from act_plus_plus_b200.policy import _init_like_reference

    global B_PER_GPU, K_CAM
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    strong = args.global_batch > 0
    if strong:
        assert args.global_batch % world == 0, "--global-batch must be divisible by the number of ranks"
    B_PER_GPU, K_CAM = (args.global_batch // world if strong else args.batch), args.cameras
    flop_train = FLOP_TRAIN.get(K_CAM, 115.6e9 * K_CAM)
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
        group = dist.group.WORLD
    dev = torch.device(f"cuda:{local}")
    B = B_PER_GPU
    lib = _lib.lib()
    peaks = measured_peaks()
    eng = ActEngine(batch_max=B, num_cameras=K_CAM, img_h=H, img_w=W, train=True, device=dev)
    _init_like_reference(eng, seed=0)                 # random-init weights of the reference architecture, same on every rank
    if args.dropout > 0:
        eng.set_dropout(args.dropout, seed=1234 + 1000003 * rank)
    # Gradient exchange: fused into the optimiser kernel over NVLink peer memory (csrc/dp_p2p.cu); ACT_DP_P2P=0 selects
    # NCCL all-reduce + AdamW per range, which is also the fallback when a rank cannot map its peers' buffers.
    p2p = world > 1 and os.environ.get("ACT_DP_P2P", "1") == "1"
    if p2p:
        try:
            eng.dp_connect(group)                    # collective: raises on every rank if any rank failed
        except RuntimeError as ex:
            print(f"[bench] rank {rank}: peer-memory exchange unavailable ({str(ex)[:160]}); using NCCL", file=sys.stderr)
            p2p = False
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    NB = 4                                            # rotate input batches (4 x 29.5 MB images at B=8) so nothing is cached
    host = []
    for i in range(NB):
        n_valid = torch.randint(1, NQ + 1, (B,), generator=g)
        host.append(dict(image=torch.rand((B, K_CAM, 3, H, W), generator=g).pin_memory(),
                         qpos=torch.randn((B, S_DIM), generator=g).pin_memory(),
                         actions=torch.randn((B, NQ, A_DIM), generator=g).pin_memory(),
                         is_pad=(torch.arange(NQ)[None] >= n_valid[:, None]).pin_memory(),
                         eps=torch.randn((B, Z_DIM), generator=g).pin_memory()))
    devb = [{k: v.to(dev) for k, v in hb.items()} for hb in host]

    def step_direct(b):
        return train_step(eng, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"], group)

    # ---------------- warm-up + CUDA graph of the whole step (owned by the package: train.GraphedTrainStep)
    for i in range(max(3, args.warmup)):
        step_direct(devb[i % NB])
    torch.cuda.synchronize()
    gs = None
    if not args.no_graph and (world == 1 or os.environ.get("ACT_DP_GRAPH", "1") == "1"):
        gs = GraphedTrainStep(eng, devb[0], group, warmup=1, stream_priority=int(os.environ.get("ACT_STREAM_PRIO", "-1")))
        if gs.graph is None:
            gs = None
    eng.launches(reset=True)
    step_direct(devb[0])
    launches_per_step = eng.launches(reset=True)
    torch.cuda.synchronize()

    def step(i):
        return gs.step(devb[i % NB]) if gs is not None else step_direct(devb[i % NB])

    for i in range(args.warmup):
        step(i)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- timed region 1: device-resident inputs
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        losses = step(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # ---------------- timed region 2, end to end: pinned host inputs -> H2D -> step -> loss D2H, every step
    h2d = sum(v.numel() * v.element_size() for v in host[0].values())
    pf = InputPrefetcher(host[0], dev, depth=2)                      # staging buffers / copy stream: created once
    loss_pin = [torch.zeros(3).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]

    def e2e_steps(n):
        # the step's losses go to pinned host memory EVERY step; the host reads them once the next step is enqueued (the
        # reference loop only detaches the loss dict per step, imitate_episodes.py:608: nothing makes it wait for the step
        # it has just launched), so H2D of step i+1, the compute of step i and the read of step i-1 overlap
        pf.submit(host[0])
        last = 0.0
        for i in range(n):
            if i + 1 < n:
                pf.submit(host[(i + 1) % NB])                        # H2D of the next step's inputs runs under this step
            db = pf.next()
            ls = gs.step(db) if gs is not None else step_direct(db)
            pf.consumed()
            loss_pin[i & 1].copy_(ls, non_blocking=True)             # device->host read of the step's losses
            loss_ev[i & 1].record()
            if i > 0:
                loss_ev[(i - 1) & 1].synchronize()
                last = float(loss_pin[(i - 1) & 1][2])
        loss_ev[(n - 1) & 1].synchronize()
        return float(loss_pin[(n - 1) & 1][2]) if n > 0 else last

    e2e_steps(max(3, args.warmup))                                   # untimed warm-up of the host-driven loop
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    loss_host = e2e_steps(args.steps)
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    if p2p and eng.dp_status() != 0:
        raise RuntimeError("peer-memory data-parallel step: a flag barrier timed out")
    n_samples = B * world * args.steps
    value = n_samples / (ms / 1e3)
    line = dict(metric="ACT train samples/s", value=value, unit="samples/s", n_gpus=world, steps=args.steps,
                warmup=args.warmup, ms_per_step=ms / args.steps, higher_is_better=True,
                scaling="strong" if strong else "weak", vs_baseline=None, dtype="bf16", data="synthetic",
                config=dict(workload=workload_name(B, K_CAM), global_batch=B * world, parallelism=f"dp{world}"),
                notes=dict(cuda_graph=gs is not None,
                           grad_exchange=("none" if world == 1 else
                                          ("NVLS reduce+AdamW+broadcast kernel (multimem.ld_reduce / multimem.st through the NVSwitch, csrc/dp_p2p.cu)"
                                           if getattr(eng, "_dp_nvls", False) else
                                           "peer-memory reduce+AdamW+broadcast kernel over NVLink (csrc/dp_p2p.cu)"
                                           + (f"; NVLS unavailable: {getattr(eng, '_dp_nvls_error', '')[:80]}" if getattr(eng, "_dp_nvls_error", None) else ""))
                                          if p2p else "NCCL all-reduce per range, overlapped with backward (ACT_DP_P2P=0)"),
                           l2="4 rotating input batches; parameters+Adam state (1.3 GB) exceed the 126 MB L2 every step",
                           dropout=args.dropout, last_loss=loss_host,
                           flop_per_sample=flop_train, whole_step_tflops=flop_train * B * world / (ms / args.steps / 1e3) / 1e12),
                e2e=dict(value=n_samples / (ms_e2e / 1e3), unit="samples/s", h2d_bytes_per_step=h2d, d2h_bytes_per_step=12),
                gpu_launches=launches_per_step * args.steps, clocks=clk)
    # ---------------- roofline leg: device time of the tensor-core kernels of ONE step by differential graph timing
    # (engine concurrency off = one kernel at a time in program order; serial CUDA-graph replay of the step with each
    # kernel class skipped in turn -- no per-launch event pairs in the measurement; the steps run this way produce garbage,
    # so parameters and Adam state are saved and restored around it)
    try:
        torch.cuda.synchronize()
        eng.set_concurrent(False)
        sb = devb[0]

        def save():
            return [x.clone() for x in (eng.flat_params, eng.exp_avg, eng.exp_avg_sq)], eng._shared["step"]

        def restore(st):
            for dst, src in zip((eng.flat_params, eng.exp_avg, eng.exp_avg_sq), st[0]):
                dst.copy_(src)
            eng._shared["step"] = st[1]
            eng.mark_params_changed(); eng.refresh()

        tt = tensor_time_by_class(lib, lambda: train_step(eng, sb["qpos"], sb["image"], sb["actions"], sb["is_pad"], sb["eps"],
                                                          None, advance_on_device=True), (save, restore))
        eng.set_concurrent(True)
        line["roofline"] = dict(
            bound="tensor", achieved=tt["tensor_tflops"], peak=peaks["tf_sust"], unit="TFLOP/s", frac=tt["tensor_tflops"] / peaks["tf_sust"],
            traffic=TC_DRAM_WARM.get((B, K_CAM)), traffic_source=TC_DRAM_SOURCE, peak_source=peaks["src"] + ", sustained bf16",
            kernel="all tcgen05 launches of one step (tc_kernel GEMM / conv / wgrad, conv3_kernel, attn_scores kernels)",
            timing="T(serial step) - T(serial step without the class), both CUDA-graph replays with engine concurrency off "
                   "(act_tc_set_skip_mask); algorithmic FLOPs = 2*M*N*K per launch (convs: real taps only)",
            tensor_ms_per_step=tt["tensor_ms"], non_tensor_ms_per_step=tt["non_tensor_ms"], serial_step_ms=tt["serial_step_ms"],
            tensor_gflop_per_step=tt["tensor_gflop"], by_kernel_class=tt["by_class"],
            overlap_factor=tt["serial_step_ms"] / (ms / args.steps),
            whole_step_tflops=flop_train * B / (ms / args.steps / 1e3) / 1e12)
    except Exception as ex:
        eng.set_concurrent(True)
        lib.act_tc_set_skip_mask(0)
        line["roofline"] = dict(error=str(ex)[:300])
    # ---------------- HBM-bound side of the step: the fused AdamW (+ bf16 shadow, + conv re-pack), timed alone
    try:
        nb = {r: eng.region(r)[1] for r in (R_MAIN, R_CVAE, R_BACKBONE, R_DEAD)}
        adam_bytes = 30.0 * (nb[R_MAIN] + nb[R_CVAE]) + 28.0 * nb[R_BACKBONE] + 8.0 * nb[R_DEAD]
        ts = []
        for _ in range(5):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(); eng.adamw_step(); a1.record(); torch.cuda.synchronize()
            ts.append(a0.elapsed_time(a1))
        t_adam = sorted(ts)[2]
        line["roofline_hbm"] = dict(kernel="adamw_kernel x4 regions + conv re-pack (one optimiser step, timed alone)", bound="hbm",
                                    achieved=adam_bytes / (t_adam / 1e3) / 1e9, peak=peaks["hbm"], unit="GB/s",
                                    frac=adam_bytes / (t_adam / 1e3) / 1e9 / peaks["hbm"], ms=t_adam,
                                    algorithmic_bytes=adam_bytes, peak_source=peaks["src"] + ", copy bandwidth")
    except Exception as ex:
        line["roofline_hbm"] = dict(error=str(ex)[:200])
    # ---------------- the same step through the drop-in API (rank 0's shard only; single-GPU semantics)
    if not args.no_policy_api and world == 1:
        try:
            line["e2e_policy_api"] = policy_api_leg(dev, B, args.steps, args.warmup, eng.flat_params, host, fused=True)
            line["e2e_policy_api"]["mode"] = ("ACTPolicy(fused_step=True): loss.backward() deferred into optimizer.step(), step replayed "
                                              "from two CUDA graphs; pinned host batch -> H2D, loss D2H every step (pinned, read one step later)")
            line["e2e_policy_api_unfused"] = policy_api_leg(dev, B, max(5, args.steps // 2), 3, eng.flat_params, host, fused=False)
            line["e2e_policy_api_unfused"]["mode"] = "ACTPolicy default: backward at loss.backward(), AdamW at optimizer.step(), direct launches"
        except Exception as ex:
            line["e2e_policy_api"] = dict(error=str(ex)[:300])
    if args.eager_gpu and rank == 0:
        try:
            line["eager_gpu_baseline"] = eager_gpu_leg(dev, B)
        except Exception as ex:
            line["eager_gpu_baseline"] = dict(error=str(ex)[:300])
    # ---------------- batched rollout inference: every rank runs its shard of the environments
    if not args.no_infer:
        try:
            inf = infer_legs(eng, dev, args.infer_envs, max(3, min(args.steps, 10)), rank, world, peaks, lib)
            if world > 1:
                # global figure = all environments / slowest rank's step; actions gathered on the host side
                tt2 = torch.tensor([inf["ms_per_env_step_batch"], inf["e2e"]["ms_per_env_step_batch"]], device=dev, dtype=torch.float64)
                dist.all_reduce(tt2, op=dist.ReduceOp.MAX)
                from act_plus_plus_b200.dp import gather_env_actions
                acts = gather_env_actions(torch.zeros((args.infer_envs, A_DIM), dtype=torch.float64, device=dev),
                                          [args.infer_envs] * world, group)
                inf.update(n_gpus=world, envs_global=args.infer_envs * world, gathered_actions_shape=list(acts.shape),
                           chunks_per_s_global=args.infer_envs * world / float(tt2[0]) * 1e3,
                           e2e_chunks_per_s_global=args.infer_envs * world / float(tt2[1]) * 1e3)
            else:
                inf.update(n_gpus=1, envs_global=args.infer_envs, chunks_per_s_global=inf["chunks_per_s"],
                           e2e_chunks_per_s_global=inf["e2e"]["value"])
            line["inference"] = inf
        except Exception as ex:
            line["inference"] = dict(error=str(ex)[:300])
    if rank != 0:
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)       # a captured graph that contains NCCL kernels keeps the communicator busy on teardown
    if not args.no_cpu_baseline and world == 1:
        try:
            base, tr = cpu_train_sample()
            line["cpu_baseline"] = base
            if "inference" in line and "error" not in line["inference"]:
                line["inference"]["cpu_baseline"] = cpu_infer_sample(tr)
        except Exception as ex:
            line["cpu_baseline"] = dict(value=None, unit="samples/s", cores=os.cpu_count(), kind="port", sample=f"failed: {ex}")
    print(json.dumps(line))
    if world > 1:
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)


# DRAM traffic of the tensor-core kernels measured inside a WARM step (ncu --cache-control none, second step of
# tools/profile_step.py): mean dram__bytes_read.sum + dram__bytes_write.sum per tcgen05 launch; keyed by (batch, cameras)
TC_DRAM_WARM = {(8, 1): 7974248.0}     # bytes per tcgen05 launch (277 launches, 2208.9 MB per step; algorithmic 3149.6 MB)
TC_DRAM_SOURCE = "profiles/r02_tc_dram_warm.md (ncu --cache-control none, warm step)"


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
