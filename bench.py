#!/usr/bin/env python
"""Benchmark of the ACT DETRVAE hot path (BASELINE.json metric: ACT train samples/s).

  python bench.py --gpus N --steps K --warmup W            our CUDA engine (one process per GPU under torchrun for N>1)
  python bench.py --impl reference ...                     the reference algorithm's CPU path (oracle port) on host cores

A step = one optimisation step (zero_grad, forward+loss, backward, AdamW) of ACT sim_transfer_cube_scripted at per-GPU
batch 8, one 480x640 camera, qpos 14, chunk 100, hidden 512, ff 3200, enc 4 / dec 7, kl_weight 10 (BASELINE configs[1];
at N GPUs the global batch is 8*N, i.e. configs[2]'s batch 64 at N=8).  Synthetic inputs, random-init weights.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU, H, W, K_CAM, NQ, A_DIM, S_DIM, Z_DIM = 8, 480, 640, 1, 100, 16, 14, 32
# mean DRAM bytes per tcgen05 launch of the step (ncu, profiles/r01_tc_dram_per_step.md): 2787.8 MB / 277 launches
TC_DRAM_BYTES_PER_LAUNCH = 2787.8e6 / 277
FLOP_PER_SAMPLE_TRAIN = 115.6e9        # live path (decoder layer 0 only), 3x forward
ADAMW_BYTES = 83.9e6 * 28


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="launch kernels directly instead of replaying a CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-infer", action="store_true", help="skip the batched-inference / latency legs")
    ap.add_argument("--infer-envs", type=int, default=128, help="environments per GPU in the rollout leg (1024/8)")
    ap.add_argument("--batch", type=int, default=8, help="per-GPU batch (default 8 = BASELINE configs[1]; 64/N for configs[2] "
                    "at fixed global batch, 16 for configs[3])")
    ap.add_argument("--cameras", type=int, default=1, help="cameras per sample (3 = BASELINE configs[3], Mobile-ALOHA)")
    ap.add_argument("--eager-gpu", action="store_true", help="also time the reference algorithm (oracle port) as eager PyTorch "
                    "on this GPU, fp32 and bf16 autocast: the practical bar of SURVEY.md 8(d), not a reference-published number")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.rows, self.stop_flag = [], False
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        self.cmd = ["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(index), "-lms", "100"]
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(self.cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))


def workload_name(b, k):
    if b == 8 and k == 1:
        tag = "ACT sim_transfer_cube_scripted training step (fwd+loss+bwd+AdamW), per-GPU batch 8, 1 camera 480x640"
        cfgs = "BASELINE configs[1]"
    elif k == 1:
        tag = f"ACT sim_insertion_scripted training step (fwd+loss+bwd+AdamW), per-GPU batch {b}, 1 camera 480x640"
        cfgs = "BASELINE configs[2] shape, non-default --batch"
    else:
        tag = f"Mobile-ALOHA {k}-camera ACT training step (fwd+loss+bwd+AdamW), per-GPU batch {b}, {k} cameras 480x640"
        cfgs = "BASELINE configs[3] shape, non-default --cameras/--batch"
    return f"{tag}, qpos 14, chunk 100, hidden 512, ff 3200, enc 4/dec 7, kl_weight 10 ({cfgs})"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


class CpuTrainer:
    """Reference algorithm on host cores (oracle port = torch CPU fp32 restatement of the reference modules, validated
    against the unmodified reference by oracle/make_golden.py): zero_grad + fwd + loss + backward + AdamW steps of the
    same 480x640 workload, built once and stepped repeatedly."""

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

    def step(self, batch):
        """one optimisation step at `batch` samples; returns seconds"""
        O = self.O
        if batch not in self.inputs:
            self.inputs[batch] = O.make_inputs(self.cfg, batch, H, W, 0)
        b = self.inputs[batch]
        t0 = time.perf_counter()
        self.opt.zero_grad()
        out = O.policy_forward(self.leaf, self.cfg, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
        out["loss"].backward()
        self.opt.step()
        return time.perf_counter() - t0

    def pick_batch(self, steps, budget_s):
        """warm up with one batch-2 step and pick the largest per-step batch (<= the workload's 8) whose `steps`
        steps fit the time budget"""
        dt2 = self.step(2)
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
    return b * n / tot, tr.threads, (f"{n} optimisation steps (zero_grad+fwd+loss+bwd+AdamW) at batch {b} of the 480x640 "
                                     f"1-camera workload, {tot:.1f} s, torch {tr.torch.__version__} CPU fp32")


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port), all host threads.  Each timed
    step is one optimisation step at the largest batch <= 8 for which `--steps` steps end within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    tr = CpuTrainer()
    b = tr.pick_batch(args.steps + args.warmup, 150.0)
    for _ in range(args.warmup):
        tr.step(b)
    tot = 0.0
    for _ in range(args.steps):
        tot += tr.step(b)
    value = b * args.steps / tot
    sample = (f"{args.steps} optimisation steps at batch {b} of the 480x640 1-camera workload, {tot:.1f} s, "
              f"torch {torch.__version__} CPU fp32")
    line = dict(impl="reference", metric="ACT train samples/s", value=value, unit="samples/s", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=tot / args.steps * 1e3, higher_is_better=True,
                scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload="ACT sim_transfer_cube_scripted training step (fwd+loss+bwd+AdamW), per-GPU batch 8, "
                            "1 camera 480x640, qpos 14, chunk 100, hidden 512, ff 3200, enc 4/dec 7, kl_weight 10 (BASELINE configs[1])",
                            impl_note=f"reference algorithm on host CPU (oracle port), bounded sample: batch {b} per step",
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
    res["note"] = (f"eager PyTorch {torch.__version__} (cuDNN/cuBLAS, torch default TF32 settings) running the reference's op "
                   f"sequence incl. all 7 decoder layers at batch {batch}; not a reference-published number")
    return res


def infer_legs(eng, dev, envs, steps):
    """Batched rollout inference (BASELINE configs[4]: 1024 envs over 8 GPUs = 128 per GPU, one policy query per env per
    step + temporal_agg ensemble) -> chunks/s on this GPU; and the B=1 query latency (p50 over synchronised calls)."""
    import torch
    from act_plus_plus_b200.engine import ActEngine
    from act_plus_plus_b200.rollout import TemporalEnsembler
    out = {}
    g = torch.Generator(device="cpu").manual_seed(7)
    big = ActEngine.sharing(eng, envs, H, W, False)
    ens = TemporalEnsembler(envs, NQ, A_DIM, device=dev)
    imgs = [torch.rand((envs, K_CAM, 3, H, W), generator=g).to(dev) for _ in range(2)]     # 2 x 472 MB > L2
    qpos = torch.randn((envs, S_DIM), generator=g).to(dev)
    for i in range(3):
        ens.step(big.forward(qpos, imgs[i % 2])["a_hat"], i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        act = ens.step(big.forward(qpos, imgs[i % 2])["a_hat"], 3 + i)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    out.update(envs_per_gpu=envs, ms_per_env_step_batch=ms, chunks_per_s=envs / ms * 1e3,
               tflops={1: 34.9e9, 3: 107.1e9}.get(K_CAM, 34.9e9 * K_CAM) * envs / ms / 1e9, action_checksum=float(act.sum()))
    del big, imgs
    one = ActEngine.sharing(eng, 1, H, W, False)
    img1 = torch.rand((1, K_CAM, 3, H, W), generator=g).to(dev)
    q1 = qpos[:1].contiguous()
    lat = []
    for i in range(30):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        a = one.forward(q1, img1)["a_hat"]
        torch.cuda.synchronize()
        lat.append((time.perf_counter() - t0) * 1e3)
    lat = sorted(lat[5:])
    out.update(p50_latency_ms_b1=lat[len(lat) // 2], p90_latency_ms_b1=lat[int(len(lat) * 0.9)])
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from act_plus_plus_b200.engine import ActEngine
    from act_plus_plus_b200 import _lib
    from act_plus_plus_b200.train import train_step
    import ctypes as C

    global B_PER_GPU, K_CAM, FLOP_PER_SAMPLE_TRAIN
    B_PER_GPU, K_CAM = args.batch, args.cameras
    FLOP_PER_SAMPLE_TRAIN = {1: 115.6e9, 3: 332.0e9}.get(K_CAM, 115.6e9 * K_CAM)      # SURVEY.md 8(d), live path
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
        group = dist.group.WORLD
    dev = torch.device(f"cuda:{local}")
    B = B_PER_GPU
    eng = ActEngine(batch_max=B, num_cameras=K_CAM, img_h=H, img_w=W, train=True, device=dev)
    # random-init weights of the reference architecture, identical on every rank
    from act_plus_plus_b200.policy import _init_like_reference
    _init_like_reference(eng, seed=0)
    # Gradient exchange: by default fused into the optimiser kernel over NVLink peer memory (csrc/dp_p2p.cu; checked
    # against the full-batch step and measured beside the NCCL path at N = 2, 4, 8: profiles/r01_dp_p2p_vs_nccl.md).
    # ACT_DP_P2P=0 selects NCCL all-reduce + AdamW per range; it is also the fallback when any rank cannot map its
    # peers' buffers (CUDA IPC unavailable).
    p2p = world > 1 and os.environ.get("ACT_DP_P2P", "1") == "1"
    if p2p:
        try:
            eng.dp_connect(group)                    # collective: raises on every rank if any rank failed
        except RuntimeError as ex:
            print(f"[bench] rank {rank}: peer-memory exchange unavailable ({str(ex)[:160]}); using NCCL", file=sys.stderr)
            p2p = False
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    NB = 4                                            # rotate input batches (4 x 29.5 MB images) so nothing is cached
    host = []
    for i in range(NB):
        n_valid = torch.randint(1, NQ + 1, (B,), generator=g)
        host.append(dict(image=torch.rand((B, K_CAM, 3, H, W), generator=g).pin_memory(),
                         qpos=torch.randn((B, S_DIM), generator=g).pin_memory(),
                         actions=torch.randn((B, NQ, A_DIM), generator=g).pin_memory(),
                         is_pad=(torch.arange(NQ)[None] >= n_valid[:, None]).pin_memory(),
                         eps=torch.randn((B, Z_DIM), generator=g).pin_memory()))
    devb = [{k: v.to(dev) for k, v in hb.items()} for hb in host]
    stat = {k: torch.empty_like(v) for k, v in devb[0].items()}        # static inputs for graph replay

    def step_direct(b):
        return train_step(eng, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"], group)

    # ---------------- warm-up (also compiles nothing: all kernels are prebuilt) + optional CUDA graph capture
    for i in range(max(3, args.warmup)):
        losses = step_direct(devb[i % NB])
    torch.cuda.synchronize()
    use_graph = not args.no_graph and (world == 1 or os.environ.get("ACT_DP_GRAPH", "1") == "1")
    graph = None
    if use_graph:
        try:
            for k in stat:
                stat[k].copy_(devb[0][k])
            s = torch.cuda.Stream(priority=int(os.environ.get('ACT_STREAM_PRIO', '-1')))
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                step_direct(stat)
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                g_losses = step_direct(stat)
            torch.cuda.synchronize()
        except Exception as ex:                                     # capture is an optimisation, never a requirement
            print(f"[bench] CUDA graph capture failed ({ex}); launching directly", file=sys.stderr)
            graph = None
    launches_per_step = 0
    eng.launches(reset=True)
    step_direct(devb[0])
    launches_per_step = eng.launches(reset=True)
    torch.cuda.synchronize()

    def step(i):
        if graph is not None:
            for k in stat:
                stat[k].copy_(devb[i % NB][k], non_blocking=True)      # device->device, part of the step
            graph.replay()
            return g_losses
        return step_direct(devb[i % NB])

    for i in range(args.warmup):
        step(i)
    # ---------------- timed region: device-resident inputs
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
    # ---------------- end-to-end: pinned host inputs -> H2D -> step -> loss D2H, every step
    h2d = sum(v.numel() * v.element_size() for v in host[0].values())
    from act_plus_plus_b200.train import InputPrefetcher
    pf = InputPrefetcher(host[0], dev, depth=2)                      # staging buffers / copy stream: created once

    def e2e_steps(n):
        last = 0.0
        pf.submit(host[0])
        for i in range(n):
            if i + 1 < n:
                pf.submit(host[(i + 1) % NB])                        # H2D of the next step's inputs runs under this step
            db = pf.next()
            if graph is not None:
                for k in stat:
                    stat[k].copy_(db[k], non_blocking=True)
                pf.consumed()
                graph.replay()
                ls = g_losses
            else:
                ls = step_direct(db)
                pf.consumed()
            last = float(ls[2].item())                               # device->host read of the step's loss
        return last

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
    # ---------------- roofline leg: CUDA-event time of every tcgen05 kernel over two more (direct-launch) steps
    lib = _lib.lib()
    torch.cuda.synchronize()
    eng.set_concurrent(False)                       # one kernel at a time, program order, on this stream
    step_direct(devb[0])
    torch.cuda.synchronize()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(200)]
    for ev in evs:
        ev.record()
    torch.cuda.synchronize()
    ev_pair_ms = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(0, 200, 2))[50]   # median empty event pair
    dump_path = os.path.join(ROOT, "gpurun_out", f"tc_launches_rank{rank}.csv")
    try:
        os.makedirs(os.path.dirname(dump_path), exist_ok=True)
        lib.act_tc_profile_set_dump(dump_path.encode())
    except Exception:
        dump_path = None
    lib.act_tc_profile_begin()
    PSTEPS = 2
    for i in range(PSTEPS):
        step_direct(devb[i % NB])
    torch.cuda.synchronize()
    tot_ms, tot_fl, nl, mx = C.c_double(), C.c_double(), C.c_int(), C.c_double()
    _lib.check(lib.act_tc_profile_end(C.byref(tot_ms), C.byref(tot_fl), C.byref(nl), C.byref(mx)), "tc_profile_end")
    lib.act_tc_profile_set_dump(b"")
    eng.set_concurrent(True)
    by_class = {}
    if dump_path and os.path.exists(dump_path):
        import csv
        for row in csv.DictReader(open(dump_path)):
            c = row["label"].split(" ")[0]
            d = by_class.setdefault(c, dict(launches=0, ms=0.0, gflop=0.0))
            d["launches"] += 1; d["ms"] += float(row["ms"]); d["gflop"] += float(row["gflop"])
        for d in by_class.values():
            net = max(1e-9, d["ms"] - d["launches"] * ev_pair_ms)
            d.update(launches=d["launches"] // PSTEPS, ms=d["ms"] / PSTEPS, gflop=d["gflop"] / PSTEPS,
                     tflops=d["gflop"] / d["ms"], tflops_net_of_event_cost=d["gflop"] / net)
    if rank != 0:
        # a captured graph that contains NCCL kernels keeps the communicator busy on teardown: leave without it
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)
    peaks = measured_peaks()
    n_samples = B * world * args.steps
    value = n_samples / (ms / 1e3)
    tc_ms_step = tot_ms.value / PSTEPS
    tc_tflops = (tot_fl.value / PSTEPS) / (tc_ms_step / 1e3) / 1e12
    line = dict(metric="ACT train samples/s", value=value, unit="samples/s", n_gpus=world, steps=args.steps,
                warmup=args.warmup, ms_per_step=ms / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="bf16", data="synthetic",
                config=dict(workload=workload_name(B, K_CAM),
                            global_batch=B * world, parallelism=f"dp{world}", cuda_graph=graph is not None,
                            grad_exchange=("none" if world == 1 else "peer-memory reduce+AdamW+broadcast kernel over NVLink (csrc/dp_p2p.cu)"
                                           if p2p else "NCCL all-reduce per range, overlapped with backward (ACT_DP_P2P=0)"),
                            l2="4 rotating input batches; parameters+Adam state (1.3 GB) exceed the 126 MB L2 every step",
                            dropout="0 (reference validation path; see DESIGN.md)", last_loss=loss_host),
                e2e=dict(value=n_samples / (ms_e2e / 1e3), unit="samples/s", h2d_bytes_per_step=h2d, d2h_bytes_per_step=4),
                gpu_launches=launches_per_step * args.steps,
                roofline=dict(bound="tensor", achieved=tc_tflops, peak=peaks["tf_sust"], unit="TFLOP/s",
                              frac=tc_tflops / peaks["tf_sust"], traffic=TC_DRAM_BYTES_PER_LAUNCH if (B == 8 and K_CAM == 1) else None, peak_source=peaks["src"] + " (sustained bf16)",
                              traffic_source="profiles/r01_tc_dram_per_step.md: ncu dram__bytes_read+write summed over the 277 tcgen05 launches "
                                             "of one step (cold cache per launch) / 277",
                              kernel="tc_kernel<*> (all tcgen05 GEMM/conv/wgrad launches of one step)",
                              launches_per_step=nl.value // PSTEPS, tc_ms_per_step=tc_ms_step,
                              timing="CUDA events around every launch, serial direct-launch replay of the step "
                                     "(ACT engine concurrency off); includes the event-pair cost below",
                              event_pair_us=ev_pair_ms * 1e3,
                              achieved_net_of_event_cost=(tot_fl.value / PSTEPS) / max(1e-9, (tc_ms_step - (nl.value // PSTEPS) * ev_pair_ms) / 1e3) / 1e12,
                              by_kernel_class=by_class, longest_launch_ms=mx.value,
                              whole_step_tflops=FLOP_PER_SAMPLE_TRAIN * B / (ms / args.steps / 1e3) / 1e12),
                clocks=clk)
    # ---------------- HBM-bound side of the step: the fused AdamW (+ bf16 shadow, + conv re-pack), timed alone
    try:
        from act_plus_plus_b200.engine import R_BACKBONE, R_CVAE, R_DEAD, R_MAIN
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
                                    algorithmic_bytes=adam_bytes, peak_source=peaks["src"] + " (copy bandwidth)")
    except Exception as ex:
        line["roofline_hbm"] = dict(error=str(ex)[:200])
    if args.eager_gpu:
        try:
            line["eager_gpu_baseline"] = eager_gpu_leg(dev, B)
        except Exception as ex:
            line["eager_gpu_baseline"] = dict(error=str(ex)[:300])
    if not args.no_infer:
        try:
            line["inference"] = infer_legs(eng, dev, args.infer_envs, max(3, min(args.steps, 10)))
        except Exception as ex:
            line["inference"] = dict(error=str(ex)[:200])
    if not args.no_cpu_baseline and world == 1:
        try:
            v, threads, sample = cpu_train_sample()
            line["cpu_baseline"] = dict(value=v, unit="samples/s", cores=threads, kind="port", sample=sample)
        except Exception as ex:
            line["cpu_baseline"] = dict(value=None, unit="samples/s", cores=os.cpu_count(), kind="port", sample=f"failed: {ex}")
    print(json.dumps(line))
    if world > 1:
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
