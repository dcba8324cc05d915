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


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d.get("bf16_tflops_sustained", d["bf16_tflops"]), src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def cpu_train_sample(batch=2, threads=None):
    """Reference algorithm on host cores (oracle port = torch CPU fp32 restatement of the reference modules):
    one fwd + loss + backward + AdamW step on a bounded sample (batch `batch` of the same 480x640 workload)."""
    import torch
    from oracle import act_oracle as O
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    cfg = O.make_cfg(num_cameras=K_CAM)
    sd = O.init_state_dict(cfg, 0)
    b = O.make_inputs(cfg, batch, H, W, 0)
    names = O.trainable_names(cfg)
    params = [torch.nn.Parameter(sd[n]) for n in names]
    opt = torch.optim.AdamW(params, lr=1e-5, weight_decay=1e-4)
    t0 = time.time()
    leaf = dict(sd)
    leaf.update({n: p for n, p in zip(names, params)})
    opt.zero_grad()
    out = O.policy_forward(leaf, cfg, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"])
    out["loss"].backward()
    opt.step()
    dt = time.time() - t0
    return batch / dt, threads, f"1 optimisation step (fwd+loss+bwd+AdamW) at batch {batch} of the 480x640 1-camera workload, {dt:.1f} s"


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    vals = []
    for _ in range(max(1, min(args.warmup, 1))):
        cpu_train_sample(2)
    for _ in range(max(1, args.steps if args.steps <= 3 else 3)):
        v, threads, sample = cpu_train_sample(2)
        vals.append(v)
    value = sum(vals) / len(vals)
    line = dict(impl="reference", metric="ACT train samples/s", value=value, unit="samples/s", n_gpus=args.gpus,
                steps=len(vals), warmup=1, ms_per_step=2.0 / value * 1e3, higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload="ACT sim_transfer_cube_scripted train step, batch 8/GPU, 1 camera 480x640, chunk 100 "
                            "(reference algorithm on host CPU, bounded sample: batch 2 per step)", torch=torch.__version__),
                cpu_baseline=dict(value=value, unit="samples/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=value, unit="samples/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line))


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
               tflops=34.9e9 * envs / ms / 1e9, action_checksum=float(act.sum()))
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
    lib.act_tc_profile_begin()
    PSTEPS = 2
    for i in range(PSTEPS):
        step_direct(devb[i % NB])
    torch.cuda.synchronize()
    tot_ms, tot_fl, nl, mx = C.c_double(), C.c_double(), C.c_int(), C.c_double()
    _lib.check(lib.act_tc_profile_end(C.byref(tot_ms), C.byref(tot_fl), C.byref(nl), C.byref(mx)), "tc_profile_end")
    eng.set_concurrent(True)
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
                config=dict(workload="ACT sim_transfer_cube_scripted training step (fwd+loss+bwd+AdamW), per-GPU batch 8, "
                            "1 camera 480x640, qpos 14, chunk 100, hidden 512, ff 3200, enc 4/dec 7, kl_weight 10 (BASELINE configs[1])",
                            global_batch=B * world, parallelism=f"dp{world}", cuda_graph=graph is not None,
                            l2="4 rotating input batches; parameters+Adam state (1.3 GB) exceed the 126 MB L2 every step",
                            dropout="0 (reference validation path; see DESIGN.md)", last_loss=loss_host),
                e2e=dict(value=n_samples / (ms_e2e / 1e3), unit="samples/s", h2d_bytes_per_step=h2d, d2h_bytes_per_step=4),
                gpu_launches=launches_per_step * args.steps,
                roofline=dict(bound="tensor", achieved=tc_tflops, peak=peaks["tf_sust"], unit="TFLOP/s",
                              frac=tc_tflops / peaks["tf_sust"], traffic=TC_DRAM_BYTES_PER_LAUNCH, peak_source=peaks["src"] + " (sustained bf16)",
                              traffic_source="profiles/r01_tc_dram_per_step.md: ncu dram__bytes_read+write summed over the 277 tcgen05 launches "
                                             "of one step (cold cache per launch) / 277",
                              kernel="tc_kernel<*> (all tcgen05 GEMM/conv/wgrad launches of one step)",
                              launches_per_step=nl.value // PSTEPS, tc_ms_per_step=tc_ms_step,
                              timing="CUDA events around every launch, serial direct-launch replay of the step "
                                     "(ACT engine concurrency off); includes the event-pair cost below",
                              event_pair_us=ev_pair_ms * 1e3,
                              achieved_net_of_event_cost=(tot_fl.value / PSTEPS) / max(1e-9, (tc_ms_step - (nl.value // PSTEPS) * ev_pair_ms) / 1e3) / 1e12,
                              tc_share_of_serial_step=None, longest_launch_ms=mx.value,
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
    if not args.no_infer:
        try:
            line["inference"] = infer_legs(eng, dev, args.infer_envs, max(3, min(args.steps, 10)))
        except Exception as ex:
            line["inference"] = dict(error=str(ex)[:200])
    if not args.no_cpu_baseline and world == 1:
        try:
            v, threads, sample = cpu_train_sample(2)
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
