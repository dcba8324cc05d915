"""Measurement helpers shared by bench.py and tools/: clocks sampled under load, the box's measured peaks, and the
differential tensor-kernel timing (serial CUDA-graph replays of a step with one kernel class skipped at a time)."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import threading

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TC_CLASSES = (("gemm", 1), ("conv", 2), ("conv3", 4), ("wgrad", 8), ("attn", 16))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.rows = []
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        self.cmd = ["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(index), "-lms", "50"]
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
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback (B200_PROFILING.md)")


def graph_time_ms(fn, replays: int = 5):
    """Median device time (CUDA events around one replay) of `fn` captured into a CUDA graph; fn runs once eagerly first."""
    import torch
    fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(replays):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    del g
    return sorted(ts)[len(ts) // 2]


def tensor_time_by_class(lib, step_fn, save_restore=None, replays: int = 5):
    """Device time and algorithmic FLOPs of every tensor-core kernel class inside `step_fn` (a serial, single-stream
    sequence of launches: engine concurrency off).  T(class) = T(serial step) - T(serial step with that class skipped),
    both measured as CUDA-graph replays: no per-launch event pairs, no launch gaps.  `save_restore` = (save, restore)
    callables protecting state the garbage steps would corrupt (parameters, Adam moments)."""
    state = save_restore[0]() if save_restore else None
    out = {}
    try:
        lib.act_tc_set_skip_mask(0)
        t_full = graph_time_ms(step_fn, replays)
        lib.act_tc_set_skip_mask(31)
        t_none = graph_time_ms(step_fn, replays)                         # the step without any tensor-core kernel
        for name, bit in TC_CLASSES:
            lib.act_tc_set_skip_mask(bit)
            step_fn()                                                    # eager pass: count the skipped FLOPs of ONE step
            flops = float(lib.act_tc_skipped_flops())
            if flops <= 0.0:
                continue
            lib.act_tc_set_skip_mask(bit)
            t_wo = graph_time_ms(step_fn, replays)
            ms = max(1e-6, t_full - t_wo)
            out[name] = dict(ms=ms, gflop=flops / 1e9, tflops=flops / (ms * 1e9))
    finally:
        lib.act_tc_set_skip_mask(0)
        if save_restore:
            save_restore[1](state)
    tot_fl = sum(v["gflop"] for v in out.values())
    tot_ms = max(1e-6, t_full - t_none)
    return dict(serial_step_ms=t_full, non_tensor_ms=t_none, tensor_ms=tot_ms, tensor_gflop=tot_fl,
                tensor_tflops=tot_fl / tot_ms, by_class=out)
