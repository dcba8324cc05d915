"""CUDA-graph time of the step's phases at the bench workload: forward, forward+backward, full step.
usage: python tools/phase_times.py [B]"""
import os
import sys
import torch
sys.path.insert(0, ".")
from act_plus_plus_b200.engine import ActEngine
from act_plus_plus_b200.policy import _init_like_reference

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
eng = ActEngine(batch_max=B, train=True)
_init_like_reference(eng, seed=0)
g = torch.Generator(device="cuda").manual_seed(0)
image = torch.rand((B, 1, 3, 480, 640), generator=g, device="cuda")
qpos = torch.randn((B, 14), generator=g, device="cuda")
actions = torch.randn((B, 100, 16), generator=g, device="cuda")
is_pad = torch.arange(100, device="cuda")[None].repeat(B, 1) >= 60
eps = torch.randn((B, 32), generator=g, device="cuda")


def phase(n):
    eng.zero_grad(overlap_forward=n >= 2)
    eng.forward(qpos, image, actions, is_pad, eps)
    if n == 2:
        eng.backward(0, 1.0)
    if n == 3:
        eng.backward(0, 1.0)
        eng.adamw_step()
    if n == 4:
        eng.backward_and_step(1.0)


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream(priority=int(os.environ.get('ACT_STREAM_PRIO', '-1')))
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        fn()
    gr.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        gr.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


t1 = timed(lambda: phase(1)); t2 = timed(lambda: phase(2)); t3 = timed(lambda: phase(3)); t4 = timed(lambda: phase(4))
print(f"B={B} forward {t1:.3f} ms  backward {t2 - t1:.3f} ms  adamw {t3 - t2:.3f} ms  step {t3:.3f} ms  step with overlapped optimiser {t4:.3f} ms")
