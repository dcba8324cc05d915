"""Kernel timeline of one graph-replayed training step via torch.profiler (CUPTI activity records keep the real
concurrency).  Writes gpurun_out/timeline.json (name, stream, start us, dur us per kernel of the last replay).
usage: python tools/timeline.py [B]"""
import json
import os
import sys
import torch
sys.path.insert(0, ".")
from act_plus_plus_b200.engine import ActEngine
from act_plus_plus_b200.policy import _init_like_reference
from act_plus_plus_b200.train import train_step

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
group, rank = None, 0
if int(os.environ.get("WORLD_SIZE", "1")) > 1:          # under torchrun: data-parallel step, rank 0 writes the timeline
    import torch.distributed as dist
    rank = int(os.environ["RANK"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    group = dist.group.WORLD
eng = ActEngine(batch_max=B, train=True)
_init_like_reference(eng, seed=0)
if group is not None and os.environ.get("ACT_DP_P2P", "0") == "1":
    eng.dp_connect(group)
g = torch.Generator(device="cuda").manual_seed(0)
image = torch.rand((B, 1, 3, 480, 640), generator=g, device="cuda")
qpos = torch.randn((B, 14), generator=g, device="cuda")
actions = torch.randn((B, 100, 16), generator=g, device="cuda")
is_pad = torch.arange(100, device="cuda")[None].repeat(B, 1) >= 60
eps = torch.randn((B, 32), generator=g, device="cuda")
fn = lambda: train_step(eng, qpos, image, actions, is_pad, eps, group)
for _ in range(3):
    fn()
torch.cuda.synchronize()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    fn()
torch.cuda.current_stream().wait_stream(s)
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr):
    fn()
for _ in range(3):
    gr.replay()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
if rank != 0:
    torch.cuda.synchronize()
    os._exit(0)
os.makedirs("gpurun_out", exist_ok=True)
prof.export_chrome_trace("gpurun_out/trace_full.json")
tr = json.load(open("gpurun_out/trace_full.json"))
ev = [e for e in tr["traceEvents"] if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
ev.sort(key=lambda e: e["ts"])
n = len(ev) // 3
ev = ev[-n:]
t0 = ev[0]["ts"]
out = [dict(name=e["name"][:90], stream=e["args"].get("stream"), ts=round(e["ts"] - t0, 2), dur=round(e["dur"], 2),
            grid=e["args"].get("grid"), block=e["args"].get("block")) for e in ev]
json.dump(out, open("gpurun_out/timeline.json", "w"))
os.remove("gpurun_out/trace_full.json")
print("kernels", len(out), "span us", out[-1]["ts"] + out[-1]["dur"])
sys.stdout.flush()
if group is not None:
    os._exit(0)
