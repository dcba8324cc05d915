"""Device time of dp_adamw_kernel with a same-GPU ghost peer (no NVLink in the path): separates the kernel's own cost
from the peer-memory traffic.  usage: python tools/dp_ghost_time.py"""
import sys
import torch
sys.path.insert(0, ".")
from act_plus_plus_b200 import _lib
from act_plus_plus_b200.engine import ActEngine
from act_plus_plus_b200.policy import _init_like_reference
from torch.profiler import profile, ProfilerActivity

B = 8
eng = ActEngine(batch_max=B, train=True)
_init_like_reference(eng, seed=0)
n = eng.flat_params.numel()
gg, gp = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
gs, gsig = torch.zeros(n, device="cuda", dtype=torch.bfloat16), torch.zeros(512, device="cuda", dtype=torch.int32)
mine = eng.dp_local_ptrs()
ghost = (gg.data_ptr(), gp.data_ptr(), gs.data_ptr(), gsig.data_ptr())
eng.dp_set_peers(0, 2, [[mine[i], ghost[i]] for i in range(4)])
_lib.check(eng.lib.act_engine_dp_preset_signal(eng._h, 1, 1 << 30), "preset")
g = torch.Generator(device="cuda").manual_seed(0)
image = torch.rand((B, 1, 3, 480, 640), generator=g, device="cuda")
qpos = torch.randn((B, 14), generator=g, device="cuda")
actions = torch.randn((B, 100, 16), generator=g, device="cuda")
is_pad = torch.arange(100, device="cuda")[None].repeat(B, 1) >= 60
eps = torch.randn((B, 32), generator=g, device="cuda")
eng.set_concurrent(False)


def step():
    eng.zero_grad()
    eng.forward(qpos, image, actions, is_pad, eps)
    eng.backward_and_step_p2p(1.0)


for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
for e in prof.events():
    if "dp_" in e.name or "adamw" in e.name:
        print(f"{e.name[:60]:60s} {e.device_time:8.1f} us")
for off, cnt in eng.dp_slices():
    print("range", off, cnt)
