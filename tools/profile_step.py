"""Run a few direct-launch training steps of the bench workload (for ncu) and dump per-launch CUDA-event timings."""
import ctypes as C
import sys
import torch
sys.path.insert(0, ".")
from act_plus_plus_b200 import _lib
from act_plus_plus_b200.engine import ActEngine
from act_plus_plus_b200.policy import _init_like_reference
from act_plus_plus_b200.train import train_step

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dump = sys.argv[3] if len(sys.argv) > 3 else ""
eng = ActEngine(batch_max=B, train=True)
_init_like_reference(eng, seed=0)
g = torch.Generator(device="cuda").manual_seed(0)
image = torch.rand((B, 1, 3, 480, 640), generator=g, device="cuda")
qpos = torch.randn((B, 14), generator=g, device="cuda")
actions = torch.randn((B, 100, 16), generator=g, device="cuda")
is_pad = torch.arange(100, device="cuda")[None].repeat(B, 1) >= 60
eps = torch.randn((B, 32), generator=g, device="cuda")
lib = _lib.lib()
for i in range(steps):
    if dump and i == steps - 1:
        torch.cuda.synchronize()
        lib.act_tc_profile_set_dump(dump.encode())
        lib.act_tc_profile_begin()
    train_step(eng, qpos, image, actions, is_pad, eps)
torch.cuda.synchronize()
if dump:
    a, b, c, d = C.c_double(), C.c_double(), C.c_int(), C.c_double()
    lib.act_tc_profile_end(C.byref(a), C.byref(b), C.byref(c), C.byref(d))
    print("tc ms", a.value, "launches", c.value, "TFLOP/s", b.value / a.value / 1e9)
print("launches/step", eng.launches() // steps)
