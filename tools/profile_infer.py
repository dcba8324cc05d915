"""Run a few batched-inference steps (E environments on one GPU: policy query + temporal ensemble) for ncu, and dump
per-launch CUDA-event timings of the tensor-core kernels.  usage: python tools/profile_infer.py [E] [steps] [dump.csv]"""
import ctypes as C
import sys
import torch
sys.path.insert(0, ".")
from act_plus_plus_b200 import _lib
from act_plus_plus_b200.engine import ActEngine
from act_plus_plus_b200.policy import _init_like_reference
from act_plus_plus_b200.rollout import TemporalEnsembler

E = int(sys.argv[1]) if len(sys.argv) > 1 else 128
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dump = sys.argv[3] if len(sys.argv) > 3 else ""
eng = ActEngine(batch_max=E, train=False)
_init_like_reference(eng, seed=0)
ens = TemporalEnsembler(E, 100, 16, device=torch.device("cuda"))
g = torch.Generator(device="cuda").manual_seed(0)
image = torch.rand((E, 1, 3, 480, 640), generator=g, device="cuda")
qpos = torch.randn((E, 14), generator=g, device="cuda")
lib = _lib.lib()
eng.launches(reset=True)
for i in range(steps):
    if dump and i == steps - 1:
        torch.cuda.synchronize()
        eng.set_concurrent(False)
        lib.act_tc_profile_set_dump(dump.encode())
        lib.act_tc_profile_begin()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    ens.step(eng.forward(qpos, image)["a_hat"], i)
    t1.record()
    torch.cuda.synchronize()
    print(f"step {i}: {t0.elapsed_time(t1):.3f} ms")
if dump:
    a, b, c, d = C.c_double(), C.c_double(), C.c_int(), C.c_double()
    lib.act_tc_profile_end(C.byref(a), C.byref(b), C.byref(c), C.byref(d))
    print("tc ms", a.value, "launches", c.value, "TFLOP/s", b.value / a.value / 1e9)
print("launches/step", eng.launches() // steps)
