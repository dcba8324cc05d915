"""One forward + one backward call of the chunked attention-scores kernel at the 3-camera shape (for ncu)."""
import ctypes as C, sys, torch
sys.path.insert(0, ".")
from act_plus_plus_b200 import ops
lib = ops._lib.lib()
B, h, Lq, Lk, d = 16, 8, 902, 902, 512
Lkp = (Lk + 7) // 8 * 8
st = lambda: torch.cuda.current_stream().cuda_stream
qk = torch.randn(B, Lq, 2 * d, device="cuda").bfloat16()
P = torch.zeros(B, h, Lq, Lkp, device="cuda", dtype=torch.bfloat16)
dO = torch.randn(B, Lq, d, device="cuda").bfloat16(); V = torch.randn(B, Lk, d, device="cuda").bfloat16()
dS = torch.zeros_like(P)
for _ in range(2):
    ops.check(lib.act_attn_scores(0, qk.data_ptr(), 2 * d, 64, Lq * 2 * d, qk.data_ptr() + d * 2, 2 * d, 64, Lk * 2 * d, Lq, Lk, Lkp, h, B,
                                  C.c_float(0.125), None, None, P.data_ptr(), st()), "fwd")
    ops.check(lib.act_attn_scores(1, dO.data_ptr(), d, 64, Lq * d, V.data_ptr(), d, 64, Lk * d, Lq, Lk, Lkp, h, B, C.c_float(0.125), None,
                                  P.data_ptr(), dS.data_ptr(), st()), "bwd")
torch.cuda.synchronize()
print("ok", float(P.float().sum()))
