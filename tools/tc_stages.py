"""Where one small tc_kernel launch spends its time: %globaltimer stamps of CTA 0 (act_tc_debug_timestamps) for a CUDA
graph of back-to-back launches of one GEMM shape.  usage: python tools/tc_stages.py [M N K]"""
import sys
import torch
sys.path.insert(0, ".")
from act_plus_plus_b200 import ops, _lib

shapes = [tuple(int(x) for x in sys.argv[1:4])] if len(sys.argv) >= 4 else [(816, 512, 512), (2416, 512, 512), (2416, 3200, 512), (2416, 512, 3200)]
lib = _lib.lib()
names = ["entry", "setup done", "dep wait over", "operands landed", "last MMA issued", "acc ready", "epilogue done", "exit"]
for (M, N, K) in shapes:
    a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    epi = ops.make_epilogue(out_bf16=out, ldc=N)
    fn = lambda: ops.tc_gemm_raw(M, N, K, 1, 1, a, 0, K, 0, 0, b, 0, K, 0, 0, epi, 1, 0, 0)
    fn(); torch.cuda.synchronize()
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    ts = torch.zeros(1 + 8 * 64, dtype=torch.int64, device="cuda")
    lib.act_tc_debug_timestamps(ts.data_ptr())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            fn()
    lib.act_tc_debug_timestamps(None)
    g.replay(); torch.cuda.synchronize()
    ts.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    t = ts.cpu()
    n = int(t[0])
    st = t[1:1 + 8 * n].view(n, 8).double()
    st = st[st[:, 0].argsort()]
    print(f"\n## GEMM {M}x{N}x{K}: {e0.elapsed_time(e1) / 20 * 1e3:.2f} us per launch (graph of 20), {n} stamped launches")
    d = (st[5:15, 1:] - st[5:15, :-1]).mean(0)
    for i in range(7):
        print(f"  {names[i]:>16s} -> {names[i + 1]:<16s} {d[i]:8.0f} ns")
    print(f"  entry -> exit {float((st[5:15, 7] - st[5:15, 0]).mean()):.0f} ns; exit(i) -> entry(i+1) {float((st[6:16, 0] - st[5:15, 7]).mean()):.0f} ns; "
          f"exit(i) -> dep wait over(i+1) {float((st[6:16, 2] - st[5:15, 7]).mean()):.0f} ns; period {float((st[6:16, 0] - st[5:15, 0]).mean()):.0f} ns")
