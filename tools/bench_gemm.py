"""Micro-benchmark of the tcgen05 GEMM on the ACT path's shapes (CUDA events, L2 flushed between runs)."""
import json
import sys
import torch
sys.path.insert(0, ".")
from act_plus_plus_b200 import ops

SHAPES = [  # (name, M, N, K)
    ("qk_proj", 2416, 1024, 512), ("v_proj", 2416, 512, 512), ("ffn1", 2416, 3200, 512), ("ffn2", 2416, 512, 3200),
    ("ffn1_b64", 64 * 302, 3200, 512), ("ffn2_b64", 64 * 302, 512, 3200), ("big", 8192, 8192, 8192),
]


def timeit(fn, iters=20):
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    rows = []
    for name, M, N, K in SHAPES:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        epi = ops.make_epilogue(out_bf16=out, ldc=N)
        best = None
        for bn in (64, 128, 256):
            if bn > N: continue
            for st in (3, 4, 6):
                try:
                    t = timeit(lambda: ops.tc_gemm_raw(M, N, K, 1, 1, a, 0, K, 0, 0, b, 0, K, 0, 0, epi, 1, bn, st))
                except Exception as ex:
                    print("fail", name, bn, st, ex); continue
                if best is None or t < best[0]: best = (t, bn, st)
        tt = timeit(lambda: torch.matmul(a, b.t()))
        fl = 2.0 * M * N * K
        rows.append(dict(name=name, M=M, N=N, K=K, ours_ms=best[0], bn=best[1], stages=best[2],
                         ours_tflops=fl / best[0] / 1e9, cublas_ms=tt, cublas_tflops=fl / tt / 1e9))
        print(json.dumps(rows[-1]))
    json.dump(rows, open("gpurun_out/bench_gemm.json", "w"), indent=1)


if __name__ == "__main__":
    main()
