"""Per-kernel micro-benchmark of the tcgen05 GEMM / conv / wgrad launches on the ACT step's shapes.

Each case is captured into a CUDA graph of REP back-to-back launches (so host launch cost and tensor-map encoding are
excluded) and timed with CUDA events; cuBLAS / cuDNN (torch) timings of the same contraction are printed beside it as
the library bar.  usage: python tools/bench_kernels.py [gemm|conv|wgrad|all] [out.json]"""
import json
import sys

import torch

sys.path.insert(0, ".")
from act_plus_plus_b200 import ops  # noqa: E402

REP = 20


def graph_time(fn, rep=REP, iters=5):
    fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(rep):
            fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / rep)
    return best * 1e3   # us


GEMMS = [  # name, M, N, K, a_mn, b_mn, ksplit
    ("enc_qk", 2416, 1024, 512, 0, 0, 1), ("enc_v/out", 2416, 512, 512, 0, 0, 1), ("enc_ffn1", 2416, 3200, 512, 0, 0, 1),
    ("enc_ffn2", 2416, 512, 3200, 0, 0, 1), ("cvae_out", 816, 512, 512, 0, 0, 1), ("cvae_ffn1", 816, 3200, 512, 0, 0, 1),
    ("cvae_ffn2", 816, 512, 3200, 0, 0, 1), ("dx_ffn2", 2416, 3200, 512, 0, 1, 1), ("dx_ffn1", 2416, 512, 3200, 0, 1, 1),
    ("dw_ffn1", 3200, 512, 2416, 1, 1, 2), ("dw_ffn2", 512, 3200, 2416, 1, 1, 3), ("dw_attn", 512, 512, 2416, 1, 1, 10),
    ("stem_gemm", 614400, 64, 192, 0, 0, 1), ("big", 8192, 8192, 8192, 0, 0, 1),
]
ATTN = [("enc_qkT", 302, 304, 64, 64, 0, 0), ("enc_pv", 302, 64, 304, 64, 0, 1), ("cvae_qkT", 102, 104, 64, 64, 0, 0)]
CONVS = [(8, 120, 160, 64, 64, 3, 1), (8, 60, 80, 128, 128, 3, 1), (8, 30, 40, 256, 256, 3, 1), (8, 15, 20, 512, 512, 3, 1),
         (8, 120, 160, 64, 128, 3, 2), (8, 60, 80, 128, 256, 3, 2)]


def bench_gemm(rows):
    for name, M, N, K, amn, bmn, ks in GEMMS:
        a = torch.randn((K, M) if amn else (M, K), device="cuda").bfloat16()
        b = torch.randn((K, N) if bmn else (N, K), device="cuda").bfloat16()
        atomic = ks > 1
        out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if atomic else torch.bfloat16)
        epi = ops.make_epilogue(out_f32=out, ldc=N, atomic=True) if atomic else ops.make_epilogue(out_bf16=out, ldc=N)
        lda, ldb = (M if amn else K), (N if bmn else K)
        res = {}
        for bn in (0, 64, 128, 256):
            if bn > max(N, 64):
                continue
            try:
                res[bn] = graph_time(lambda: ops.tc_gemm_raw(M, N, K, 1, 1, a, amn, lda, 0, 0, b, bmn, ldb, 0, 0, epi, ks, bn, 0))
            except Exception as ex:
                res[bn] = None
                print("fail", name, bn, str(ex)[:100])
        A = a.t() if amn else a
        Bt = b if bmn else b.t()
        ref = graph_time(lambda: torch.matmul(A, Bt))
        fl = 2.0 * M * N * K
        rows.append(dict(kind="gemm", name=name, M=M, N=N, K=K, a_mn=amn, b_mn=bmn, ksplit=ks,
                         us={str(k): (round(v, 2) if v else None) for k, v in res.items()},
                         tflops_auto=round(fl / res[0] / 1e6, 1) if res.get(0) else None,
                         cublas_us=round(ref, 2), cublas_tflops=round(fl / ref / 1e6, 1)))
        print(json.dumps(rows[-1]), flush=True)
    for name, M, N, K, Z, amn, bmn in ATTN:
        a = torch.randn((Z, K, M) if amn else (Z, M, K), device="cuda").bfloat16()
        b = torch.randn((Z, K, N) if bmn else (Z, N, K), device="cuda").bfloat16()
        out = torch.zeros(Z, M, N, device="cuda", dtype=torch.float32)
        epi = ops.make_epilogue(out_f32=out, ldc=N, osz0=M * N)
        t = graph_time(lambda: ops.tc_gemm_raw(M, N, K, Z, 1, a, amn, a.stride(1), a.stride(0), 0, b, bmn, b.stride(1), b.stride(0), 0, epi, 1, 0, 0))
        A = a.transpose(1, 2) if amn else a
        Bt = b if bmn else b.transpose(1, 2)
        ref = graph_time(lambda: torch.bmm(A, Bt))
        fl = 2.0 * M * N * K * Z
        rows.append(dict(kind="bgemm", name=name, M=M, N=N, K=K, Z=Z, us=round(t, 2), tflops=round(fl / t / 1e6, 1),
                         cublas_us=round(ref, 2)))
        print(json.dumps(rows[-1]), flush=True)


def bench_conv(rows, wgrad=False):
    for (N, H, W, Cin, Cout, KH, stride) in CONVS:
        x = torch.randn(N, H, W, Cin, device="cuda").bfloat16()
        w = torch.randn(Cout, Cin, KH, KH, device="cuda") * 0.05
        wpk, wpkt = ops.pack_conv_weights(w)
        Ho, Wo = (H + 2 - KH) // stride + 1, (W + 2 - KH) // stride + 1
        fl = 2.0 * N * Ho * Wo * Cout * Cin * KH * KH
        xt = x.permute(0, 3, 1, 2)                       # NCHW view of channels-last memory
        wt = w.bfloat16().contiguous(memory_format=torch.channels_last)
        if not wgrad:
            y = torch.empty(N, Ho, Wo, Cout, device="cuda", dtype=torch.bfloat16)
            lib = ops._lib.lib()
            epi = ops.make_epilogue(out_bf16=y, relu=True)
            import ctypes as C

            def run():
                ops.check(lib.act_conv_fwd(N, H, W, Cin, Cout, KH, stride, 1, x.data_ptr(), wpk.data_ptr(), C.byref(epi),
                                           torch.cuda.current_stream().cuda_stream), "conv_fwd")
            t = graph_time(run)
            ref = graph_time(lambda: torch.nn.functional.conv2d(xt, wt, stride=stride, padding=1))
            rows.append(dict(kind="conv_fwd", shape=[N, H, W, Cin, Cout, KH, stride], us=round(t, 2),
                             tflops=round(fl / t / 1e6, 1), cudnn_us=round(ref, 2), cudnn_tflops=round(fl / ref / 1e6, 1)))
        else:
            dy = torch.randn(N, Ho, Wo, Cout, device="cuda").bfloat16()
            dw = torch.zeros(Cout, Cin, KH, KH, device="cuda")
            t = graph_time(lambda: ops.conv_wgrad(x, dy, KH, stride, 1, dw=dw))
            rows.append(dict(kind="conv_wgrad", shape=[N, H, W, Cin, Cout, KH, stride], us=round(t, 2),
                             tflops=round(fl / t / 1e6, 1)))
        print(json.dumps(rows[-1]), flush=True)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    rows = []
    if what in ("gemm", "all"):
        bench_gemm(rows)
    if what in ("conv", "all"):
        bench_conv(rows)
    if what in ("wgrad", "all"):
        bench_conv(rows, wgrad=True)
    if len(sys.argv) > 2:
        json.dump(rows, open(sys.argv[2], "w"), indent=1)


def bench_floor():
    """MMA issue floor per operand layout / tile width: run with ACT_TC_DEBUG=3 (no loads, no epilogue)."""
    M = N = K = 4096
    for amn in (0, 1):
        for bmn in (0, 1):
            a = torch.randn((K, M) if amn else (M, K), device="cuda").bfloat16()
            b = torch.randn((K, N) if bmn else (N, K), device="cuda").bfloat16()
            out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
            epi = ops.make_epilogue(out_bf16=out, ldc=N)
            for bn in (64, 128, 256):
                t = graph_time(lambda: ops.tc_gemm_raw(M, N, K, 1, 1, a, amn, M if amn else K, 0, 0, b, bmn, N if bmn else K, 0, 0, epi, 1, bn, 0), rep=5)
                tiles = (M // 128) * (N // bn)
                rounds = -(-tiles // 148)
                mmas = rounds * (K // 16)
                print(json.dumps(dict(kind="floor", a_mn=amn, b_mn=bmn, bn=bn, us=round(t, 1), tflops=round(2.0 * M * N * K / t / 1e6, 1),
                                      ns_per_mma=round(t * 1e3 / mmas, 1))), flush=True)


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "floor":
    bench_floor()


def bench_attn():
    """Fused scores+softmax kernel vs the unfused pair (tc_gemm fp32 scores + softmax kernel) and torch SDPA-free math."""
    import ctypes as C
    lib = ops._lib.lib()
    st = lambda: torch.cuda.current_stream().cuda_stream
    for (B, h, Lq, Lk) in ((8, 8, 302, 302), (8, 8, 102, 102), (8, 8, 100, 302), (8, 8, 100, 100), (16, 8, 902, 902), (16, 8, 100, 902),
                           (2, 8, 902, 902)):
        d = 512
        Lkp = (Lk + 7) // 8 * 8
        qk = torch.randn(B, Lq, 2 * d, device="cuda").bfloat16()
        kk = torch.randn(B, Lk, 2 * d, device="cuda").bfloat16()
        P = torch.zeros(B, h, Lq, Lkp, device="cuda", dtype=torch.bfloat16)
        S = torch.zeros(B, h, Lq, Lkp, device="cuda", dtype=torch.float32)

        def fused():
            ops.check(lib.act_attn_scores(0, qk.data_ptr(), 2 * d, 64, Lq * 2 * d, kk.data_ptr() + d * 2, 2 * d, 64, Lk * 2 * d, Lq, Lk, Lkp,
                                          h, B, C.c_float(0.125), None, None, P.data_ptr(), st()), "attn_scores")

        epi = ops.make_epilogue(out_f32=S, ldc=Lkp, osz0=Lq * Lkp, osz1=h * Lq * Lkp, alpha=0.125)

        def unfused():
            ops.tc_gemm_raw(Lq, Lk, 64, h, B, qk, 0, 2 * d, 64, Lq * 2 * d, kk[:, :, d:], 0, 2 * d, 64, Lk * 2 * d, epi, 1, 0, 0)
            ops.check(lib.act_softmax_fwd(S.data_ptr(), None, P.data_ptr(), B * h * Lq, Lk, Lkp, h * Lq, st()), "softmax")
        dO = torch.randn(B, Lq, d, device="cuda").bfloat16()
        V = torch.randn(B, Lk, d, device="cuda").bfloat16()
        dS = torch.zeros(B, h, Lq, Lkp, device="cuda", dtype=torch.bfloat16)

        def fused_bwd():
            ops.check(lib.act_attn_scores(1, dO.data_ptr(), d, 64, Lq * d, V.data_ptr(), d, 64, Lk * d, Lq, Lk, Lkp,
                                          h, B, C.c_float(0.125), None, P.data_ptr(), dS.data_ptr(), st()), "attn_scores bwd")

        epi2 = ops.make_epilogue(out_f32=S, ldc=Lkp, osz0=Lq * Lkp, osz1=h * Lq * Lkp, alpha=1.0)

        def unfused_bwd():
            ops.tc_gemm_raw(Lq, Lk, 64, h, B, dO, 0, d, 64, Lq * d, V, 0, d, 64, Lk * d, epi2, 1, 0, 0)
            ops.check(lib.act_softmax_bwd(P.data_ptr(), S.data_ptr(), dS.data_ptr(), B * h * Lq, Lk, Lkp, C.c_float(0.125), st()), "softmax bwd")
        tf, tu = graph_time(fused), graph_time(unfused)
        Oo = torch.randn(B, Lq, d, device="cuda").bfloat16()

        def fused_bwd_o():
            ops.check(lib.act_attn_scores_bwd(dO.data_ptr(), d, 64, Lq * d, V.data_ptr(), d, 64, Lk * d, Oo.data_ptr(), d, 64, Lq * d, Lq, Lk,
                                              Lkp, h, B, C.c_float(0.125), P.data_ptr(), dS.data_ptr(), st()), "attn_scores_bwd")
        tfb, tub, tfo = graph_time(fused_bwd), graph_time(unfused_bwd), graph_time(fused_bwd_o)
        print(json.dumps(dict(kind="attn_scores", B=B, h=h, Lq=Lq, Lk=Lk, fused_us=round(tf, 2), unfused_us=round(tu, 2),
                              fused_bwd_us=round(tfb, 2), unfused_bwd_us=round(tub, 2), fused_bwd_with_O_us=round(tfo, 2))), flush=True)


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "attn":
    bench_attn()


def bench_dw_sweep():
    """Weight-gradient GEMMs (dW = dY^T X, both operands MN-major, fp32 atomic accumulation): time over split-K / tile width,
    the data behind plan_splitk's cost model."""
    shapes = [("dw_ffn1", 3200, 512, 2416), ("dw_ffn2", 512, 3200, 2416), ("dw_attn", 512, 512, 2416), ("dw_qk", 1024, 512, 2416),
              ("dw_cvae_ffn1", 3200, 512, 816), ("dw_cvae_attn", 512, 512, 816), ("dw_dec_ffn1", 3200, 512, 800)]
    for name, M, N, K in shapes:
        a = torch.randn(K, M, device="cuda").bfloat16(); b = torch.randn(K, N, device="cuda").bfloat16()
        out = torch.zeros(M, N, device="cuda", dtype=torch.float32)
        epi = ops.make_epilogue(out_f32=out, ldc=N, atomic=True)
        best = None
        row = {}
        for bn in (64, 128, 256):
            for ks in (1, 2, 3, 4, 6, 8, 12, 16):
                if ks > (K + 63) // 64:
                    continue
                t = graph_time(lambda: ops.tc_gemm_raw(M, N, K, 1, 1, a, 1, M, 0, 0, b, 1, N, 0, 0, epi, ks, bn, 0))
                row[f"bn{bn}_ks{ks}"] = round(t, 2)
                if best is None or t < best[0]:
                    best = (t, bn, ks)
        A, Bt = a.t(), b
        ref = graph_time(lambda: torch.matmul(A, Bt))
        print(json.dumps(dict(kind="dw_sweep", name=name, M=M, N=N, K=K, best_us=round(best[0], 2), best_bn=best[1], best_ks=best[2],
                              cublas_us=round(ref, 2), us=row)), flush=True)


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "dw":
    bench_dw_sweep()
