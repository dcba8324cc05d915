"""Render tools/bench_kernels.py JSON rows as a markdown table.  usage: python tools/microbench_md.py in.json out.md"""
import json
import sys

rows = json.load(open(sys.argv[1]))
out = ["# Kernel micro-benchmark (CUDA-graph of 20 back-to-back launches, CUDA events, best of 5; B200)", "",
       "Library columns are torch.matmul / torch.bmm (cuBLAS) and F.conv2d channels-last bf16 (cuDNN) on the same contraction.", ""]
g = [r for r in rows if r["kind"] == "gemm"]
if g:
    out += ["## GEMM (tc_kernel MODE_GEMM; `auto` = planner's tile, then forced BN)", "",
            "| case | M | N | K | A/B major | split | auto us | BN64 | BN128 | BN256 | auto TFLOP/s | cuBLAS us | cuBLAS TFLOP/s |", "|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    for r in g:
        u = r["us"]
        out.append(f"| {r['name']} | {r['M']} | {r['N']} | {r['K']} | {'MN' if r['a_mn'] else 'K'}/{'MN' if r['b_mn'] else 'K'} | {r['ksplit']} | "
                   f"{u.get('0')} | {u.get('64')} | {u.get('128')} | {u.get('256')} | {r['tflops_auto']} | {r['cublas_us']} | {r['cublas_tflops']} |")
    out.append("")
g = [r for r in rows if r["kind"] == "bgemm"]
if g:
    out += ["## Batched attention GEMMs (Z = batch x heads)", "", "| case | M | N | K | Z | us | TFLOP/s | cuBLAS us |", "|---|---|---|---|---|---|---|---|"]
    for r in g:
        out.append(f"| {r['name']} | {r['M']} | {r['N']} | {r['K']} | {r['Z']} | {r['us']} | {r['tflops']} | {r['cublas_us']} |")
    out.append("")
g = [r for r in rows if r["kind"] == "conv_fwd"]
if g:
    out += ["## Conv forward (3x3/s1 -> conv3_kernel, 3x3/s2 -> tc_kernel MODE_CONV); shape = N,H,W,Cin,Cout,k,stride", "",
            "| shape | us | TFLOP/s | cuDNN us | cuDNN TFLOP/s |", "|---|---|---|---|---|"]
    for r in g:
        out.append(f"| {r['shape']} | {r['us']} | {r['tflops']} | {r['cudnn_us']} | {r['cudnn_tflops']} |")
    out.append("")
g = [r for r in rows if r["kind"] == "conv_wgrad"]
if g:
    out += ["## Conv weight gradient (MODE_WGRAD into staged T[KK][Cout][Cin] + fold)", "", "| shape | us | TFLOP/s |", "|---|---|---|"]
    for r in g:
        out.append(f"| {r['shape']} | {r['us']} | {r['tflops']} |")
    out.append("")
open(sys.argv[2], "w").write("\n".join(out))
print("\n".join(out))
