"""DRAM traffic of the tcgen05 launches inside a WARM training step against their algorithmic (compulsory) bytes.
Joins the ncu pass `--cache-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum`
(tools/gpu_profile_r02.sh, long CSV: one row per launch and metric) with the labelled launch list of the same program
(tools/profile_step.py dump: one label per launch, same order).
usage: python tools/tc_dram_warm.py ncu_tc_dram_warm.csv tc_launches.csv out.md"""
import collections
import csv
import re
import sys

ncu_path, lab_path, out = sys.argv[1:4]
labels = [r["label"] for r in csv.DictReader(open(lab_path))]
per_step = len(labels)
launch = collections.OrderedDict()
for r in csv.reader(open(ncu_path)):
    if len(r) < 15 or r[0] == "ID":
        continue
    d = launch.setdefault(int(r[0]), dict(name=re.sub(r"\(.*", "", r[4]).replace("void ", "")))
    d[r[12]] = float(r[14].replace(",", ""))
    d["unit_" + r[12]] = r[13]
rows = list(launch.values())[-per_step:]
assert len(rows) == per_step, (len(rows), per_step)


def num(label, key):
    m = re.search(key + r"=(\d+)", label)
    return int(m.group(1)) if m else 0


def algorithmic_bytes(label):
    """operands read once + result written once (bf16 operands; bf16 outputs except fp32 weight gradients)"""
    kind = label.split()[0]
    if kind == "gemm":
        M, N, K, z = (num(label, k) for k in ("M", "N", "K", "z"))
        return z * 2 * (M * K + N * K + M * N)
    if kind.startswith("attn"):
        Lq, Lk, z = (num(label, k) for k in ("Lq", "Lk", "z"))
        extra = 2 * z * Lq * Lk if kind == "attn_dS" else 0          # dS also reads P
        return z * 2 * (Lq * 64 + Lk * 64 + Lq * Lk) + extra
    n, h, w = (int(x) for x in label.split()[1].split("x"))
    cin, cout, taps = (num(label, k) for k in ("Cin", "Cout", "taps"))
    pix = n * h * w
    if kind == "wgrad":
        return 2 * pix * (cin + cout) + 4 * taps * cin * cout
    return 2 * pix * (cin + cout) + 2 * taps * cin * cout             # conv / conv3 (fwd or dgrad)


agg = collections.OrderedDict()
mismatch = 0
for lab, r in zip(labels, rows):
    kind = lab.split()[0]
    if (kind == "conv3") != ("conv3_kernel" in r["name"]) or kind.startswith("attn") != ("attn_scores" in r["name"]):
        mismatch += 1
    a = agg.setdefault(kind, dict(n=0, rd=0.0, wr=0.0, alg=0.0, us=0.0))
    a["n"] += 1
    a["rd"] += r["dram__bytes_read.sum"] * (1e6 if r["unit_dram__bytes_read.sum"] == "Mbyte" else 1e3 if r["unit_dram__bytes_read.sum"] == "Kbyte" else 1)
    a["wr"] += r["dram__bytes_write.sum"] * (1e6 if r["unit_dram__bytes_write.sum"] == "Mbyte" else 1e3 if r["unit_dram__bytes_write.sum"] == "Kbyte" else 1)
    a["alg"] += algorithmic_bytes(lab)
    t = r["gpu__time_duration.sum"]
    u = r["unit_gpu__time_duration.sum"]
    a["us"] += t / 1e3 if u in ("ns", "nsecond") else t if u in ("us", "usecond") else t * 1e3
tot = dict(n=0, rd=0.0, wr=0.0, alg=0.0, us=0.0)
for a in agg.values():
    for k in tot:
        tot[k] += a[k]
lines = ["# DRAM traffic of the tcgen05 kernels inside a warm training step (batch 8, 1 camera)", "",
         "`ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"
         " -k regex:tc_kernel|conv3_kernel|attn_scores python tools/profile_step.py 8 3` -- last step "
         f"({per_step} launches, direct launch, serialised by ncu but WITHOUT ncu's cache flush: L2 keeps what the previous kernels "
         "left, as in the real step).  Algorithmic bytes = every operand read once + the result written once (bf16 operands; "
         "fp32 for staged weight gradients); split-K partial sums, residual / bias reads and fp32 second outputs are not counted, "
         f"so a ratio below 1 means L2 served the operands.  Label/kernel order mismatches: {mismatch}.", "",
         "| class | launches | dram read MB | dram write MB | dram total MB | algorithmic MB | dram / algorithmic | sum of durations us |",
         "|---|---|---|---|---|---|---|---|"]
for k, a in list(agg.items()) + [("all", tot)]:
    lines.append(f"| {k} | {a['n']} | {a['rd'] / 1e6:.1f} | {a['wr'] / 1e6:.1f} | {(a['rd'] + a['wr']) / 1e6:.1f} | {a['alg'] / 1e6:.1f} | "
                 f"{(a['rd'] + a['wr']) / max(a['alg'], 1):.2f} | {a['us']:.1f} |")
lines += ["", f"Mean per launch: {(tot['rd'] + tot['wr']) / tot['n'] / 1e6:.3f} MB measured, {tot['alg'] / tot['n'] / 1e6:.3f} MB algorithmic "
          f"(this is `roofline.traffic` in bench.py, bytes per launch = {(tot['rd'] + tot['wr']) / tot['n']:.0f})."]
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
