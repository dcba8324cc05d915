"""Join an ncu launch list (gpu__time_duration) of the training step with the tc launch labels dumped by
tools/profile_step.py (same launch order) -> per-shape totals.  usage: ncu_join.py launches.csv labels.csv per_step"""
import collections, csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0] != "ID"]
per = int(sys.argv[3])
rows = rows[-per:]
tc = [r for r in rows if "tc_kernel" in r[4] or "conv3_kernel" in r[4]]
labels = [r["label"] for r in csv.DictReader(open(sys.argv[2]))]
labels = labels[:len(tc)]
def flops(k):
    m = re.search(r"gemm M=(\d+) N=(\d+) K=(\d+) z=(\d+)", k)
    if m:
        M, N, K, Z = map(int, m.groups()); return 2.0 * M * N * K * Z
    m = re.search(r"(conv3|conv|wgrad) (\d+)x(\d+)x(\d+) Cin=(\d+) Cout=(\d+) taps=(\d+)", k)
    if m:
        n, h, w, ci, co, t = map(int, m.groups()[1:]); return 2.0 * n * h * w * ci * co * t
    return 0.0
agg = collections.OrderedDict()
for r, l in zip(tc, labels):
    a = agg.setdefault(re.sub(r" grid=.*", "", l), [0, 0.0])
    a[0] += 1; a[1] += float(r[-1]) / 1e3
other = sum(float(r[-1]) / 1e3 for r in rows if "tc_kernel" not in r[4] and "conv3_kernel" not in r[4])
tot = sum(a[1] for a in agg.values())
print(f"tc launches {len(tc)} total {tot:.1f} us; other kernels {other:.1f} us")
for k, a in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    print(f"{a[0]:3d} {a[1]:7.1f}us {a[1]/a[0]:6.1f}/l {flops(k)/(a[1]/a[0])/1e6:7.1f}TF  {k}")
