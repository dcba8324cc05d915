"""Concurrency summary of gpurun_out/timeline.json (tools/timeline.py): per-stream busy time, union of busy intervals,
overlap factor (sum of kernel time / step span), idle gaps, and time by kernel family.  Markdown on stdout."""
import collections
import json
import re
import sys

ev = json.load(open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/timeline.json"))
span = max(e["ts"] + e["dur"] for e in ev)
tot = sum(e["dur"] for e in ev)
iv = sorted((e["ts"], e["ts"] + e["dur"]) for e in ev)
union, cur_s, cur_e, gaps = 0.0, iv[0][0], iv[0][1], []
for s, e in iv[1:]:
    if s > cur_e:
        union += cur_e - cur_s
        gaps.append((cur_e, s - cur_e))
        cur_s, cur_e = s, e
    else:
        cur_e = max(cur_e, e)
union += cur_e - cur_s
print(f"# Graph-replayed training step: kernel timeline summary\n")
print(f"kernels {len(ev)}, step span {span:.1f} us, summed kernel time {tot:.1f} us, busy (union) {union:.1f} us, "
      f"overlap factor {tot / span:.2f}, idle inside the step {span - union:.1f} us in {len(gaps)} gaps "
      f"(largest {max([g[1] for g in gaps] or [0]):.1f} us)\n")
by_stream = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    by_stream[e["stream"]][0] += 1; by_stream[e["stream"]][1] += e["dur"]
print("| stream | kernels | busy us | share of span |\n|---|---|---|---|")
for s, (n, d) in sorted(by_stream.items(), key=lambda kv: -kv[1][1]):
    print(f"| {s} | {n} | {d:.1f} | {d / span:.2f} |")
fam = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    m = re.search(r"act::(\w+)", e["name"])
    k = m.group(1) if m else e["name"].split("(")[0][:40]
    fam[k][0] += 1; fam[k][1] += e["dur"]
print("\n| kernel | launches | total us | mean us |\n|---|---|---|---|")
for k, (n, d) in sorted(fam.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"| {k} | {n} | {d:.1f} | {d / n:.2f} |")
