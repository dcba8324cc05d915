"""Print gpurun_out/timeline.json (from tools/timeline.py): start, END, stream, grid, name.  usage: timeline_print.py [t0 t1]"""
import json, re, sys
ev = json.load(open("gpurun_out/timeline.json"))
t0 = float(sys.argv[1]) if len(sys.argv) > 1 else 0
t1 = float(sys.argv[2]) if len(sys.argv) > 2 else 1e9
main = max(set(e["stream"] for e in ev), key=lambda s: sum(1 for e in ev if e["stream"] == s))
for e in ev:
    if e["ts"] + e["dur"] < t0 or e["ts"] > t1:
        continue
    n = re.sub(r"^void ", "", e["name"]); n = n.replace("act::", "")
    print(f"{e['ts']:8.1f} ->{e['ts'] + e['dur']:8.1f} ({e['dur']:6.1f}) s{'M' if e['stream'] == main else e['stream']:>3} g{str(e['grid']):<15} {n[:40]}")
