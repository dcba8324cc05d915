"""Last part of the step timeline (gpurun_out/timeline.json): what runs in the final N microseconds and on which stream.
usage: python tools/timeline_tail.py [us=500]"""
import json, re, sys
ev = json.load(open("gpurun_out/timeline.json"))
win = float(sys.argv[1]) if len(sys.argv) > 1 else 500.0
end = max(e["ts"] + e["dur"] for e in ev)
print(f"step span {end:.1f} us; kernels active in the last {win:.0f} us:")
for e in sorted(ev, key=lambda e: e["ts"]):
    if e["ts"] + e["dur"] >= end - win:
        m = re.search(r"act::(\w+)(<[^>]*>)?", e["name"])
        nm = (m.group(1) + (m.group(2) or "")) if m else e["name"][:50]
        print(f"  t={e['ts']:8.1f} dur={e['dur']:7.1f} end-{end - e['ts'] - e['dur']:6.1f} stream {e['stream']:>4} grid {e['grid']} {nm}")
