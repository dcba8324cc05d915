"""Per-kernel SASS evidence for the shipped library: counts of the tcgen05 / TMEM / TMA instructions in every kernel of
libact_b200.so (cuobjdump -sass).  Markdown on stdout.  usage: python tools/sass_summary.py > profiles/rNN_sass_summary.md"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "act_plus_plus_b200/libact_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = dict(UTCHMMA="tcgen05.mma", LDTM="tcgen05.ld (TMEM->reg)", STTM="tcgen05.st", UTMALDG="TMA tensor load", UTMASTG="TMA tensor store",
            UTCBAR="tcgen05.commit", SYNCS="mbarrier ops", UBLKCP="bulk copy", ELECT="elect.sync")
rows, tot = [], collections.Counter()
for blk in re.split(r"\n\s*Function : ", txt)[1:]:
    name = blk.split("\n")[0].strip()
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    short = re.sub(r"\(.*", "", dem)
    c = {k: len(re.findall(r"\b" + k + r"\b", blk)) for k in pats}
    n_ins = len(re.findall(r"/\*[0-9a-f]{4,5}\*/", blk))
    rows.append((short, n_ins, c))
    tot.update(c)
print("# SASS summary of libact_b200.so (cuobjdump -sass, sm_100a)\n")
print("Totals: " + ", ".join(f"{k} {tot[k]}" for k in pats) + "\n")
print("| kernel | SASS instr | " + " | ".join(pats) + " |\n|---|---|" + "---|" * len(pats))
for short, n, c in sorted(rows, key=lambda r: -r[2]["UTCHMMA"] * 100000 - r[1]):
    if c["UTCHMMA"] or c["UTMALDG"] or c["LDTM"]:
        print(f"| `{short}` | {n} | " + " | ".join(str(c[k]) for k in pats) + " |")
print(f"\n{sum(1 for r in rows if not (r[2]['UTCHMMA'] or r[2]['UTMALDG'] or r[2]['LDTM']))} further kernels (LayerNorm, softmax, loss, AdamW, "
      "packing, temporal ensemble, data-parallel exchange ...) use no tensor-core / TMA instructions: bandwidth-bound CUDA-core kernels.")
print("\nLegend: " + "; ".join(f"{k} = {v}" for k, v in pats.items()))
