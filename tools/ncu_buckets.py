#!/usr/bin/env python
"""Development aid: warp-instructions and stall samples of one kernel grouped by source section ('// ---- ' markers and
function heads of csrc/track.cu):  python tools/ncu_buckets.py <rep> <kernel regex> <units> [source file]"""
import csv, io, subprocess, sys
rep, pat, units = sys.argv[1], sys.argv[2], float(sys.argv[3])
srcfile = sys.argv[4] if len(sys.argv) > 4 else "pbhc_b200/csrc/track.cu"
base = srcfile.split("/")[-1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
lines = open(srcfile).read().split("\n")
marks = [(i + 1, l.strip()[:70]) for i, l in enumerate(lines)
         if "// ---- " in l or l.startswith("__device__") or l.startswith("template") or l.startswith("__global__")]
cur, agg = None, {}
for r in csv.reader(io.StringIO(src)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) < 8 or r[0] in ("Line No", "Function Name", ""):
        continue
    try:
        inst, samp = int(r[7]), int(r[6])
    except ValueError:
        continue
    a = agg.setdefault((cur, int(r[0])), [0, 0])
    a[0] += inst
    a[1] += samp
out = {}
for (f, l), v in agg.items():
    if f == base:
        m = [x for x in marks if x[0] <= l]
        name = f"{m[-1][0]}:{m[-1][1]}" if m else "header"
    else:
        name = f
    o = out.setdefault(name, [0, 0])
    o[0] += v[0]
    o[1] += v[1]
ts = max(1, sum(v[1] for v in out.values()))
print("total", sum(v[0] for v in out.values()) / units)
for k, v in sorted(out.items(), key=lambda kv: -kv[1][0])[:30]:
    print(f"{v[0] / units:8.1f} inst/unit {100 * v[1] / ts:5.1f}% smp  {k}")
