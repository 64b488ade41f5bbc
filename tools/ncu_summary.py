#!/usr/bin/env python
"""Writes profiles/track_kernel_traffic.json (DRAM bytes per launch of the dominant kernel, read by bench.py's
roofline.traffic) from an `ncu --set full` report:  python tools/ncu_summary.py <rep> [kernel regex]"""
import csv
import io
import json
import os
import subprocess
import sys

rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else "track_kernel"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, un = rows[0], rows[1]
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tot, n, dur = 0.0, 0, 0.0
for r in rows[2:]:
    d = dict(zip(hdr, r))
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        tot += float(d[k]) * scale[un[hdr.index(k)]]
    dur += float(d["gpu__time_duration.sum"]) * {"us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(un[hdr.index("gpu__time_duration.sum")], 1e-6)
    n += 1
out = {"kernel": pat, "launches": n, "dram_bytes_per_launch": tot / n, "ncu_duration_s_per_launch": dur / n, "report": os.path.basename(rep)}
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "track_kernel_traffic.json")
with open(p, "w") as f:
    json.dump(out, f, indent=1)
print(out)
