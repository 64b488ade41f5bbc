#!/usr/bin/env python
"""Development aid: executed warp-instructions per SASS opcode of one kernel from an `ncu --set full --import-source on`
report:  python tools/ncu_opmix.py <rep> <kernel regex> [units] [top]"""
import csv
import io
import re
import subprocess
import sys


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    units = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + pat],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = next(r for r in rows if r and r[0] == "Address")
    ci, cs, ct = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
    agg, static = {}, {}
    for r in rows:
        if len(r) <= ci or not r[0].startswith("0x"):
            continue
        text = re.sub(r"^@!?U?P\d+\s+", "", r[1].strip())
        op = text.split()[0].rstrip(";") if text else "?"
        a = agg.setdefault(op, [0, 0, 0])
        a[0] += int(r[ci]); a[1] += int(r[cs]); a[2] += int(r[ct])
        static[op] = static.get(op, 0) + 1
    tot = sum(v[0] for v in agg.values())
    smp = max(1, sum(v[1] for v in agg.values()))
    print(f"{pat}: {tot} warp-instructions, {tot / units:.1f} per unit, {sum(static.values())} static SASS instructions")
    for op, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{v[0] / units:9.2f} /unit {100 * v[0] / tot:5.1f}%  lanes {v[2] / max(1, v[0]):5.1f}  samples {100 * v[1] / smp:5.1f}%  static {static[op]:5d}  {op}")


if __name__ == "__main__":
    main()
