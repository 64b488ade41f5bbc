#!/usr/bin/env python
"""Development aid: per-source-line instruction counts and stall samples of one kernel from an .ncu-rep
(needs -lineinfo).  python tools/ncu_lines.py <rep> <kernel regex> [n_units] [top]"""
import csv
import io
import subprocess
import sys


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    units = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, un = rows[0], rows[1]
    keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
            "launch__occupancy_limit_shared_mem", "launch__grid_size", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__t_sector_hit_rate.pct"]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(d.get("Kernel Name"))
        for k in keys:
            if k in d:
                print("  ", k, d[k], un[hdr.index(k)])
        for k in hdr:
            if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
                try:
                    if float(d[k]) > 0.25:
                        print("     stall", k.split("issue_stalled_")[1].split("_per_issue")[0], d[k])
                except ValueError:
                    pass
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + pat],
                         capture_output=True, text=True).stdout
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
        a = agg.setdefault((cur, int(r[0]), r[1].strip()[:100]), [0, 0])
        a[0] += inst
        a[1] += samp
    tot, tots = sum(v[0] for v in agg.values()), max(1, sum(v[1] for v in agg.values()))
    print(f"total inst {tot}  per unit {tot / units:.1f}  samples {tots}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{v[0] / units:8.1f} inst/unit {100 * v[1] / tots:5.1f}% smp  {k[0]}:{k[1]}  {k[2]}")


if __name__ == "__main__":
    main()
