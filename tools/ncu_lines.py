#!/usr/bin/env python
"""tools/ncu_lines.py REPORT.ncu-rep [top] ['name:lo-hi,...'] -- per-CUDA-line warp-stall samples of a
`ncu --set full --import-source on` capture (kernels built with -lineinfo).  Prints the hottest
source lines with their top stall reasons, plus totals over optional named line ranges."""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = None
    data = []
    fname = ""
    for r in rows:
        if len(r) >= 2 and r[0] == "File Name":
            fname = r[1].split("/")[-1]
        if len(r) > 10 and r[0] == "Line No":
            h = r
            si = h.index("# Samples"); ie = h.index("Instructions Executed"); bi = h.index("stall_barrier")
            continue
        if h and len(r) > 10 and r[0].isdigit():
            try:
                data.append((int(r[si]), fname, int(r[0]), r[1].strip()[:110], int(r[ie]),
                             [(int(r[bi + j] or 0), h[bi + j]) for j in range(17)]))
            except ValueError:
                pass
    tot = sum(d[0] for d in data) or 1
    print("total samples %d" % tot)
    reasons = {}
    for d in data:
        for c, nm in d[5]:
            reasons[nm] = reasons.get(nm, 0) + c
    print("stall mix: " + ", ".join("%s %.1f%%" % (k, 100.0 * v / tot) for k, v in sorted(reasons.items(), key=lambda kv: -kv[1])[:8]))
    if len(sys.argv) > 3:
        for part in sys.argv[3].split(","):
            nm, rg = part.split(":")
            lo, hi = map(int, rg.split("-"))
            s = sum(d[0] for d in data if lo <= d[2] <= hi)
            ins = sum(d[4] for d in data if lo <= d[2] <= hi)
            print("  %-22s lines %4d-%4d  %5.1f%% samples  %6.1f M warp-inst" % (nm, lo, hi, 100.0 * s / tot, ins / 1e6))
    for d in sorted(data, key=lambda d: -d[0])[:top]:
        st = ", ".join("%s %d" % (nm.replace("stall_", ""), c) for c, nm in sorted(d[5], reverse=True)[:3])
        print("%5.1f%% %s:%-4d %-110s | %s" % (100.0 * d[0] / tot, d[1], d[2], d[3], st))


if __name__ == "__main__":
    main()
