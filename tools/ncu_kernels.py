#!/usr/bin/env python
"""tools/ncu_kernels.py <raw.csv> [...] -- per-kernel table from `ncu -i X.ncu-rep --page raw --csv`:
launches captured, mean duration, DRAM bytes per launch, achieved DRAM GB/s (bytes / duration), DRAM
throughput % of peak, FP64 pipe % (DMMA shares the FP64 pipe's issue port on sm_100), DMMA active
cycles, warps active %.  Units in row 2 of the CSV are honoured (byte / Kbyte / Mbyte / Gbyte, us / ms)."""
import csv
import re
import sys
from collections import OrderedDict

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6,
         "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}
WANT = OrderedDict([
    ("dur", "gpu__time_duration.sum"),
    ("rd", "dram__bytes_read.sum"),
    ("wr", "dram__bytes_write.sum"),
    ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("fp64_pct", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
    ("fp64_rt", "TPC.TriageCompute.sm__pipe_fp64_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"),
    ("dmma", "SM_C.TriageCompute.smsp__pipe_tensor_subpipe_dmma_cycles_active.avg"),
    ("cyc", "sm__cycles_elapsed.avg"),
    ("warps", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("issue", "sm__inst_issued.avg.pct_of_peak_sustained_active"),
])


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return float("nan")


def main(paths):
    for path in paths:
        rows = list(csv.reader(open(path)))
        rows = [r for r in rows if len(r) > 10]
        hdr, units, body = rows[0], rows[1], rows[2:]
        col = {h: i for i, h in enumerate(hdr)}
        kn = col["Kernel Name"]
        agg = OrderedDict()
        for r in body:
            name = re.sub(r"\(.*", "", r[kn]).replace("void ", "")
            a = agg.setdefault(name, {"n": 0, **{k: 0.0 for k in WANT}})
            a["n"] += 1
            for k, m in WANT.items():
                if m in col:
                    v = num(r[col[m]]) * SCALE.get(units[col[m]], 1.0)
                    if v == v:
                        a[k] += v
        print("## %s\n" % path)
        print("| kernel | launches | avg us | DRAM read+write per launch | achieved GB/s | DRAM %% of peak | FP64 pipe %% | DMMA cycles / elapsed | warps active %% | issue %% |")
        print("|---|---|---|---|---|---|---|---|---|---|")
        for name, a in agg.items():
            n = a["n"]
            dur = a["dur"] / n
            byt = (a["rd"] + a["wr"]) / n
            fp = a["fp64_pct"] / n if a["fp64_pct"] else a["fp64_rt"] / n
            dm = (a["dmma"] / a["cyc"]) if a["cyc"] else float("nan")
            print("| `%s` | %d | %.1f | %.2f MB | %.0f | %.1f | %.1f | %.3f | %.1f | %.1f |" % (
                name, n, dur, byt / 1e6, byt / dur / 1e3 if dur else 0.0, a["dram_pct"] / n, fp, dm, a["warps"] / n, a["issue"] / n))
        print()


if __name__ == "__main__":
    main(sys.argv[1:])
