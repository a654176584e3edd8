#!/usr/bin/env python
"""tools/ncu_c3_record.py <raw.csv> <newton steps captured> [commit] -- from `ncu --page raw --csv` of the k_fwd / k_bwd
(and factor) launches of a C3 run: DRAM bytes per Newton step of the triangular solves and the FP64 pipe figures of the
frontal kernels, written to profiles/c3_ncu.json (bench.py reads `solve_dram_bytes_per_newton` for roofline.traffic)."""
import csv
import json
import os
import subprocess
import sys
from collections import defaultdict

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    path, newton = sys.argv[1], float(sys.argv[2])
    commit = sys.argv[3] if len(sys.argv) > 3 else subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = dict((h, i) for i, h in enumerate(hdr))

    def val(r, name):
        if name not in col:
            return 0.0
        try:
            return float(r[col[name]].replace(",", "")) * SCALE.get(units[col[name]], 1.0)
        except ValueError:
            return 0.0
    agg = defaultdict(lambda: dict(launches=0, dram=0.0, dur=0.0, fp64=0.0))
    for r in body:
        name = r[col["Kernel Name"]].split("(")[0].replace("void ", "")
        a = agg[name]
        a["launches"] += 1
        a["dram"] += val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
        a["dur"] += val(r, "gpu__time_duration.sum")
        a["fp64"] += val(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active")
    solve = sum(a["dram"] for k, a in agg.items() if k.startswith("k_fwd") or k.startswith("k_bwd"))
    out = {"commit": commit, "source": os.path.basename(path), "newton_steps_captured": newton,
           "solve_dram_bytes_per_newton": solve / newton,
           "kernels": dict((k, {"launches": a["launches"], "dram_bytes_per_launch": a["dram"] / a["launches"], "avg_us": a["dur"] / a["launches"] * 1e6,
                                "fp64_pipe_pct": a["fp64"] / a["launches"]}) for k, a in sorted(agg.items()))}
    json.dump(out, open(os.path.join(ROOT, "profiles", "c3_ncu.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
