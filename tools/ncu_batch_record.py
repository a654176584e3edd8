#!/usr/bin/env python
"""tools/ncu_batch_record.py <raw.csv> <qps in the captured launch> [commit] -- from `ncu -i X.ncu-rep --page raw --csv`
of one k_batch_solve launch: DRAM bytes and executed warp instructions per QP and the headline utilisation figures,
written to profiles/c5_ncu.json with the commit the capture was taken at.  bench.py reads that file for
`roofline.traffic` and `roofline_issue` (a measured record with its provenance instead of a constant in the code)."""
import csv
import json
import os
import subprocess
import sys

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "inst": 1.0, "cycle": 1.0, "%": 1.0, "": 1.0,
         "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    path, qps = sys.argv[1], int(sys.argv[2])
    commit = sys.argv[3] if len(sys.argv) > 3 else subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = dict((h, i) for i, h in enumerate(hdr))
    row = [r for r in body if "k_batch_solve" in r[col["Kernel Name"]]][0]

    def get(name, default=None):
        if name not in col:
            return default
        try:
            return float(row[col[name]].replace(",", "")) * SCALE.get(units[col[name]], 1.0)
        except ValueError:
            return default
    rd, wr = get("dram__bytes_read.sum", 0.0), get("dram__bytes_write.sum", 0.0)
    inst = get("smsp__inst_executed.sum") or get("sm__inst_executed.sum")
    out = {"commit": commit, "kernel": "k_batch_solve", "qps_in_launch": qps, "source": os.path.basename(path),
           "duration_ms": (get("gpu__time_duration.sum") or 0.0) * 1e3,
           "dram_bytes_per_qp": (rd + wr) / qps, "dram_read_bytes": rd, "dram_write_bytes": wr,
           "warp_inst_per_qp": inst / qps if inst else None,
           "issue_slots_busy_pct": get("sm__inst_issued.avg.pct_of_peak_sustained_active") or get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
           "warps_active_pct": get("sm__warps_active.avg.pct_of_peak_sustained_active"),
           "fp64_pipe_pct": get("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
           "registers_per_thread": get("launch__registers_per_thread"),
           "shared_mem_per_block": (get("launch__shared_mem_per_block_dynamic") or 0) + (get("launch__shared_mem_per_block_static") or 0)}
    json.dump(out, open(os.path.join(ROOT, "profiles", "c5_ncu.json"), "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
