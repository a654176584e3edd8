#!/bin/bash
# ncu --set full with source on the k_big_panel launches of one C1 factorization, dense sampling; stall mix over the LDL' lines
O=gpurun_out
LO=${1:-575}; HI=${2:-640}
python tools/factor_check.py c1 > $O/ncu_panel_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_big_panel$ -s 30 -c 30 -f -o $O/panel python tools/factor_check.py c1 > $O/ncu_panel.log 2>&1
echo "ncu rc=$?"
python - $O/panel.ncu-rep $LO $HI <<'PY' > $O/panel_ldl_stalls.txt 2>&1
import csv, subprocess, sys
rep, lo, hi = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = None; tot = {}; allr = {}; nsamp = 0; nall = 0; lines = {}
for r in rows:
    if len(r) > 10 and r[0] == "Line No":
        h = r; si = h.index("# Samples"); bi = h.index("stall_barrier"); continue
    if h and len(r) > 10 and r[0].isdigit():
        ln = int(r[0])
        try: s = int(r[si])
        except ValueError: continue
        nall += s
        for j in range(17):
            c = int(r[bi + j] or 0)
            allr[h[bi + j]] = allr.get(h[bi + j], 0) + c
            if lo <= ln <= hi: tot[h[bi + j]] = tot.get(h[bi + j], 0) + c
        if lo <= ln <= hi:
            nsamp += s; lines[ln] = lines.get(ln, 0) + s
print("all lines: %d samples; " % nall + ", ".join("%s %.1f%%" % (k.replace("stall_", ""), 100.0 * v / max(nall, 1)) for k, v in sorted(allr.items(), key=lambda kv: -kv[1])[:8]))
print("lines %d-%d (diagonal-block chain): %d samples; " % (lo, hi, nsamp) + ", ".join("%s %.1f%%" % (k.replace("stall_", ""), 100.0 * v / max(nsamp, 1)) for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]))
for ln, s in sorted(lines.items(), key=lambda kv: -kv[1])[:12]: print("   line %d: %d samples" % (ln, s))
PY
cat $O/panel_ldl_stalls.txt
