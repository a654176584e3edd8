#!/usr/bin/env python
"""tools/ncu_launches.py LAUNCHES.csv [top] -- per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list"""
import collections
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
h = rows[hi]; kn = h.index('Kernel Name'); mv = h.index('Metric Value'); mu = h.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    v = float(r[mv].replace(',', '')); u = r[mu]
    v = v / 1e3 if u == 'ns' else v * 1e3 if u == 'ms' else v
    name = r[kn].split('(')[0]
    agg[name][0] += 1; agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print("| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 20]:
    print("| `%s` | %d | %.0f | %.1f%% | %.1f |" % (k, v[0], v[1], 100 * v[1] / tot, v[1] / v[0]))
print("\nTotal %.0f us over %d launches." % (tot, sum(v[0] for v in agg.values())))
