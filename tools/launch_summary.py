"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel. usage: launch_summary.py <csv> [first-n-to-list]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
h = rows[hi]; k = h.index('Kernel Name'); v = h.index('Metric Value')
agg = collections.defaultdict(lambda: [0, 0.0]); seq = []
for r in rows[hi + 1:]:
    if len(r) <= v: continue
    n = r[k].split('(')[0]; t = float(r[v].replace(',', ''))
    agg[n][0] += 1; agg[n][1] += t; seq.append((n, t))
tot = sum(a[1] for a in agg.values())
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-40s %5d launches %10.3f ms %5.1f%%" % (n, a[0], a[1] / 1e6, 100 * a[1] / tot))
for n, t in seq[:int(sys.argv[2]) if len(sys.argv) > 2 else 0]:
    print("   %-40s %10.1f us" % (n, t / 1e3))
