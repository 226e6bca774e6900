"""Dev tool: aggregate an ncu --csv launch list (gpu__time_duration.sum) per kernel."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
h = rows[hdr]; ki = h.index('Kernel Name'); vi = h.index('Metric Value'); ui = h.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= vi: continue
    n = r[ki].split('(')[0]; v = float(r[vi].replace(',', '')); u = r[ui]
    v = v / 1e3 if u == 'ns' else v * 1e3 if u == 'ms' else v * 1e6 if u == 's' else v
    agg[n][0] += 1; agg[n][1] += v
tot = sum(t for _, t in agg.values())
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{n:44s} {c:5d} {t:12.1f} us {100*t/tot:6.2f}%  {t/c:10.1f} us/launch")
