"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel: time share per kernel."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
i0 = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
H = rows[i0]; D = rows[i0 + 1:]
kn, mv = H.index("Kernel Name"), H.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in D:
    if len(r) <= mv:
        continue
    name = r[kn].split('(')[0][:60]
    agg[name][0] += 1; agg[name][1] += float(r[mv].replace(',', '')) / 1e6
tot = sum(v[1] for v in agg.values())
print(f"{len(D)} launches, {tot:.1f} ms under ncu (serialised, cold cache)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print(f"{v[1]:9.2f} ms {100 * v[1] / tot:5.1f}%  n={v[0]:4d} avg {v[1] / v[0]:7.3f} ms  {k}")
