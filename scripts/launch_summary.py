"""Per-kernel totals of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv): count, total ms, share.
usage: launch_summary.py <launches.csv>"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
h = rows[hi]; kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hi + 1:]:
    if len(r) <= mv: continue
    v = float(r[mv].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[r[mu]]
    agg[r[kn][:70]][0] += 1; agg[r[kn][:70]][1] += v
tot = sum(v[1] for v in agg.values())
print("%-72s %6s %12s %7s" % ("kernel", "n", "total ms", "share"))
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-72s %6d %12.3f %6.1f%%" % (k, n, t, 100 * t / tot))
