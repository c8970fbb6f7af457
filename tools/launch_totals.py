"""Totals per kernel of an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file <csv>)."""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    u = r[ui]
    v = v / 1e3 if u in ("usecond", "us") else v / 1e6 if u in ("nsecond", "ns") else v
    a = agg.setdefault(r[ki][:70], [0, 0.0])
    a[0] += 1; a[1] += v
tot = sum(t for _, t in agg.values())
for k, (n, t) in agg.items():
    print("%-72s n=%4d  %9.3f ms  %5.1f %%" % (k, n, t, 100 * t / tot))
