"""Per-kernel mean duration from an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]; ki = h.index("Kernel Name"); vi = h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[hdr + 2:]:
    if len(r) <= vi:
        continue
    a = agg.setdefault(r[ki][:70], [0, 0.0]); a[0] += 1; a[1] += float(r[vi].replace(",", ""))
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print("%9.3f ms/launch x%3d %s" % (t / c / 1e6, c, n))
