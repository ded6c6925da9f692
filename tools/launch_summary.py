"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/launch_summary.py launches.csv"""
import csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
agg = {}
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki]); name = re.sub(r"^void ", "", name)
    v = float(r[vi].replace(",", "")); v = v / 1e6 if r[ui] in ("ns", "nsecond") else (v / 1e3 if r[ui].startswith("us") else v)
    e = agg.setdefault(name, [0, 0.0]); e[0] += 1; e[1] += v
tot = sum(e[1] for e in agg.values())
print("%d launches, %.2f ms under ncu (cold caches, serialised)" % (sum(e[0] for e in agg.values()), tot))
for name, (n, ms) in sorted(agg.items(), key=lambda x: -x[1][1])[:25]:
    print("%6.2f%% %9.3f ms %5d  %s" % (100 * ms / tot, ms, n, name[:110]))
