"""Aggregate an ncu report's executed instructions and stall samples by file and by line ranges.
usage: python tools/ncu_ranges.py report.ncu-rep file:lo-hi:name ..."""
import csv, subprocess, sys
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname = None; agg = {}; smp = {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No": ci = r.index("Instructions Executed"); si = r.index("# Samples"); continue
    if len(r) > 8 and r[0] not in ("", "Line No"):
        try:
            k = (fname, int(r[0])); n = int(float(r[ci])); s = int(float(r[si]))
        except ValueError: continue
        agg[k] = agg.get(k, 0) + n; smp[k] = smp.get(k, 0) + s
tot = sum(agg.values()); ts = sum(smp.values())
print("total warp-instructions %d, samples %d" % (tot, ts))
byfile = {}
for (f, l), n in agg.items():
    e = byfile.setdefault(f, [0, 0]); e[0] += n; e[1] += smp.get((f, l), 0)
for f, (n, s) in sorted(byfile.items(), key=lambda x: -x[1][0])[:12]: print("%-32s %5.1f%% inst %5.1f%% smp" % (f, 100 * n / tot, 100 * s / ts))
for a in sys.argv[2:]:
    f, r, name = a.split(":"); lo, hi = map(int, r.split("-"))
    n = sum(v for (ff, l), v in agg.items() if ff == f and lo <= l < hi); s = sum(v for (ff, l), v in smp.items() if ff == f and lo <= l < hi)
    print("  %-14s %4d-%4d %5.1f%% inst %5.1f%% smp  %s" % (f, lo, hi, 100 * n / tot, 100 * s / ts, name))
