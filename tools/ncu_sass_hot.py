"""Executed-SASS view of an ncu report (captured with --import-source on): the opcode mix weighted by executed
instructions, then the SASS lines above a threshold in address order (loops show up as runs of equal counts).
usage: ncu_sass_hot.py report.ncu-rep [min_executed_per_line]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 20_000_000
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
hdr = None
lines = []
for r in csv.reader(txt.splitlines()):
    if "Source" in r and "Instructions Executed" in r:
        hdr, ci, si = r, r.index("Instructions Executed"), r.index("Source")
        ai = r.index("Address") if "Address" in r else 0
        continue
    if hdr and len(r) == len(hdr):
        try:
            lines.append((int(float(r[ci])), r[ai][-5:], r[si]))
        except ValueError:
            pass
tot = sum(n for n, _, _ in lines)
print("%d SASS lines, %d instructions executed" % (len(lines), tot))
mix = collections.Counter()
for n, _, s in lines:
    t = s.split()
    op = t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")
    mix[op.split(".")[0]] += n
for op, n in mix.most_common(24):
    print("  %-10s %5.1f%%" % (op, 100.0 * n / tot))
for n, a, s in lines:
    if n >= thr:
        print("%12d  %s  %s" % (n, a, s[:100]))
