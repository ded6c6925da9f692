"""Aggregate an ncu report's executed instructions by file and by marker comments ('// ----') in a source file (argv[2], default knn_fast.cuh)."""
import csv, subprocess, sys, os
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname = None; agg = {}; smp = {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split('/')[-1]; continue
    if len(r) > 8 and r[0] == "Line No": ci = r.index("Instructions Executed"); si = r.index("# Samples"); continue
    if len(r) > 8 and r[0] not in ("", "Line No"):
        try:
            k = (fname, int(r[0])); agg[k] = agg.get(k, 0) + int(float(r[ci])); smp[k] = smp.get(k, 0) + int(float(r[si]))
        except ValueError: pass
tot = sum(agg.values()); ts = sum(smp.values())
byfile = {}
for (f, l), n in agg.items(): byfile[f] = byfile.get(f, 0) + n
for f, n in sorted(byfile.items(), key=lambda x: -x[1]): print("%-30s %5.1f%%" % (f, 100 * n / tot))
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
target = sys.argv[2] if len(sys.argv) > 2 else "knn_fast.cuh"
src = open(os.path.join(root, "pointcloudprocess_b200/csrc", target)).read().split('\n')
marks = [(i + 1, l.strip()[:70]) for i, l in enumerate(src) if l.strip().startswith("// ----") or l.startswith("template") or l.startswith("static __global__") or l.startswith("__device__")]
for i, (l0, name) in enumerate(marks):
    l1 = marks[i + 1][0] if i + 1 < len(marks) else 10 ** 9
    n = sum(v for (f, l), v in agg.items() if f == target and l0 <= l < l1)
    s = sum(v for (f, l), v in smp.items() if f == target and l0 <= l < l1)
    if n: print("  L%4d %5.1f%% inst %5.1f%% smp  %s" % (l0, 100 * n / tot, 100 * s / ts, name))
