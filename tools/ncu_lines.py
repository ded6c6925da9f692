"""Summarise an ncu report per CUDA source line: instructions executed and stall samples.
usage: python tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname = None; hdr = None; out = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0] == "Line No": hdr = r; ci = r.index("Instructions Executed"); si = r.index("# Samples"); continue
    if hdr and len(r) > ci and r[0] not in ("", "Line No"):
        try: out.append((int(float(r[ci])), int(float(r[si])), fname, r[0], r[1].strip()[:100]))
        except ValueError: pass
tot = sum(o[0] for o in out); ts = sum(o[1] for o in out)
print("total warp-instructions %d, samples %d" % (tot, ts))
for n, s, f, l, src in sorted(out, reverse=True)[:top]:
    print("%5.1f%% inst %5.1f%% smp  %s:%s  %s" % (100.0 * n / tot, 100.0 * s / max(ts, 1), f, l, src))
