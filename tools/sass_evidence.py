"""Per kernel of libpcp_cuda.so: how many SASS instructions of the Blackwell-specific kinds it holds.
usage: python tools/sass_evidence.py > profiles/rNN_sass_evidence.txt   (needs cuobjdump and c++filt; no GPU)"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "pointcloudprocess_b200", "libpcp_cuda.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
pat = {"FFMA2": r"\bFFMA2\b", "FADD2": r"\bFADD2\b", "FMUL2": r"\bFMUL2\b", "LDG.256": r"LDG\.E\.(ENL2\.)?256",
       "STG.256": r"STG\.E\.(ENL2\.)?256", "REDUX": r"\bREDUX\b", "MATCH": r"\bMATCH\b", "ATOMS": r"\bATOMS\b",
       "instr": r"^\s+/\*[0-9a-f]{4,6}\*/"}
fn = None
cnt = collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        continue
    if fn:
        for k, p in pat.items():
            if re.search(p, line):
                cnt[fn][k] += 1
names = list(cnt)
dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
rows = [(d, cnt[n]) for n, d in zip(names, dem)
        if sum(cnt[n][k] for k in ("FFMA2", "FADD2", "LDG.256", "STG.256", "REDUX", "MATCH")) > 0 and "cub::" not in d]
rows.sort(key=lambda r: -r[1]["instr"])
print("# SASS of libpcp_cuda.so (cuobjdump -sass, sm_100a): instructions per kernel, and how many of them are packed FP32")
print("# (FFMA2 / FADD2 / FMUL2: two candidates per lane and instruction in the dense search), 256-bit global loads and")
print("# stores (LDG.E.256 / STG.E.256: one per 32-byte point record), warp reductions and match.any (the radix sort and")
print("# the scans), shared-memory atomics (the histograms).  tools/sass_evidence.py; CUB's kernels (behind PCP_SORT_CUB) left out.")
cols = ("instr", "FFMA2", "FADD2", "FMUL2", "LDG.256", "STG.256", "REDUX", "MATCH", "ATOMS")
print("%-96s" % "kernel" + "".join("%8s" % c for c in cols))
for d, c in rows:
    print("%-96s" % d[:96] + "".join("%8d" % c[k] for k in cols))
