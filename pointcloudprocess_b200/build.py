"""Build recipe for libpcp_cuda.so: every csrc/*.cu compiled for sm_100a only, linked in-tree.

nvcc cross-compiles without a GPU; the resulting .so travels to the GPU box with the repo snapshot.
-fmad=false: no FMA contraction, so squared distances, scatter sums and the Jacobi rotations round
exactly like the reference's non-fused x64 build (and like the oracle's -ffp-contract=off).
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libpcp_cuda.so")
OBJ = os.path.join(HERE, "build")
NVCC = os.environ.get("PCP_NVCC", "/usr/local/cuda/bin/nvcc")
HOSTCXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
         "-Xcompiler", "-fPIC,-fvisibility=hidden,-O3", "-ccbin", HOSTCXX, "--expt-relaxed-constexpr"]
# measurement builds only (e.g. PCP_NVCC_DEFS="PCP_FD_REASONS"); objects are cached by time stamp, not by flags: use force
FLAGS += ["-D" + d for d in os.environ.get("PCP_NVCC_DEFS", "").split()]


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "pcp_cuda.h"))
    return hs


def _compile(src, verbose):
    obj = os.path.join(OBJ, src[:-3] + ".o")
    srcp = os.path.join(CSRC, src)
    newest = max([os.path.getmtime(srcp)] + [os.path.getmtime(h) for h in _headers()])
    if os.path.exists(obj) and os.path.getmtime(obj) > newest:
        return obj, ""
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", srcp, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    return obj, r.stderr


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in os.listdir(OBJ):
            os.remove(os.path.join(OBJ, f))
    srcs = _sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile(s, verbose), srcs))
    objs = [o for o, _ in results]
    log = "".join(l for _, l in results)
    if (not os.path.exists(OUT)) or any(os.path.getmtime(o) > os.path.getmtime(OUT) for o in objs):
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", HOSTCXX, "-o", OUT] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    if verbose:
        sys.stderr.write(log)
    return OUT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
