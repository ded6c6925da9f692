"""Generates tests/golden/*.json in THIS container (needs cv2, scipy and, for LZF, oracle/_ref built from
/root/reference/lzf.cpp).  The fixtures are committed; the GPU box never runs this script.

  eigen_cv2.json     : cv2.eigen outputs (same Jacobi lineage as the reference's cvEigenVV,
                       calculate_feature.cpp:165) for un-normalised scatter matrices of random neighbourhoods
  knn_scipy.json     : scipy.spatial.cKDTree(leafsize=15).query index sets on a seeded cloud (independent
                       double-precision exact kNN; kd_tree.h:795-797 uses the same leaf size)
  lzf_ref.json       : byte streams produced by the reference's own lzfCompress (lzf.cpp:86-285)
"""
import base64
import json
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")


def hexd(a):
    return [struct.pack("<d", float(v)).hex() for v in np.asarray(a, dtype=np.float64).ravel()]


def main():
    import cv2
    from scipy.spatial import cKDTree
    from oracle import oracle as orc
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(20261104)
    cases = []
    for t in range(200):
        h = int(rng.integers(4, 40))
        scale = rng.uniform(0.001, 2.0, size=(3, 1)) * (10.0 ** rng.integers(-3, 2))
        X = rng.normal(size=(3, h)) * scale
        X -= X.mean(1, keepdims=True)
        S = np.zeros((3, 3))
        for i in range(3):
            for j in range(3):
                s = 0.0
                for n in range(h):
                    s += X[i, n] * X[j, n]
                S[i, j] = s
        ok, w, v = cv2.eigen(S)
        cases.append({"S": hexd(S), "W": hexd(w), "V": hexd(v)})
    # a few structured ones: diagonal, repeated eigenvalues, rank-1, zeros
    for S in (np.diag([3.0, 2.0, 1.0]), np.diag([1.0, 1.0, 1.0]), np.outer([1, 2, 3], [1, 2, 3]).astype(float),
              np.zeros((3, 3)), np.array([[2.0, 1, 0], [1, 2, 0], [0, 0, 2.0]])):
        ok, w, v = cv2.eigen(S)
        cases.append({"S": hexd(S), "W": hexd(w), "V": hexd(v)})
    json.dump({"source": "cv2.eigen, OpenCV %s" % cv2.__version__, "cases": cases},
              open(os.path.join(OUT, "eigen_cv2.json"), "w"))

    P = rng.uniform(0, 10, size=(4000, 3)) + np.array([483189.100569, 3371741.645173, 430.094726])
    d, i = cKDTree(P, leafsize=15).query(P[:400], 20)
    json.dump({"source": "scipy.spatial.cKDTree(leafsize=15).query(k=20)", "seed": 20261104,
               "points": hexd(P), "idx": i.tolist()}, open(os.path.join(OUT, "knn_scipy.json"), "w"))

    ref = orc.ref_lzf()
    if ref is None:
        print("oracle/_ref/liblzf_ref.so missing: lzf_ref.json not regenerated")
        return
    import ctypes as C
    blobs = []
    srcs = [b"a", b"ab", b"abc", b"aaaa" * 100, bytes(range(256)) * 8,
            rng.integers(0, 4, 5000, dtype=np.uint8).tobytes(),
            np.repeat(rng.normal(size=300).astype(np.float64), 3).tobytes(),
            rng.integers(0, 256, 3000, dtype=np.uint8).tobytes()]
    for s in srcs:
        out = C.create_string_buffer(len(s) * 2 + 64)
        n = ref.ref_lzf_compress(s, len(s), out, len(out))
        blobs.append({"raw": base64.b64encode(s).decode(), "lzf": base64.b64encode(out.raw[:n]).decode()})
    json.dump({"source": "/root/reference/lzf.cpp lzfCompress via oracle/_ref/liblzf_ref.so", "cases": blobs},
              open(os.path.join(OUT, "lzf_ref.json"), "w"))


if __name__ == "__main__":
    main()
