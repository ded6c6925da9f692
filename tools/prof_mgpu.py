#!/usr/bin/env python
"""pcp_mgpu_features_knn on every GPU of the box: host cloud in, host results out, wall clock, and the bench's digest
(for the city cloud, seed 5, leaf 0.02, k 20 it must equal the `parity.digest` bench.py prints at the same size).
    python tools/prof_mgpu.py [million points] [ndev]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    n = (int(sys.argv[1]) if len(sys.argv) > 1 else 50) * 1_000_000
    ctx = pcp.Context(0)
    ndev = int(sys.argv[2]) if len(sys.argv) > 2 else ctx.L.pcp_device_count()
    cloud = np.zeros(n, dtype=pcp.PACKED_POINT_DTYPE)
    chunk = 10_000_000
    d48, d32 = ctx.alloc(chunk * 48), ctx.alloc(chunk * 32)
    for at in range(0, n, chunk):
        c = min(chunk, n - at)
        ctx.synth_cloud("city", n, first=at, n=c, seed=5, out=d48.ptr)
        ctx.cloud_pack(d48.ptr, c, d32.ptr)
        ctx.memcpy(cloud[at:at + c].ctypes.data, d32.ptr, c * 32)
    d48.free()
    d32.free()
    ctx.close()
    mg = pcp.MultiGpu(ndev)
    for rep in range(3):
        t0 = time.perf_counter()
        pts, props = mg.features_knn(cloud, 0.02, 20, 2.0)
        dt = time.perf_counter() - t0
        print("pcp_mgpu_features_knn, %d GPUs, %d points -> %d: %.1f ms host to host (%.1f M points/s; pageable host memory)"
              % (ndev, n, len(pts), dt * 1e3, len(pts) / dt / 1e6), flush=True)
    mg.close()
    dg = sharded.digest(torch, torch.from_numpy(pts.view(np.uint8).reshape(-1)), torch.from_numpy(props.view(np.uint8).reshape(-1)), len(pts))
    print("digest", dg)


if __name__ == "__main__":
    main()
