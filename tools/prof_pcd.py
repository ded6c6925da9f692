"""Loader throughput: pcp_pcd_load (host reader, then upload) against pcp_pcd_load_many (pipelined) on synthetic files."""
import os, sys, time, tempfile
import numpy as np
sys.path.insert(0, "/root/repo")
import pointcloudprocess_b200 as pcp

nfiles = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
ctx = pcp.Context(0)
d = tempfile.mkdtemp(dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
for mode, saver in (("binary_compressed", pcp.io.savePCDFileBinaryCompressed), ("binary", pcp.io.savePCDFileBinary)):
    files = []
    for i in range(nfiles):
        c = pcp.PointCloud(); c.points = ctx.synth_cloud("city", n * nfiles, first=i * n, n=n, seed=6)
        f = os.path.join(d, "%s_%d.pcd" % (mode, i)); assert saver(f, c) == 0; files.append(f)
    size = sum(os.path.getsize(f) for f in files)
    dev = ctx.alloc(n * nfiles * 48)
    t = time.time()
    at = 0
    for f in files:
        c = pcp.PointCloud(); assert pcp.io.loadPCDFile(f, c) == 0
        ctx.memcpy(dev.ptr + at * 48, c.points.ctypes.data, len(c.points) * 48); at += len(c.points)
    ctx.synchronize(); t_host = time.time() - t
    ref = dev.to_host(pcp.POINT_DTYPE, at)
    res = {}
    for threads in (1, 4, 0):
        t = time.time()
        m, off, dense = pcp.io.loadPCDFiles(files, ctx, out=dev, threads=threads)
        ctx.synchronize(); res[threads] = time.time() - t
    got = dev.to_host(pcp.POINT_DTYPE, m)
    same = all(got[f].tobytes() == ref[f].tobytes() for f in ("x", "y", "z", "w", "rgba", "stamp_id"))
    print("%s: %d files x %d points, %.0f MB on disk; host reader + upload %.2f s (%.1f M pts/s); pipelined 1 thread %.2f s, 4 threads %.2f s, all cores (%d) %.2f s (%.1f M pts/s); identical=%s"
          % (mode, nfiles, n, size / 1e6, t_host, m / t_host / 1e6, res[1], res[4], os.cpu_count(), res[0], m / res[0] / 1e6, same))
    dev.free()
    for f in files: os.remove(f)
