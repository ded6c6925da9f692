#!/usr/bin/env python
"""Times the key sort alone on device-resident keys shaped like the two sorts of the pipeline:
  voxel : 42-bit (k,j,i) cell keys of a city cloud in RANDOM point order (voxel_grid.h:948)
  morton: Morton codes of the voxelised cloud, i.e. of points that arrive in (k,j,i) cell order
    python tools/prof_sort.py [n_millions]          (PCP_SORT_CUB=1 for the library sort)
Run it under `ncu --metrics gpu__time_duration.sum` for the per-kernel split."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import pointcloudprocess_b200 as pcp
    n = int(sys.argv[1]) * 1_000_000 if len(sys.argv) > 1 else 25_000_000
    ctx = pcp.Context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(1)
    i = torch.randint(0, 30000, (n,), device="cuda", generator=g, dtype=torch.int64)
    j = torch.randint(0, 30000, (n,), device="cuda", generator=g, dtype=torch.int64)
    k = (torch.randn(n, device="cuda", generator=g) * 3 + 20).clamp(0, 2999).to(torch.int64)
    vox = i + j * 30000 + k * 30000 * 30000

    def spread(v):
        v = v & 0x1fffff
        v = (v | (v << 32)) & 0x1f00000000ffff
        v = (v | (v << 16)) & 0x1f0000ff0000ff
        v = (v | (v << 8)) & 0x100f00f00f00f00f
        v = (v | (v << 4)) & 0x10c30c30c30c30c3
        v = (v | (v << 2)) & 0x1249249249249249
        return v
    vs = torch.sort(vox).values
    ii, jj, kk = vs % 30000, (vs // 30000) % 30000, vs // (30000 * 30000)
    mort = spread(ii // 13) | (spread(jj // 13) << 1) | (spread(kk // 13) << 2)  # ~0.26 m cells of 0.02 m voxels
    for name, keys, bits in (("voxel", vox, 42), ("morton", mort, 36)):
        work = torch.empty_like(keys)
        vals = torch.empty(n, dtype=torch.int32, device="cuda")
        ts = []
        for rep in range(5):
            work.copy_(keys)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ctx.sort_pairs((work.data_ptr(), n, 8), vals.data_ptr(), bits)
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t0) * 1e3)
        ok = bool((work[1:] >= work[:-1]).all())
        print("%s: n=%d bits=%d  %.3f ms (min of %s)  sorted=%s" % (name, n, bits, min(ts), ["%.2f" % t for t in ts], ok))
    ctx.close()


if __name__ == "__main__":
    main()
