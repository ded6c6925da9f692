#!/usr/bin/env python
"""Small shapes of every kernel family, for compute-sanitizer (one tool per gpurun call):
    compute-sanitizer --tool memcheck|racecheck|initcheck python tools/sanitize_run.py
No torch: only the C ABI.  Results are checked for plausibility only (the parity tests hold the oracle comparison)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import pointcloudprocess_b200 as pcp
    ctx = pcp.Context(0)
    rng = np.random.default_rng(1)
    done = []
    # kNN + features: dense kernel, second dense pass, far kernel, streaming fallback (duplicates), batched external queries
    for kind, n in (("city", 30000), ("forest", 12000), ("strip", 15000)):
        cloud = ctx.synth_cloud(kind, n, seed=3)
        ix = pcp.Index(ctx, cloud)
        r = ix.features_knn(20, want_idx=True)
        assert (r["knn_idx"][:, 0] >= 0).all()
        q = np.stack([cloud["x"], cloud["y"], cloud["z"]], 1)[:300] + rng.normal(0, 0.3, (300, 3))
        ix.knn(q, 20)
        ix.knn(q[:5], 7)
        rr = ix.features_radius(0.3)
        assert rr["counts"].min() >= 1
        ix.radius(q[:50], 0.4)
        ix.features_rpca(20, 0.99, 0.5, 1)
        ix.free()
        done.append("knn/radius/rpca %s" % kind)
    dup = pcp.make_cloud(np.repeat(rng.uniform(0, 2, (60, 3)) + 1000.0, 50, axis=0))
    ix = pcp.Index(ctx, dup)
    ix.features_knn(20)
    ix.features_knn(40)
    ix.free()
    done.append("mass ties (streaming kernel)")
    # voxel grid / remove_duplicate (48- and 32-byte records, in place and out of place), serial-order centroid, min/max
    cloud = ctx.synth_cloud("city", 40000, seed=5)
    ctx.voxel_grid(cloud, 0.05, want_cells=True)
    a = ctx.remove_duplicate(cloud, 0.02)
    packed = np.zeros(len(cloud), dtype=pcp.PACKED_POINT_DTYPE)
    ctx.cloud_pack(cloud, len(cloud), packed)
    out = np.zeros(len(cloud), dtype=pcp.PACKED_POINT_DTYPE)
    m = ctx.remove_duplicate_to(packed, 0.02, out)
    assert m == len(a)
    ctx.centroid(cloud)
    ctx.minmax(cloud)
    ctx.cloud_stats(cloud)
    big = ctx.synth_cloud("strip", 70000, seed=6)
    s3, cnt = ctx.serial_sum(big, np.zeros(3), 0)
    done.append("voxel / dedup / centroid")
    # sort, scan, partition, pack
    k = rng.integers(0, 1 << 40, 50001, dtype=np.uint64)
    v = np.arange(len(k), dtype=np.uint32)
    ctx.sort_pairs(k, v, 40)
    assert (np.diff(k.astype(np.int64)) >= 0).all()
    k32 = rng.integers(0, 1 << 20, 9000, dtype=np.uint32)
    ctx.sort_pairs(k32, None, 20)
    ctx.exclusive_sum(rng.integers(0, 9, 70000, dtype=np.uint32))
    bounds = np.array([0, 300, 700, 100000], dtype=np.int32)
    outp = np.zeros(len(packed), dtype=pcp.PACKED_POINT_DTYPE)
    c, _ = ctx.centroid(cloud)
    counts = ctx.partition_columns(packed, 0, 0.02, c[0], -2000, bounds, outp)
    assert counts.sum() == len(packed)
    props = np.zeros(1000, dtype=pcp.PROP_DTYPE)
    pk = np.zeros(1000, dtype=pcp.PACKED_PROP_DTYPE)
    ctx.props_pack(props, 1000, pk)
    ctx.synchronize()
    done.append("sort / scan / partition / pack")
    # region growing: two blocks and the default grid
    cloud = ctx.synth_cloud("city", 25000, seed=7)
    dd = ctx.remove_duplicate(cloud, 0.04)
    ix = pcp.Index(ctx, dd)
    pr = ix.features_knn(20, want_eig=False, want_lps=False)["props"]
    seg = ix.region_growing(pr.copy(), 50, 0.5, 0.94)
    ix.free()
    done.append("region growing (%d segments, PCP_RG_BLOCKS=%s)" % (len(seg["planes"]), os.environ.get("PCP_RG_BLOCKS", "default")))
    # plane fit of segments, shard selection primitives, merge
    off = np.array([0, 10, 30, 31, 100], dtype=np.int64)
    ctx.plane_fit(cloud[:100], off)
    sel = np.zeros(len(cloud), dtype=pcp.POINT_DTYPE)
    ctx.select_range(cloud, 0, float(cloud["x"].min()), float(cloud["x"].min()) + 3.0, sel)
    done.append("plane fit / select")
    ctx.close()
    print("sanitize_run ok: " + "; ".join(done))


if __name__ == "__main__":
    main()
