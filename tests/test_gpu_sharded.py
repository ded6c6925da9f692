"""Multi-GPU path on ONE GPU: all ranks of the sharded pipeline (pointcloudprocess_b200/sharded.py) are run in lock
step in this process -- the plan collectives, the all-to-all redistribution, the halo and boundary exchanges answered by
run_lockstep instead of NCCL -- and the union of the shards' results is compared with the single-device pipeline on the
same cloud: bit for bit, and through the digest bench.py prints at every N."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(ctx, torch, world, n_total, seed, leaf, k, kind, halo):
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    jobs = [sharded.ShardedPipeline(ctx, r, world, n_total, seed=seed, leaf=leaf, k=k, kind=kind, torch=torch, halo=halo)
            for r in range(world)]
    for j in jobs:
        j.generate()
    assert sum(j.n for j in jobs) == n_total
    if world == 1:
        ms = [jobs[0].step_device()]
    else:
        ms = sharded.run_lockstep([j._step_gen(j._p(j.src)) for j in jobs])
    pts = np.concatenate([j.work[:m * 32].cpu().numpy().view(pcp.PACKED_POINT_DTYPE) for j, m in zip(jobs, ms)])
    props = np.concatenate([j.props[:m * 48].cpu().numpy().view(pcp.PROP_DTYPE) for j, m in zip(jobs, ms)])
    dg = sharded.combine_digests([sharded.digest(torch, j.work, j.props, m) for j, m in zip(jobs, ms)])
    return pts, props, dg, jobs, ms


@pytest.mark.parametrize("world,n_total,halo,kind,leaf,k", [(2, 400_000, 1.0, "city", 0.02, 20), (3, 600_000, 0.3, "city", 0.02, 20),
                                                            (4, 400_000, 0.5, "strip", 0.02, 20), (2, 300_000, 1.0, "city", None, 32)])
def test_sharded_pipeline_equals_single_device(ctx, world, n_total, halo, kind, leaf, k):
    import torch
    try:
        p1, f1, d1, _, _ = _run(ctx, torch, 1, n_total, 5, leaf, k, kind, halo)
        p2, f2, d2, jobs, ms = _run(ctx, torch, world, n_total, 5, leaf, k, kind, halo)
    finally:
        torch.cuda.synchronize()
        ctx.set_stream(None)
    stats = [j.stats for j in jobs]
    assert len(p1) == len(p2)  # every voxel belongs to exactly one shard
    o1 = np.lexsort((p1["z"], p1["y"], p1["x"]))
    o2 = np.lexsort((p2["z"], p2["y"], p2["x"]))
    for f in ("x", "y", "z", "rgba", "stamp_id"):
        assert p1[f][o1].tobytes() == p2[f][o2].tobytes(), f  # same cells, same centroids
    for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature"):
        a, b = f1[f][o1], f2[f][o2]
        bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
        assert not bad.any(), (f, int(bad.sum()), stats)
    assert d1 == d2, (d1, d2)  # what bench.py prints as `parity`
    assert sum(s["boundary_queries"] for s in stats) > 0  # the second exchange was exercised
    # the slabs are in x order and the shards' own points add up
    assert sum(s["own"] for s in stats) == len(p1)
    for r, s in enumerate(stats):
        # a flagged query goes to the side(s) it outreaches, not to both neighbours by default
        assert max(s["sent"]) <= s["boundary_queries"] <= sum(s["sent"])
        if 0 < r < world - 1 and s["boundary_queries"] > 50:
            assert sum(s["sent"]) < 1.5 * s["boundary_queries"], s


def test_e2e_and_gather_return_the_device_results(ctx):
    """The host-buffer path (packed records up, packed records down, pipelined over the copy lanes) and the final
    gather deliver exactly the device-resident step's records."""
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    try:
        job = sharded.ShardedPipeline(ctx, 0, 1, 500_000, seed=5, leaf=0.02, k=20, torch=torch)
        job.generate()
        m = job.step_device()
        ref = job.props[:m * 48].cpu().numpy().view(pcp.PROP_DTYPE).copy()
        keep = []
        ms, h2d, d2h, units = job.time_e2e(3, lambda: None, torch.cuda.synchronize, keep=keep)
        assert h2d == 500_000 * 32 and d2h == m * 24 and units == m
        got = ctx.props_unpack(keep[0].view(pcp.PACKED_PROP_DTYPE))
        for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID"):
            assert got[f].tobytes() == ref[f].tobytes(), f
        c_ms, _, c_d2h, _ = job.time_e2e(2, lambda: None, torch.cuda.synchronize, copy_only=True)
        assert c_d2h == d2h and c_ms > 0
        # gather, 3 ranks in lock step: rank 0 ends with every shard's records in slab order
        jobs = [sharded.ShardedPipeline(ctx, r, 3, 300_000, seed=5, leaf=0.02, k=20, torch=torch, halo=0.5) for r in range(3)]
        for j in jobs:
            j.generate()
        msh = sharded.run_lockstep([j._step_gen(j._p(j.src)) for j in jobs])
        res = sharded.run_lockstep([j.gather_gen(mm) for j, mm in zip(jobs, msh)])
        cat, off = res[0]
        assert res[1] is None and res[2] is None and [int(v) for v in off] == list(np.cumsum([0] + [mm * 24 for mm in msh]))
        allp = ctx.props_unpack(cat.cpu().numpy().view(pcp.PACKED_PROP_DTYPE))
        exp = np.concatenate([j.props[:mm * 48].cpu().numpy().view(pcp.PROP_DTYPE) for j, mm in zip(jobs, msh)])
        for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID"):
            assert allp[f].tobytes() == exp[f].tobytes(), f
    finally:
        torch.cuda.synchronize()
        ctx.set_stream(None)


def test_plan_centroid_equals_the_serial_order_centroid(ctx, orc):
    """The chained two-step serial sum (every rank prepares, the carry walks) reproduces compute3DCentroid of the
    whole cloud bit for bit -- on a cloud with UTM-scale coordinates, where a tree sum differs in the last bits."""
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    n = 700_001
    cloud = ctx.synth_cloud("city", n, seed=8)
    c_ref, cnt = orc.centroid(cloud)
    packed = np.zeros(n, dtype=pcp.PACKED_POINT_DTYPE)
    ctx.cloud_pack(cloud, n, packed)
    try:
        world = 5
        jobs = [sharded.ShardedPipeline(ctx, r, world, n, torch=torch, halo=0.5) for r in range(world)]
        for j in jobs:
            j.set_input(packed[j.first:j.first + j.n])
        plans = sharded.run_lockstep([j._plan_gen(j._p(j.src)) for j in jobs])
    finally:
        torch.cuda.synchronize()
        ctx.set_stream(None)
    for P in plans:
        assert P.count == cnt and P.centre.tobytes() == np.asarray(c_ref).tobytes()
        assert P.bounds == plans[0].bounds and (P.min_b == plans[0].min_b).all() and (P.div_b == plans[0].div_b).all()


def test_slabs_follow_the_longer_horizontal_axis(ctx):
    """A strip that runs along y is cut along y (ShardPlan.axis = 1): same results as the single-device run."""
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    n = 300_000
    cloud = ctx.synth_cloud("strip", n, seed=6)
    cloud["x"], cloud["y"] = cloud["y"].copy(), cloud["x"].copy()
    packed = np.zeros(n, dtype=pcp.PACKED_POINT_DTYPE)
    ctx.cloud_pack(cloud, n, packed)
    try:
        res = {}
        for world in (1, 3):
            jobs = [sharded.ShardedPipeline(ctx, r, world, n, leaf=0.02, k=20, torch=torch, halo=0.5) for r in range(world)]
            for j in jobs:
                j.set_input(packed[j.first:j.first + j.n])
            ms = [jobs[0].step_device()] if world == 1 else sharded.run_lockstep([j._step_gen(j._p(j.src)) for j in jobs])
            res[world] = sharded.combine_digests([sharded.digest(torch, j.work, j.props, m) for j, m in zip(jobs, ms)])
            if world > 1:
                assert all(j.plan.axis == 1 and j.stats["axis"] == "y" for j in jobs)
                assert sum(j.stats["boundary_queries"] for j in jobs) > 0
        assert res[1] == res[3]
    finally:
        torch.cuda.synchronize()
        ctx.set_stream(None)
