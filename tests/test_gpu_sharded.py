"""Multi-GPU path on ONE GPU: all ranks of the sharded pipeline (pointcloudprocess_b200/sharded.py) are run in
lock step in this process, their NCCL exchanges replaced by in-process hand-overs, and the union of the
shards' results is compared with the single-device pipeline on the same cloud -- bit for bit."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _single(ctx, torch, n_total, seed, leaf, k, kind):
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    job = sharded.ShardedPipeline(ctx, 0, 1, n_total, seed=seed, leaf=leaf, k=k, kind=kind, torch=torch)
    job.generate()
    m = job.step_device()
    pts = job.work[:m * 48].cpu().numpy().view(pcp.POINT_DTYPE)
    props = job.props[:m * 48].cpu().numpy().view(pcp.PROP_DTYPE)
    return pts.copy(), props.copy()


def _sharded(ctx, torch, world, n_total, seed, leaf, k, kind, halo):
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    jobs = [sharded.ShardedPipeline(ctx, r, world, n_total // world, seed=seed, leaf=leaf, k=k, kind=kind, torch=torch,
                                    halo=halo) for r in range(world)]
    for j in jobs:
        j.generate()
    assert sum(j.n for j in jobs) == n_total
    ms = sharded.run_lockstep([j._step_gen(j._p(j.cloud)) for j in jobs])
    pts = np.concatenate([j.work[:m * 48].cpu().numpy().view(pcp.POINT_DTYPE) for j, m in zip(jobs, ms)])
    props = np.concatenate([j.props[:m * 48].cpu().numpy().view(pcp.PROP_DTYPE) for j, m in zip(jobs, ms)])
    return pts, props, [j.stats for j in jobs]


@pytest.mark.parametrize("world,n_total,halo,kind", [(2, 400_000, 1.0, "city"), (3, 600_000, 0.3, "city"),
                                                      (4, 400_000, 0.5, "strip")])
def test_sharded_pipeline_equals_single_device(ctx, world, n_total, halo, kind):
    import torch
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    try:
        p1, f1 = _single(ctx, torch, n_total, 5, 0.02, 20, kind)
        p2, f2, stats = _sharded(ctx, torch, world, n_total, 5, 0.02, 20, kind, halo)
    finally:
        torch.cuda.synchronize()
        ctx.set_stream(None)
    assert len(p1) == len(p2)  # every voxel belongs to exactly one shard
    o1 = np.lexsort((p1["z"], p1["y"], p1["x"]))
    o2 = np.lexsort((p2["z"], p2["y"], p2["x"]))
    for f in ("x", "y", "z", "rgba", "stamp_id"):
        assert p1[f][o1].tobytes() == p2[f][o2].tobytes(), f  # same cells, same centroids
    for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature"):
        a, b = f1[f][o1], f2[f][o2]
        bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
        assert not bad.any(), (f, int(bad.sum()), stats)
    assert sum(s["boundary_queries"] for s in stats) > 0  # the second exchange was exercised
    for r, s in enumerate(stats):
        # a flagged query goes to the side(s) it outreaches, not to both neighbours by default
        assert max(s["sent"]) <= s["boundary_queries"] <= sum(s["sent"])
        if 0 < r < world - 1 and s["boundary_queries"] > 50:
            assert sum(s["sent"]) < 1.5 * s["boundary_queries"], s
