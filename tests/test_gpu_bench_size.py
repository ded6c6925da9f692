"""Parity AT THE SIZES THE BENCH TIMES: the 25 M-point city shard through the bench's own step
(second dense pass, ~1.2 M far queries, the streaming fallback), configs[1] in full (10 M-point strip, leaf 0.05: the
64-bit-key path), configs[2] sampled (20 M-point forest, r = 0.5) and k = 32 on the 25 M shard -- against the oracle's
own kd-tree (leaf 15) over the FULL cloud.  Voxel outputs are compared in full; neighbour lists and records on
samples of queries that include the ones with the largest k-th distances (far kernel / streaming kernel territory).
Bit-exact throughout.  About two minutes on the GPU box's host cores."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _fields_equal(a, b, fields):
    assert len(a) == len(b)
    for f in fields:
        assert a[f].tobytes() == b[f].tobytes(), f


def _fit_rows(orc, xyz, idx_rows):
    """calculate_plan_parameter_h_points on every row of neighbour indices -> (normals float32 n x 3, Distance, curvature)."""
    out = np.zeros((len(idx_rows), 5), dtype=np.float64)
    for j, row in enumerate(idx_rows):
        pl = orc.plane_fit(xyz[row])
        out[j] = (pl["normal_x"], pl["normal_y"], pl["normal_z"], pl["Distance"], pl["curvature"])
    return out


def _check_props(props, sel, fit):
    got = np.stack([props["normal_x"][sel], props["normal_y"][sel], props["normal_z"][sel]], 1).astype(np.float64)
    assert np.array_equal(got, fit[:, :3])
    # LAS_POINT_PROPERTY keeps PlanSegment's float Distance / curvature widened to double (calculate_feature.cpp:344-345)
    assert np.array_equal(props["Distance"][sel], fit[:, 3]) and np.array_equal(props["curvature"][sel], fit[:, 4], equal_nan=True)


@pytest.fixture(scope="module")
def city25(ctx, orc):
    """The bench's 25 M-point shard: cloud, the bench step's results, the oracle's voxelised cloud and its kd-tree."""
    import torch
    import pointcloudprocess_b200 as pcp
    from pointcloudprocess_b200 import sharded
    n = 25_000_000
    try:
        job = sharded.ShardedPipeline(ctx, 0, 1, n, seed=5, leaf=0.02, k=20, torch=torch)
        job.generate()
        m = job.step_device()
        work = job.work[:m * 32].cpu().numpy().view(pcp.PACKED_POINT_DTYPE).copy()
        props = job.props[:m * 48].cpu().numpy().view(pcp.PROP_DTYPE).copy()
        job.close()
        del job
        torch.cuda.empty_cache()
    finally:
        torch.cuda.synchronize()
        ctx.set_stream(None)
    cloud = ctx.synth_cloud("city", n, seed=5)          # the same points the pipeline generated (host copy)
    ref, _ = orc.remove_duplicate(cloud, 0.02)
    del cloud
    tree = orc.KdTree(ref)
    return {"m": m, "work": work, "props": props, "ref": ref, "tree": tree}


def test_bench_step_voxel_outputs_in_full(city25):
    """remove_duplicate(0.02) of the 25 M shard: every cell, centroid, colour and stamp, in order."""
    assert city25["m"] == len(city25["ref"])
    _fields_equal(city25["work"], city25["ref"], ("x", "y", "z", "rgba", "stamp_id"))


def test_bench_step_neighbours_and_records_sampled(ctx, orc, city25):
    """220 k queries of the bench step against the oracle's kd-tree over all 24.7 M voxelised points."""
    import pointcloudprocess_b200 as pcp
    ref, tree, m = city25["ref"], city25["tree"], city25["m"]
    k = 20
    ix = pcp.Index(ctx, ref)
    kth = np.zeros(m, dtype=np.float32)
    got = ix.features_knn(k, want_eig=False, want_lps=False, want_idx=True, kth_d2=kth)
    ix.free()
    # the standalone call and the bench's step agree on every record
    for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature"):
        assert np.array_equal(got["props"][f], city25["props"][f], equal_nan=True), f
    rng = np.random.default_rng(7)
    sel = np.unique(np.concatenate([rng.integers(0, m, 200_000), np.argsort(kth)[-20_000:]]))  # random + the sparsest
    q = np.stack([ref["x"][sel], ref["y"][sel], ref["z"][sel]], 1)
    ridx, rd2, rfound = tree.knn(q, k)
    assert (rfound == k).all()
    assert np.array_equal(got["knn_idx"][sel], ridx)
    assert tree.boundary_ties(q, k) == 0  # no exact tie at rank k in this cloud: the index sets are well defined
    xyz = np.stack([ref["x"], ref["y"], ref["z"]], 1)
    sub = rng.permutation(len(sel))[:20_000]
    _check_props(got["props"], sel[sub], _fit_rows(orc, xyz, ridx[sub]))


def test_k32_on_the_25m_shard_sampled(ctx, orc, city25):
    """configs[3]'s k = 32 at shard size (the wider lists and staging of the dense kernel, k > 24)."""
    import pointcloudprocess_b200 as pcp
    ref, tree, m = city25["ref"], city25["tree"], city25["m"]
    k = 32
    ix = pcp.Index(ctx, ref)
    kth = np.zeros(m, dtype=np.float32)
    props = ix.features_knn(k, want_eig=False, want_lps=False, kth_d2=kth)["props"]
    rng = np.random.default_rng(11)
    sel = np.unique(np.concatenate([rng.integers(0, m, 40_000), np.argsort(kth)[-10_000:]]))
    q = np.stack([ref["x"][sel], ref["y"][sel], ref["z"][sel]], 1)
    idx, d2, found = ix.knn(q, k)  # the batched external-query path on the same index
    ix.free()
    ridx, rd2, rfound = tree.knn(q, k)
    assert np.array_equal(idx, ridx) and np.array_equal(d2, rd2) and np.array_equal(found, rfound)
    xyz = np.stack([ref["x"], ref["y"], ref["z"]], 1)
    sub = rng.permutation(len(sel))[:10_000]
    _check_props(props, sel[sub], _fit_rows(orc, xyz, ridx[sub]))


def test_config1_voxel_grid_10m_strip_in_full(ctx, orc):
    """configs[1]: VoxelGrid leaf 0.05 on the 10 M-point strip -- 7e9 cells, so the reference's 32-bit cell index would
    overflow (the oracle says so) and the 64-bit-key path runs."""
    n = 10_000_000
    cloud = ctx.synth_cloud("strip", n, seed=2)
    ref = orc.voxel_grid(cloud, 0.05)
    got = ctx.voxel_grid(cloud, 0.05, want_cells=True)
    assert ref["overflow32"]
    assert got["m"] == len(ref["points"])
    assert np.array_equal(got["min_b"], ref["min_b"]) and np.array_equal(got["div_b"], ref["div_b"])
    assert np.array_equal(got["point_cell"], ref["point_cell"])
    assert np.array_equal(got["cell_count"], ref["cell_count"])
    _fields_equal(got["points"], ref["points"], ("x", "y", "z", "w", "rgba", "stamp_id"))


def test_config2_radius_features_20m_forest_sampled(ctx, orc):
    """configs[2]: radius (r = 0.5) neighbourhoods on the 20 M-point forest: neighbour counts and records of 20 k
    queries against the oracle's kd-tree over the whole plot."""
    import pointcloudprocess_b200 as pcp
    n, r = 20_000_000, 0.5
    cloud = ctx.synth_cloud("forest", n, seed=3)
    ix = pcp.Index(ctx, cloud)
    got = ix.features_radius(r)
    ix.free()
    tree = orc.KdTree(cloud)
    rng = np.random.default_rng(13)
    sel = np.unique(np.concatenate([rng.integers(0, n, 15_000), np.argsort(got["counts"])[:2_500], np.argsort(got["counts"])[-2_500:]]))
    q = np.stack([cloud["x"][sel], cloud["y"][sel], cloud["z"][sel]], 1)
    off, idx, d2 = tree.radius(q, r)
    assert np.array_equal(got["counts"][sel], np.diff(off).astype(np.int32))
    # records within the north star's tolerances (the radius kernel sums its moments cell by cell, not in neighbour
    # order: <= 1e-5 relative on eigenvalues, <= 1e-4 rad on normals, as tests/test_gpu_parity.py)
    xyz = np.stack([cloud["x"], cloud["y"], cloud["z"]], 1)
    sub = rng.permutation(len(sel))[:4_000]
    worst_eig = worst_ang = 0.0
    for j in sub:
        rows = idx[off[j]:off[j + 1]]
        if len(rows) <= 3:
            assert got["props"]["curvature"][sel[j]] == 1.0  # the N <= 3 fallback (calculate_feature.cpp:356-363)
            continue
        pl = orc.plane_fit(xyz[rows])
        e_ref, e_got = np.asarray(pl["eig"]), got["eig"][sel[j]]
        worst_eig = max(worst_eig, float(np.abs(e_got - e_ref).max() / max(abs(e_ref[0]), 1e-300)))
        if (e_ref[1] - e_ref[2]) / e_ref[0] > 1e-6:  # the normal of an isotropic neighbourhood is undefined
            n1 = np.array([got["props"][f][sel[j]] for f in ("normal_x", "normal_y", "normal_z")], dtype=np.float64)
            n2 = np.array([pl[f] for f in ("normal_x", "normal_y", "normal_z")], dtype=np.float64)
            c = abs(float(n1 @ n2)) / max(np.linalg.norm(n1) * np.linalg.norm(n2), 1e-300)
            worst_ang = max(worst_ang, float(np.arccos(min(c, 1.0))))
    assert worst_eig <= 1e-5 and worst_ang <= 1e-4, (worst_eig, worst_ang)
