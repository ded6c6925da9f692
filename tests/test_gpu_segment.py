"""TreeExtration::region_growning (extraction_tree.cpp:170-272) on the GPU against the CPU restatement: labels,
membership, member order and segment planes must be identical (integer / index work: bit-exact)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

UTM = np.array([483189.100569, 3371741.645173, 430.094726])


def _equal(got, ref, props_gpu):
    assert np.array_equal(props_gpu["SegmentID"], ref["props"]["SegmentID"])
    assert np.array_equal(got["offsets"], ref["offsets"])
    assert np.array_equal(got["points"], ref["points"])  # PlanSegment::PointID order = the reference's FIFO order
    assert got["planes"].tobytes() == ref["planes"].tobytes()


def _run(ctx, orc, cloud, props, **kw):
    import pointcloudprocess_b200 as pcp
    ix = pcp.Index(ctx, cloud)
    pg = props.copy()
    got = ix.region_growing(pg, want_stats=True, **kw)
    ix.free()
    ref = orc.region_growing(cloud, props, **kw)
    _equal(got, ref, pg)
    return got, ref


@pytest.mark.parametrize("kind,n,seed", [("city", 150000, 4), ("strip", 80000, 2), ("forest", 60000, 3)])
def test_region_growing_matches_oracle_on_synthetic_clouds(ctx, orc, kind, n, seed):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud(kind, n, seed=seed)
    ix = pcp.Index(ctx, cloud)
    props = ix.features_rpca(20, 0.99, 0.5, 7)  # what point_segment feeds it (static.cpp:17-20)
    ix.free()
    got, ref = _run(ctx, orc, cloud, props)
    assert len(ref["planes"]) > 0 and got["stats"]["seeds"] >= len(ref["planes"])


def test_two_planes_noise_and_all_block_passes(ctx, orc):
    rng = np.random.default_rng(0)
    n = 120000
    P = np.c_[rng.uniform(0, 20, (n, 2)), rng.normal(0, 0.0005, n)]
    P[n // 2:, 2] = P[n // 2:, 0] + 30.0          # a second plane at 45 degrees
    P[::1000, 2] += rng.uniform(1, 3, len(P[::1000]))  # isolated outliers never join
    P += UTM
    props = orc.features_knn(P, 20, want_eig=False, want_lps=False)[0]
    got, ref = _run(ctx, orc, P, props)
    assert len(ref["planes"]) >= 2
    assert got["stats"]["wide_passes"] > 0 and got["stats"]["solo_passes"] > 0
    # a single block must find the same thing
    os.environ["PCP_RG_BLOCKS"] = "1"
    try:
        got1, _ = _run(ctx, orc, P, props)
    finally:
        del os.environ["PCP_RG_BLOCKS"]
    assert got1["stats"]["wide_passes"] == 0 and np.array_equal(got1["points"], got["points"])
    os.environ["PCP_RG_BLOCKS"] = "3"
    try:
        _run(ctx, orc, P, props)
    finally:
        del os.environ["PCP_RG_BLOCKS"]


def test_small_regions_are_dissolved_and_absorbed_later(ctx, orc):
    # clusters of 1..9 coplanar points far apart: only those with more than min_points members survive; then the same
    # with other thresholds
    rng = np.random.default_rng(5)
    pts = []
    for c in range(300):
        m = 1 + c % 9
        base = np.array([c * 10.0, 0.0, 0.0])
        pts.append(base + np.c_[rng.uniform(0, 0.05, (m, 2)), np.zeros(m)])
    P = np.concatenate(pts)
    props = np.zeros(len(P), dtype=orc.PROP_DTYPE)
    props["normal_z"], props["PointID"], props["curvature"] = 1.0, np.arange(len(P)), 0.001
    for mp in (5, 0, 3, 8, 100):
        got, ref = _run(ctx, orc, P, props, k=10, min_points=mp)
    # curvature threshold: nothing seeds
    p2 = props.copy()
    p2["curvature"] = 0.005
    got, ref = _run(ctx, orc, P, p2, k=10)
    assert len(ref["planes"]) == 0 and (ref["props"]["SegmentID"] == -1).all()


def test_order_dependence_is_reproduced(ctx, orc):
    # a lattice (exact distance ties, broken by index) with normals that drift: which region takes a point depends
    # on the seed order and on the FIRST seed's normal
    g = np.arange(60)
    X, Y = np.meshgrid(g * 0.1, g * 0.1, indexing="ij")
    P = np.c_[X.ravel(), Y.ravel(), 0.02 * np.sin(X.ravel())]
    rng = np.random.default_rng(9)
    perm = rng.permutation(len(P))
    P = P[perm]
    props = np.zeros(len(P), dtype=orc.PROP_DTYPE)
    ang = 0.25 * P[:, 0]
    props["normal_x"], props["normal_z"] = np.sin(ang), np.cos(ang)
    props["Distance"] = (-(props["normal_x"] * P[:, 0] + props["normal_z"] * P[:, 2])).astype(np.float32)
    props["PointID"] = np.arange(len(P))
    props["curvature"] = np.where(rng.random(len(P)) < 0.3, 0.001, 0.01)
    got, ref = _run(ctx, orc, P, props, k=12, distanceT=0.05, cosfaT=0.97)
    assert len(ref["planes"]) > 3


def test_nan_points_empty_cloud_and_errors(ctx, orc):
    import pointcloudprocess_b200 as pcp
    rng = np.random.default_rng(2)
    n = 5000
    P = np.c_[rng.uniform(0, 5, (n, 2)), rng.normal(0, 0.001, n)]
    P[::50] = np.nan
    props = orc.features_knn(P, 20, want_eig=False, want_lps=False)[0]
    _run(ctx, orc, P, props)
    ix = pcp.Index(ctx, np.zeros(0, dtype=pcp.POINT_DTYPE))
    r = ix.region_growing(np.zeros(0, dtype=pcp.PROP_DTYPE))
    assert list(r["offsets"]) == [0] and len(r["points"]) == 0
    ix.free()
    bad = props.copy()
    bad["PointID"][7] = n
    ix = pcp.Index(ctx, P)
    with pytest.raises(pcp.PcpError):
        ix.region_growing(bad)
    with pytest.raises(pcp.PcpError):
        ix.region_growing(props.copy(), k=129)
    ix.free()


def test_tree_extration_mirror_and_point_segment(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = pcp.PointCloud()
    cloud.points = ctx.synth_cloud("city", 60000, seed=8)
    segs, props = pcp.point_segment(cloud, ctx, seed=3)
    ref_props = orc.features_rpca(cloud.points, 20, 0.99, 0.5, 3)
    ref = orc.region_growing(cloud.points, ref_props, 50, 0.5, 0.940)
    assert len(segs) == len(ref["planes"]) and np.array_equal(props["SegmentID"], ref["props"]["SegmentID"])
    for s, sg in enumerate(segs):
        assert sg["SegmentID"] == s
        assert np.array_equal(sg["PointID"], ref["points"][ref["offsets"][s]:ref["offsets"][s + 1]])
        assert (sg["normal_x"], sg["normal_y"], sg["normal_z"], sg["Distance"]) == tuple(ref["planes"][s])
    # tree_filter keeps the segments with more than 50 points (static.cpp:24-36)
    kept = sum(len(sg["PointID"]) for sg in segs if len(sg["PointID"]) > 50)
    assert 0 < kept <= len(cloud.points)
