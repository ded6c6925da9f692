"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same inputs.
Bars: kNN index sets, squared distances, voxel cells/counts/order bit-exact; features within the
north-star tolerances (<= 1e-5 relative on eigenvalues, <= 1e-4 rad on normals) -- and, because both sides
run the same non-fused arithmetic in the same order, in fact bit-identical, which is asserted too."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

UTM = np.array([483189.100569, 3371741.645173, 430.094726])


def cloud_uniform(n, seed, offset=True, side=10.0):
    rng = np.random.default_rng(seed)
    xyz = rng.uniform(0, side, size=(n, 3)) + (UTM if offset else 0.0)
    import pointcloudprocess_b200 as pcp
    c = pcp.make_cloud(xyz, rgba=rng.integers(0, 1 << 24, n, dtype=np.uint32), stamp_id=np.arange(n) % 10)
    return c


def assert_clouds_equal(a, b):
    """Field-wise bit equality (numpy does not preserve the padding bytes of a structured copy)."""
    assert len(a) == len(b)
    for f in ("x", "y", "z", "w", "rgba", "stamp_id"):
        assert a[f].tobytes() == b[f].tobytes(), f


def assert_props_equal(got, ref):
    for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID"):
        assert np.array_equal(got[f], ref[f], equal_nan=True), f


def check_features_tolerance(got_props, got_eig, ref_props, ref_eig):
    fitted = ref_eig[:, 0] > 0
    rel = np.abs(got_eig[fitted] - ref_eig[fitted]) / np.maximum(np.abs(ref_eig[fitted][:, :1]), 1e-300)
    assert rel.max(initial=0.0) <= 1e-5  # eigenvalues, relative to lambda1
    n1 = np.stack([got_props[f] for f in ("normal_x", "normal_y", "normal_z")], 1).astype(np.float64)[fitted]
    n2 = np.stack([ref_props[f] for f in ("normal_x", "normal_y", "normal_z")], 1).astype(np.float64)[fitted]
    # sign-canonicalised angle
    cosang = np.abs(np.sum(n1 * n2, 1)) / np.maximum(np.linalg.norm(n1, axis=1) * np.linalg.norm(n2, axis=1), 1e-300)
    # isotropic neighbourhoods (lambda2 ~ lambda3) have no defined normal: exclude near-degenerate ones
    sep = (ref_eig[fitted][:, 1] - ref_eig[fitted][:, 2]) / ref_eig[fitted][:, 0] > 1e-6
    ang = np.arccos(np.clip(cosang[sep], -1, 1))
    assert ang.max(initial=0.0) <= 1e-4


# ---- the reference's own (derivable) known answers: main_test.cpp:126-188 ---------------------------------
def test_kd_tree_known_answer(ctx):
    import pointcloudprocess_b200 as pcp
    cloud = pcp.PointCloud()
    cloud.points = pcp.make_cloud(np.array([[0.1 * i] * 3 for i in range(100)]))
    k = pcp.KdTreeFLANN(ctx)
    k.setInputCloud(cloud)
    n, index, dis2 = k.nearestKSearch(np.array([0.0, 0.0, 0.0]), 10)
    assert n == 10
    assert list(index) == list(range(10))
    exp = []
    for i in range(10):
        d = (0.1 * i) - 0.0
        exp.append((d * d + d * d) + d * d)
    assert list(dis2) == exp


def test_voxel_grid_known_answer(ctx):
    import pointcloudprocess_b200 as pcp
    pts = [[float(i)] * 3 for i in range(10)] + [[0.1 * i] * 3 for i in range(100)]
    cloud = pcp.PointCloud(pcp.make_cloud(np.array(pts)))
    vg = pcp.VoxelGrid(ctx)
    vg.setInputCloud(cloud)
    vg.setLeafSize(1, 1, 1)
    out = vg.filter()
    assert out.size() == 10
    for c in range(10):
        vals = [np.float32(c)] + [np.float32(0.1 * i) for i in range(10 * c, 10 * c + 10)]
        exp = sum(float(v) for v in vals) / 11.0
        assert out.points["x"][c] == exp and out.points["y"][c] == exp and out.points["z"][c] == exp
    assert (out.points["rgba"] == 0).all() and (out.points["stamp_id"] == 0).all()


# ---- kNN ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,k,offset", [(100, 10, False), (1000, 20, True), (10000, 20, True), (20000, 32, True),
                                        (5000, 50, False), (3000, 1, True), (300, 100, True)])
def test_knn_matches_oracle(ctx, orc, n, k, offset):
    import pointcloudprocess_b200 as pcp
    cloud = cloud_uniform(n, 7 + n, offset)
    ix = pcp.Index(ctx, cloud)
    rng = np.random.default_rng(1)
    # members and non-members (some outside the bounding box) as queries
    q = np.concatenate([np.stack([cloud["x"], cloud["y"], cloud["z"]], 1)[:500],
                        rng.uniform(-3, 13, size=(200, 3)) + (UTM if offset else 0.0)])
    idx, d2, found = ix.knn(q, k)
    ridx, rd2, rfound = orc.knn_brute(cloud, q, k)
    assert np.array_equal(found, rfound)
    assert np.array_equal(idx, ridx)
    assert np.array_equal(d2, rd2)
    ix.free()


def test_knn_small_and_degenerate(ctx, orc):
    import pointcloudprocess_b200 as pcp
    for n in (1, 2, 3, 4, 5, 19, 20, 21):
        cloud = cloud_uniform(n, 100 + n, True)
        ix = pcp.Index(ctx, cloud)
        idx, d2, found = ix.knn(cloud, 20)
        ridx, rd2, rfound = orc.knn_brute(cloud, cloud, 20)
        assert np.array_equal(found, rfound) and np.array_equal(idx, ridx) and np.array_equal(d2, rd2)
        got = ix.features_knn(20, want_idx=True)
        props, eig, lps, kidx = orc.features_knn(cloud, 20, want_idx=True)
        assert np.array_equal(got["knn_idx"], kidx)
        assert_props_equal(got["props"], props)
        ix.free()
    # all points identical (zero scatter -> NaN curvature on both sides)
    cloud = pcp.make_cloud(np.tile(UTM, (50, 1)))
    ix = pcp.Index(ctx, cloud)
    got = ix.features_knn(20, want_idx=True)
    props, eig, lps, kidx = orc.features_knn(cloud, 20, want_idx=True)
    assert np.array_equal(got["knn_idx"], kidx)
    assert_props_equal(got["props"], props)
    ix.free()


def test_knn_with_nan_points_remaps_indices(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = cloud_uniform(5000, 9, True)
    bad = np.random.default_rng(3).choice(5000, 50, replace=False)
    cloud["x"][bad[:20]] = np.nan
    cloud["y"][bad[20:35]] = np.inf
    cloud["z"][bad[35:]] = -np.inf
    ix = pcp.Index(ctx, cloud)
    assert ix.size() == 4950
    got = ix.features_knn(20, want_idx=True)
    props, eig, lps, kidx = orc.features_knn(cloud, 20, want_idx=True)
    assert np.array_equal(got["knn_idx"], kidx)
    assert_props_equal(got["props"], props)
    assert np.array_equal(got["eig"], eig)
    ix.free()


def test_knn_lattice_ties_follow_index_order(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("lattice", 4096, seed=8, centred=True)
    ix = pcp.Index(ctx, cloud)
    got = ix.features_knn(20, want_idx=True)
    props, eig, lps, kidx = orc.features_knn(cloud, 20, want_idx=True)
    assert orc.KdTree(cloud).boundary_ties(cloud, 20) > 0  # the case really has rank-k ties
    assert np.array_equal(got["knn_idx"], kidx)
    assert_props_equal(got["props"], props)
    idx, d2, found = ix.knn(cloud, 20)
    ridx, rd2, _ = orc.knn_brute(cloud, cloud, 20)
    assert np.array_equal(idx, ridx) and np.array_equal(d2, rd2)
    ix.free()


# ---- features ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,n,k", [("strip", 50000, 20), ("forest", 50000, 20), ("city", 100000, 20),
                                      ("city", 60000, 32), ("uniform_cube", 30000, 10), ("strip", 30000, 50)])
def test_features_knn_match_oracle(ctx, orc, kind, n, k):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud(kind, n, seed=21)
    ix = pcp.Index(ctx, cloud)
    got = ix.features_knn(k, want_idx=True)
    props, eig, lps, kidx = orc.features_knn(cloud, k, want_idx=True)
    assert np.array_equal(got["knn_idx"], kidx)
    check_features_tolerance(got["props"], got["eig"], props, eig)
    assert_props_equal(got["props"], props)            # bit-identical, beyond the required tolerance
    assert np.array_equal(got["eig"], eig, equal_nan=True)
    assert np.array_equal(got["lps"], lps, equal_nan=True)
    ix.free()


def test_plane_fit_segments(ctx, orc):
    rng = np.random.default_rng(5)
    sizes = [3, 4, 13, 20, 50, 1000]
    pts = np.concatenate([rng.normal(size=(s, 3)) * [5, 2, 0.01] + UTM for s in sizes])
    off = np.concatenate([[0], np.cumsum(sizes)])
    got = ctx.plane_fit(pts, off)
    for s in range(len(sizes)):
        ref = orc.plane_fit(pts[off[s]:off[s + 1]])
        assert got[s].tobytes() == ref.tobytes()


# ---- radius ----------------------------------------------------------------------------------------------
def test_radius_search_matches_oracle(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("forest", 40000, seed=3)
    ix = pcp.Index(ctx, cloud)
    t = orc.KdTree(cloud)
    q = cloud[:2000]
    for r, max_nn in ((0.5, 0), (0.25, 0), (1.0, 7), (0.0, 0)):
        off, idx, d2 = ix.radius(q, r, max_nn)
        roff, ridx, rd2 = t.radius(q, r, max_nn)
        assert np.array_equal(off, roff)
        assert np.array_equal(idx, ridx)
        assert np.array_equal(d2, rd2)
    ix.free()


def test_features_radius_match_oracle(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("forest", 40000, seed=4)
    ix = pcp.Index(ctx, cloud)
    got = ix.features_radius(0.5)
    props, eig, lps, counts = orc.features_radius(cloud, 0.5)
    assert np.array_equal(got["counts"], counts)          # neighbour sets are exact
    check_features_tolerance(got["props"], got["eig"], props, eig)
    ix.free()


@pytest.mark.parametrize("kind,n,r", [("city", 60000, 0.3), ("uniform_cube", 20000, 1.0), ("strip", 50000, 2.5),
                                      ("forest", 30000, 0.05)])
def test_features_radius_other_radii(ctx, orc, kind, n, r):
    """the cell-cooperative kernel (block radius 1 and 2), and the warp-per-query kernel when the ball outgrows it"""
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud(kind, n, seed=6)
    ix = pcp.Index(ctx, cloud)
    got = ix.features_radius(r)
    props, eig, lps, counts = orc.features_radius(cloud, r)
    assert np.array_equal(got["counts"], counts)
    check_features_tolerance(got["props"], got["eig"], props, eig)
    ix.free()


def test_features_radius_exact_boundary_on_a_lattice(ctx, orc):
    """distances that equal r exactly (strict d2 < r2, kd_tree.h:888) and pairs inside the FP32 decision band"""
    import pointcloudprocess_b200 as pcp
    g = np.arange(24) * 0.125  # exactly representable spacing: d2 == r2 happens exactly
    X, Y, Z = np.meshgrid(g, g, g, indexing="ij")
    P = np.c_[X.ravel(), Y.ravel(), Z.ravel()] + np.array([483189.0, 3371741.5, 430.0])
    for r in (0.25, 0.375, 0.2500001, np.nextafter(0.25, 0)):
        ix = pcp.Index(ctx, P)
        got = ix.features_radius(r)
        counts = orc.features_radius(P, r)[3]
        assert np.array_equal(got["counts"], counts), r
        ix.free()


# ---- voxel grid -------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind,n,leaf,centred", [("strip", 100000, 0.05, True), ("strip", 100000, 0.04, True),
                                                 ("city", 200000, 0.02, True), ("forest", 50000, 1.0, True),
                                                 ("uniform_cube", 20000, 0.5, False)])
def test_voxel_grid_matches_oracle(ctx, orc, kind, n, leaf, centred):
    cloud = ctx.synth_cloud(kind, n, seed=2, centred=centred)
    if centred:  # straddle zero like remove_duplicate's centred clouds, incl. negative coordinates
        for f in "xyz":
            cloud[f] -= np.median(cloud[f])
    got = ctx.voxel_grid(cloud, leaf, want_cells=True)
    ref = orc.voxel_grid(cloud, leaf)
    assert got["m"] == len(ref["points"])
    assert np.array_equal(got["min_b"], ref["min_b"]) and np.array_equal(got["div_b"], ref["div_b"])
    assert np.array_equal(got["point_cell"], ref["point_cell"])
    assert np.array_equal(got["cell_count"], ref["cell_count"])
    assert_clouds_equal(got["points"], ref["points"])  # centroids, rgb repack, stamp_id, order


def test_voxel_grid_xyz_only_and_nan(ctx, orc):
    cloud = ctx.synth_cloud("strip", 30000, seed=6, centred=True)
    cloud["x"][::97] = np.nan
    got = ctx.voxel_grid(cloud, 0.1, downsample_all=False, want_cells=True)
    ref = orc.voxel_grid(cloud, 0.1, downsample_all=False, is_dense=False)
    assert got["m"] == len(ref["points"])
    assert np.array_equal(got["point_cell"], ref["point_cell"])
    assert np.array_equal(got["points"]["x"], ref["points"]["x"])
    assert np.array_equal(got["points"]["y"], ref["points"]["y"])
    assert np.array_equal(got["points"]["z"], ref["points"]["z"])
    # all-negative axis: the DBL_MIN max seed of getMinMax3D (point_cloud_helper.h:65)
    cloud = ctx.synth_cloud("uniform_cube", 5000, seed=6, centred=True)
    cloud["y"] -= 100.0
    got = ctx.voxel_grid(cloud, 0.25, want_cells=True)
    ref = orc.voxel_grid(cloud, 0.25)
    assert np.array_equal(got["div_b"], ref["div_b"])
    assert_clouds_equal(got["points"], ref["points"])
    mn, mx = ctx.minmax(cloud)
    rmn, rmx = orc.minmax(cloud)
    assert np.array_equal(mn, rmn) and np.array_equal(mx, rmx)


def test_empty_inputs(ctx):
    import pointcloudprocess_b200 as pcp
    empty = np.zeros(0, dtype=pcp.POINT_DTYPE)
    assert ctx.voxel_grid(empty, 0.1)["m"] == 0
    ix = pcp.Index(ctx, empty)
    assert ix.size() == 0
    idx, d2, found = ix.knn(np.zeros((3, 3)), 5)
    assert (idx == -1).all() and (found == 0).all()
    assert len(ix.features_knn(20)["props"]) == 0
    ix.free()


@pytest.mark.parametrize("kind,n,leaf", [("strip", 200000, 0.04), ("city", 100000, 0.02), ("forest", 50000, 0.05)])
def test_remove_duplicate_matches_oracle(ctx, orc, kind, n, leaf):
    cloud = ctx.synth_cloud(kind, n, seed=12)  # UTM-scale coordinates: centring matters
    got = ctx.remove_duplicate(cloud, leaf)
    ref, centre = orc.remove_duplicate(cloud, leaf)
    c, cnt = ctx.centroid(cloud)
    assert np.array_equal(c, centre)  # serial-order sum reproduced bit-exactly
    assert_clouds_equal(got, ref)


# ---- serial-order centroid (compute3DCentroid, point_cloud_helper.h:192-230) -----------------------------------
def _centroid_cases():
    rng = np.random.default_rng(77)
    n = 300_000
    cases = {}
    cases["utm"] = rng.uniform(0, 50, size=(n, 3)) + UTM
    cases["centred_mixed_sign"] = rng.normal(size=(n, 3)) * [30.0, 5.0, 0.5]
    cases["negative"] = -(rng.uniform(0, 50, size=(n, 3)) + UTM)
    # round-half-even ties: once the running sum passes 2^62 its ulp is 2^10 and every odd multiple of 2^9 ties
    t = 2.0 ** 45 + 2.0 ** 9 * rng.integers(0, 64, size=(n, 3))
    cases["ties"] = t
    cases["ties_negative"] = -t
    sw = t.copy()
    sw[n // 2:] *= -1.0  # climbs through the binades, then back down through them
    cases["ties_up_then_down"] = sw
    cases["huge_dynamic_range"] = rng.normal(size=(n, 3)) * 10.0 ** rng.integers(-12, 12, size=(n, 3))
    cases["tiny_then_big"] = np.concatenate([rng.uniform(0, 1e-9, size=(n // 2, 3)), rng.uniform(0, 1e9, size=(n // 2, 3))])
    z = rng.uniform(-1, 1, size=(n, 3))
    z[::3] = 0.0
    cases["zeros_and_cancellation"] = z
    return cases


@pytest.mark.parametrize("name", list(_centroid_cases().keys()))
def test_serial_centroid_bit_exact(ctx, orc, name):
    xyz = _centroid_cases()[name]
    c, cnt = ctx.centroid(xyz)
    rc, rcnt = orc.centroid(xyz, is_dense=False)
    assert cnt == rcnt
    assert c.tobytes() == rc.tobytes(), (name, c, rc)


def test_serial_centroid_nan_and_sizes(ctx, orc):
    rng = np.random.default_rng(78)
    for n in (1, 31, 255, 257, 16383, 16384, 16385, 70000, 262144 + 5):
        xyz = rng.uniform(0, 50, size=(n, 3)) + UTM
        if n > 100:
            xyz[rng.choice(n, n // 50, replace=False), rng.integers(0, 3, n // 50)] = np.nan
        c, cnt = ctx.centroid(xyz)
        rc, rcnt = orc.centroid(xyz, is_dense=False)
        assert cnt == rcnt and c.tobytes() == rc.tobytes(), n


def test_serial_centroid_full_size(ctx, orc):
    cloud = ctx.synth_cloud("city", 5_000_000, seed=5)
    c, cnt = ctx.centroid(cloud)
    rc, rcnt = orc.centroid(cloud)
    assert cnt == rcnt and c.tobytes() == rc.tobytes()


# ---- BASELINE.json configs[0]: calculate_feature k=20 on a cloud that went through a binary_compressed PCD -------
@pytest.mark.parametrize("n", [100000, 1000000])
def test_config0_features_on_a_cloud_loaded_from_compressed_pcd(ctx, orc, tmp_path, n):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("strip", n, seed=1)
    path = tmp_path / "synth.pcd"
    assert pcp.io.savePCDFileBinaryCompressed(path, cloud) == 0
    loaded = pcp.PointCloud()
    assert pcp.io.loadPCDFile(path, loaded) == 0
    assert_clouds_equal(loaded.points, cloud)  # x y z rgba stamp_id survive bit for bit; w is the constructor's 1
    ix = pcp.Index(ctx, loaded.points)
    got = ix.features_knn(20, want_idx=(n <= 100000))
    ix.free()
    if n <= 100000:
        props, eig, lps, kidx = orc.features_knn(loaded.points, 20, want_idx=True)
        assert np.array_equal(got["knn_idx"], kidx)
    else:
        props, eig, lps = orc.features_knn(loaded.points, 20)[:3]
    assert_props_equal(got["props"], props)
    assert np.array_equal(got["eig"], eig, equal_nan=True)


def test_voxel_grid_accessors_and_leaf_layout(ctx):
    """voxel_grid.h:595-706 on main_test.cpp:126-154's cloud: 10 cells along the diagonal of a 10^3 grid."""
    import pointcloudprocess_b200 as pcp
    cloud = pcp.PointCloud()
    for i in range(10):
        cloud.push_back(i, i, i)
    for i in range(100):
        cloud.push_back(0.1 * i, 0.1 * i, 0.1 * i)
    vg = pcp.VoxelGrid(ctx)
    vg.setInputCloud(cloud)
    vg.setLeafSize(1, 1, 1)
    vg.setSaveLeafLayout(True)
    out = vg.filter()
    assert out.size() == 10
    assert list(vg.getMinBoxCoordinates()) == [0, 0, 0] and list(vg.getMaxBoxCoordinates()) == [9, 9, 9]
    assert list(vg.getNrDivisions()) == [10, 10, 10] and list(vg.getDivisionMultiplier()) == [1, 10, 100]
    layout = vg.getLeafLayout()
    assert layout.shape == (1000,) and (layout >= 0).sum() == 10
    for c in range(10):
        assert layout[111 * c] == c and vg.getCentroidIndexAt([c, c, c]) == c
        assert vg.getCentroidIndex(out.points[c]) == c
    assert vg.getCentroidIndexAt([0, 1, 0]) == -1 and vg.getCentroidIndexAt([10, 0, 0]) == -1
    assert list(vg.getGridCoordinates(3.7, -0.2, 9.99)) == [3, 0, 9]
    rel = np.array([[0, 1, -1, 1], [0, 1, -1, 0], [0, 1, -1, 0]])
    assert vg.getNeighborCentroidIndices(out.points[0], rel) == [0, 1, -1, -1]
    vg.setSaveLeafLayout(False)
    vg.filter()
    assert vg.getLeafLayout().size == 0 and vg.getCentroidIndexAt([0, 0, 0]) == -1


# ---- calculate_plan_parameter_rpca (calculate_feature.cpp:208-368) ------------------------------------------
@pytest.mark.parametrize("kind,n,k,seed", [("city", 100000, 20, 1), ("strip", 50000, 20, 7), ("forest", 50000, 20, 3),
                                           ("uniform_cube", 20000, 12, 1), ("city", 40000, 32, 2**63 + 5)])
def test_features_rpca_match_oracle_bit_exact(ctx, orc, kind, n, k, seed):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud(kind, n, seed=33)
    ix = pcp.Index(ctx, cloud)
    got = ix.features_rpca(k, 0.99, 0.5, seed)
    ref = orc.features_rpca(cloud, k, 0.99, 0.5, seed)
    assert_props_equal(got, ref)
    assert got.tobytes() == ref.tobytes()
    # a different seed gives different hypotheses; the same seed the same bytes
    assert ix.features_rpca(k, 0.99, 0.5, seed).tobytes() == got.tobytes()
    assert ix.features_rpca(k, 0.99, 0.5, seed + 1 if seed < 2**63 else 1).tobytes() != got.tobytes()
    ix.free()


def test_features_rpca_other_probabilities_nan_points_and_tiny_clouds(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("city", 30000, seed=4)
    cloud["x"][::97] = np.nan  # dropped from the index: fallback record, neighbours renumbered like the reference
    ix = pcp.Index(ctx, cloud)
    for pr, epi in ((0.9, 0.5), (0.999, 0.3), (0.99, 0.0)):
        assert pcp.CalculateFeature.compute_iteration_number(pr, epi, ctx) == orc.rpca_iterations(pr, epi)
    got = ix.features_rpca(20, 0.9, 0.5, 9)
    ref = orc.features_rpca(cloud, 20, 0.9, 0.5, 9)
    assert got.tobytes() == ref.tobytes()
    assert (got["curvature"][::97] == 1.0).all()
    ix.free()
    for n in (1, 3, 4, 5, 19):
        P = cloud_uniform(n, 8)
        ix = pcp.Index(ctx, P)
        assert ix.features_rpca(20).tobytes() == orc.features_rpca(P, 20).tobytes(), n
        ix.free()
    with pytest.raises(pcp.PcpError):
        pcp.Index(ctx, cloud_uniform(100, 1)).features_rpca(33)


def test_calculate_plan_parameter_rpca_mirror(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = pcp.PointCloud(ctx.synth_cloud("city", 20000, seed=6))
    props = pcp.CalculateFeature(ctx).calculate_plan_parameter_rpca(cloud, 0.3, 0.99, 0.5, seed=4)
    assert props.tobytes() == orc.features_rpca(cloud.points, 20, 0.99, 0.5, 4).tobytes()


# ---- alternative code paths must agree bit for bit -----------------------------------------------------------
def test_fused_and_single_pass_variants_equal_the_default_path(ctx):
    """PCP_FK_FUSED (plane fit inside the search kernels), PCP_FK_ONEPASS (no second dense pass), PCP_FK_NOFLAT (the
    block staged cell by cell), PCP_FK_NOWIDE (no third pass over the box of the bounded queries), PCP_FK_NORERANK (lists
    the FP32 search cannot order go to the far kernel instead of knn_rerank_kernel) and the default (deferred fit, three
    passes, flat staging, re-rank) are the same function; PCP_RADIUS_WARP is the warp-per-query radius kernel."""
    import os
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("city", 300000, seed=12)
    ix = pcp.Index(ctx, cloud)
    ref = ix.features_knn(20, want_idx=True)
    ref32 = ix.features_knn(32, want_idx=True)
    rad = ix.features_radius(0.3)
    try:
        for env in ("PCP_FK_FUSED", "PCP_FK_ONEPASS", "PCP_FK_NOFLAT", "PCP_FK_NOWIDE", "PCP_FK_NORERANK"):
            os.environ[env] = "1"
            got = ix.features_knn(20, want_idx=True)
            got32 = ix.features_knn(32, want_idx=True)
            del os.environ[env]
            for a, b in ((got, ref), (got32, ref32)):
                assert np.array_equal(a["knn_idx"], b["knn_idx"]), env
                assert a["props"].tobytes() == b["props"].tobytes(), env
                assert a["eig"].tobytes() == b["eig"].tobytes() and a["lps"].tobytes() == b["lps"].tobytes(), env
        os.environ["PCP_RADIUS_WARP"] = "1"
        radw = ix.features_radius(0.3)
        del os.environ["PCP_RADIUS_WARP"]
        assert np.array_equal(radw["counts"], rad["counts"])
        check_features_tolerance(radw["props"], radw["eig"], rad["props"], rad["eig"])
    finally:
        for env in ("PCP_FK_FUSED", "PCP_FK_ONEPASS", "PCP_FK_NOFLAT", "PCP_FK_NOWIDE", "PCP_FK_NORERANK", "PCP_RADIUS_WARP"):
            os.environ.pop(env, None)
        ix.free()


def test_block_per_query_streaming_equals_warp_per_query(ctx, orc):
    """A handful of streaming queries take knn_block_kernel (8 warps share a query, lists merged by warp 0); PCP_KNN_NOBLOCK
    keeps them on knn_warp_kernel.  Same lists and distances, and both equal the oracle's: queries inside, far outside
    and metres above the cloud, k = 20, 50, 100."""
    import os
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("city", 400000, seed=17, centred=True)
    rng = np.random.default_rng(3)
    xyz = np.stack([cloud["x"], cloud["y"], cloud["z"]], 1)
    q = np.concatenate([xyz[rng.integers(0, len(cloud), 24)] + rng.normal(0, 0.3, (24, 3)),
                        xyz[rng.integers(0, len(cloud), 8)] + np.array([0.0, 0.0, 25.0]),
                        np.array([[1e4, -1e4, 50.0], [-500.0, 3.0, 2.0]])])
    ix = pcp.Index(ctx, cloud)
    try:
        for k in (20, 50, 100):
            a = ix.knn(q, k)                       # < 64 queries: the streaming path, block per query
            os.environ["PCP_KNN_NOBLOCK"] = "1"
            b = ix.knn(q, k)
            del os.environ["PCP_KNN_NOBLOCK"]
            ridx, rd2, rfound = orc.KdTree(cloud).knn(q, k)
            for got in (a, b):
                assert np.array_equal(got[0], ridx) and np.array_equal(got[1], rd2) and np.array_equal(got[2], rfound)
    finally:
        os.environ.pop("PCP_KNN_NOBLOCK", None)
        ix.free()


def test_rerank_and_third_pass_on_ties_and_sparse_leaves(ctx, orc):
    """The two kernels behind the dense passes against the oracle where they have most to do: a jittered lattice (FP32
    cannot order near-equal distances: knn_rerank_kernel), an exact lattice (ties ranked by index), and a forest plot
    (sparse leaves beside dense ones: knn_wide_kernel) -- neighbour lists and records, k = 20 and k = 32."""
    import pointcloudprocess_b200 as pcp
    rng = np.random.default_rng(21)
    lat = ctx.synth_cloud("lattice", 32768, seed=8, centred=True)
    jit = lat.copy()
    for f in ("x", "y", "z"):
        jit[f] = (jit[f].astype(np.float64) + rng.uniform(-1e-6, 1e-6, len(jit))).astype(jit[f].dtype)
    for cloud in (lat, jit, ctx.synth_cloud("forest", 200000, seed=14)):
        for k in (20, 32):
            ix = pcp.Index(ctx, cloud)
            got = ix.features_knn(k, want_idx=True)
            ix.free()
            props, eig, lps, kidx = orc.features_knn(cloud, k, want_idx=True)
            assert np.array_equal(got["knn_idx"], kidx)
            assert_props_equal(got["props"], props)


def test_knn_batched_external_queries_edge_cases(ctx, orc):
    """pcp_knn with >= 64 queries runs the far kernel's pruned search with the streaming kernel behind it: queries far
    outside the cloud, non-finite queries, fewer points than k, mass ties (duplicates), and both paths agree."""
    import os
    import subprocess
    import sys
    import pointcloudprocess_b200 as pcp
    rng = np.random.default_rng(5)
    # (a) far outside + NaN queries
    cloud = cloud_uniform(20000, 3, True)
    q = np.concatenate([rng.uniform(-200, 200, size=(300, 3)) + UTM, rng.uniform(0, 10, size=(300, 3)) + UTM])
    q[7] = np.nan
    q[500, 1] = np.inf
    ix = pcp.Index(ctx, cloud)
    idx, d2, found = ix.knn(q, 20)
    ok = np.isfinite(q).all(axis=1)
    ridx, rd2, rfound = orc.knn_brute(cloud, q[ok], 20)
    assert np.array_equal(idx[ok], ridx) and np.array_equal(d2[ok], rd2) and np.array_equal(found[ok], rfound)
    assert (idx[~ok] == -1).all() and (found[~ok] == 0).all() and np.isinf(d2[~ok]).all()
    ix.free()
    # (b) fewer points than k
    small = cloud_uniform(11, 4, True)
    ix = pcp.Index(ctx, small)
    q = rng.uniform(-5, 15, size=(128, 3)) + UTM
    idx, d2, found = ix.knn(q, 20)
    ridx, rd2, rfound = orc.knn_brute(small, q, 20)
    assert np.array_equal(idx, ridx) and np.array_equal(d2, rd2) and np.array_equal(found, rfound)
    ix.free()
    # (c) mass ties: every point 40 times
    base = rng.uniform(0, 3, size=(200, 3)) + UTM
    dup = pcp.make_cloud(np.repeat(base, 40, axis=0))
    ix = pcp.Index(ctx, dup)
    q = np.concatenate([base[:100], rng.uniform(0, 3, size=(100, 3)) + UTM])
    idx, d2, found = ix.knn(q, 30)
    ridx, rd2, rfound = orc.knn_brute(dup, q, 30)
    assert np.array_equal(idx, ridx) and np.array_equal(d2, rd2)
    ix.free()
    # (d) the streaming kernel alone (PCP_KNN_STREAM=1, read once per process) gives the same bytes
    code = ("import numpy as np, pointcloudprocess_b200 as pcp, hashlib\n"
            "ctx = pcp.Context(0)\n"
            "c = ctx.synth_cloud('city', 300000, seed=2)\n"
            "rng = np.random.default_rng(1)\n"
            "q = np.stack([c['x'], c['y'], c['z']], 1)[rng.integers(0, len(c), 5000)] + rng.normal(0, 0.5, (5000, 3))\n"
            "ix = pcp.Index(ctx, c)\n"
            "i, d, f = ix.knn(q, 20)\n"
            "print(hashlib.sha1(i.tobytes() + d.tobytes() + f.tobytes()).hexdigest())\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for env in ({}, {"PCP_KNN_STREAM": "1"}):
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root, env=dict(os.environ, **env), timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout.strip().splitlines()[-1])
    assert outs[0] == outs[1]
