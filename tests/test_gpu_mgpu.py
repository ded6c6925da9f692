"""pcp_mgpu_* (csrc/mgpu.cu): the pipeline over several GPUs of ONE process behind the C ABI -- host threads, peer
copies, no torch.  With fewer GPUs than shards the contexts share a device (a peer copy on one device is a plain
copy), so the whole choreography -- plan chain, partition and pulls, halo, boundary queries, merge -- runs on one GPU;
on a multi-GPU box the same test uses the real devices.  Results equal the single-device calls point for point."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _single(ctx, cloud, leaf, k):
    import pointcloudprocess_b200 as pcp
    dd = ctx.remove_duplicate(cloud, leaf) if leaf else cloud
    ix = pcp.Index(ctx, dd)
    props = ix.features_knn(k, want_eig=False, want_lps=False)["props"]
    ix.free()
    return dd, props


def _compare(p1, f1, p2, f2, point_fields):
    assert len(p1) == len(p2)
    o1 = np.lexsort((p1["z"], p1["y"], p1["x"]))
    o2 = np.lexsort((p2["z"], p2["y"], p2["x"]))
    for f in point_fields:
        assert p1[f][o1].tobytes() == p2[f][o2].tobytes(), f
    for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature"):
        a, b = f1[f][o1], f2[f][o2]
        assert not (~((a == b) | (np.isnan(a) & np.isnan(b)))).any(), f
    assert (f2["PointID"] == np.arange(len(f2))).all()  # props[i] describes out_pts[i]


@pytest.mark.parametrize("shards,kind,n,leaf,k,halo,packed", [(2, "city", 400_000, 0.02, 20, 1.0, False), (3, "city", 500_000, 0.02, 20, 0.4, True),
                                                               (4, "strip", 400_000, 0.04, 20, 0.5, False), (2, "forest", 200_000, None, 32, 0.5, True)])
def test_mgpu_equals_single_device(ctx, shards, kind, n, leaf, k, halo, packed):
    import pointcloudprocess_b200 as pcp
    ngpu = ctx.L.pcp_device_count()
    devices = [r % ngpu for r in range(shards)]
    cloud = ctx.synth_cloud(kind, n, seed=5)
    p1, f1 = _single(ctx, cloud, leaf, k)
    src = cloud
    if packed:
        src = np.zeros(n, dtype=pcp.PACKED_POINT_DTYPE)
        ctx.cloud_pack(cloud, n, src)
    mg = pcp.MultiGpu(shards, devices)
    try:
        p2, f2 = mg.features_knn(src, leaf, k, halo)
        p3, f3 = mg.features_knn(src, leaf, k, halo)  # a second call on the same handle
    finally:
        mg.close()
    fields = ("x", "y", "z", "rgba", "stamp_id") + (() if packed else ("w",))
    _compare(p1, f1, p2, f2, fields)
    assert p2.tobytes() == p3.tobytes() and f2.tobytes() == f3.tobytes()


def test_mgpu_argument_errors(ctx):
    import ctypes as C
    import pointcloudprocess_b200 as pcp
    with pytest.raises(pcp.PcpError):
        pcp.MultiGpu(2, [0, 999])
    mg = pcp.MultiGpu(1)
    try:
        m = C.c_size_t(7)
        a = np.zeros(10, dtype=np.dtype((np.void, 40)))
        assert mg.L.pcp_mgpu_features_knn(mg.h, a.ctypes.data, 10, 40, C.c_float(0.02), 20, 1.0, None, None, C.byref(m)) == -1
        # a cloud too narrow for its slabs' halos is refused, not answered wrongly
        tiny = ctx.synth_cloud("uniform_cube", 5000, seed=1)
    finally:
        mg.close()
    mg = pcp.MultiGpu(4, [0, 0, 0, 0])
    try:
        with pytest.raises(pcp.PcpError):
            mg.features_knn(tiny, 0.02, 20, halo=5.0)  # 10 m of cloud cannot hold 4 slabs at least 5 m wide
    finally:
        mg.close()
