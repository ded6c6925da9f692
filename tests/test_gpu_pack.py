"""The packed records that cross PCIe (csrc/pack.cu) and the out-of-place remove_duplicate: everything is lossless, so
every comparison is bit for bit -- with the 48-byte path and, through it, with the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _fields_equal(a, b, fields):
    assert len(a) == len(b)
    for f in fields:
        assert a[f].tobytes() == b[f].tobytes(), f


def test_cloud_pack_unpack_roundtrip(ctx):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("city", 200_000, seed=3)
    packed = np.zeros(len(cloud), dtype=pcp.PACKED_POINT_DTYPE)
    ctx.cloud_pack(cloud, len(cloud), packed)
    _fields_equal(packed, cloud, ("x", "y", "z", "rgba", "stamp_id"))
    back = np.zeros(len(cloud), dtype=pcp.POINT_DTYPE)
    ctx.cloud_unpack(packed, len(cloud), back)
    _fields_equal(back, cloud, ("x", "y", "z", "w", "rgba", "stamp_id"))


@pytest.mark.parametrize("kind,n,leaf", [("city", 300_000, 0.02), ("strip", 200_000, 0.05), ("forest", 100_000, 0.3)])
def test_remove_duplicate_on_packed_records_and_out_of_place(ctx, orc, kind, n, leaf):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud(kind, n, seed=5)
    ref, _ = orc.remove_duplicate(cloud, leaf)
    # in place, 48-byte records (the path the other tests pin to the oracle)
    got48 = ctx.remove_duplicate(cloud, leaf)
    _fields_equal(got48, ref, ("x", "y", "z", "rgba", "stamp_id"))
    # out of place, 48-byte records, host buffers
    out48 = np.zeros(n, dtype=pcp.POINT_DTYPE)
    m = ctx.remove_duplicate_to(cloud, leaf, out48)
    assert m == len(ref)
    _fields_equal(out48[:m], ref, ("x", "y", "z", "w", "rgba", "stamp_id"))
    # packed 32-byte records in and out: in place and out of place
    packed = np.zeros(n, dtype=pcp.PACKED_POINT_DTYPE)
    ctx.cloud_pack(cloud, n, packed)
    got32 = ctx.remove_duplicate(packed, leaf)
    _fields_equal(got32, ref, ("x", "y", "z", "rgba", "stamp_id"))
    out32 = np.zeros(n, dtype=pcp.PACKED_POINT_DTYPE)
    m = ctx.remove_duplicate_to(packed, leaf, out32)
    assert m == len(ref)
    _fields_equal(out32[:m], ref, ("x", "y", "z", "rgba", "stamp_id"))
    # the input of the out-of-place form is left alone
    chk = np.zeros(n, dtype=pcp.PACKED_POINT_DTYPE)
    ctx.cloud_pack(cloud, n, chk)
    assert packed.tobytes() == chk.tobytes()


def test_features_on_packed_cloud_and_packed_props(ctx, orc):
    import pointcloudprocess_b200 as pcp
    cloud = ctx.synth_cloud("strip", 50_000, seed=9)
    ref, _, _, _ = orc.features_knn(cloud, 20, want_eig=False, want_lps=False)
    packed = np.zeros(len(cloud), dtype=pcp.PACKED_POINT_DTYPE)
    ctx.cloud_pack(cloud, len(cloud), packed)
    ix = pcp.Index(ctx, packed)  # stride 32: x, y, z first, as every cloud record
    props = ix.features_knn(20, want_eig=False, want_lps=False)["props"]
    ix.free()
    for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID"):
        assert np.array_equal(props[f], ref[f], equal_nan=True), f
    pk = np.zeros(len(cloud), dtype=pcp.PACKED_PROP_DTYPE)
    ctx.props_pack(props, len(props), pk)
    ctx.synchronize()
    back = ctx.props_unpack(pk, threads=3)
    for f in ("normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID", "SegmentID", "dis_from_point_plane"):
        assert back[f].tobytes() == props[f].tobytes(), f  # lossless: Distance / curvature are widened floats
    big = ctx.props_unpack(np.tile(pk, 3))  # > 65536 records: the threaded path
    assert big["Distance"][len(pk):2 * len(pk)].tobytes() == props["Distance"].tobytes()
