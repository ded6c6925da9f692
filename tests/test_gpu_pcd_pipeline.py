"""The pipelined PCD loader (pcp_pcd_load_many): payloads decoded by host threads, uploaded as stored and
assembled into records on the device -- against the single-file host reader, byte for byte."""
import os
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def make(pcp, n, seed):
    rng = np.random.default_rng(seed)
    xyz = rng.uniform(0, 50, (n, 3)) + [483189.100569, 3371741.645173, 430.094726]
    return pcp.make_cloud(xyz, rgba=rng.integers(0, 1 << 24, n, dtype=np.uint32), stamp_id=np.arange(n) % 10)


def fields_equal(a, b):
    return all(a[f].tobytes() == b[f].tobytes() for f in ("x", "y", "z", "w", "rgba", "stamp_id"))


def test_many_files_equal_the_host_reader(ctx, tmp_path):
    import pointcloudprocess_b200 as pcp
    files, clouds = [], []
    savers = (pcp.io.savePCDFileBinaryCompressed, pcp.io.savePCDFileBinary, pcp.io.savePCDFileASCII)
    for i, n in enumerate((5000, 1, 2, 20000, 333, 12345, 7)):
        c = pcp.PointCloud()
        c.points = make(pcp, n, 10 + i)
        if i == 3:
            c.points["x"][17] = np.nan  # is_dense must come back false
        f = str(tmp_path / ("c%d.pcd" % i))
        assert savers[i % 3](f, c) == 0
        files.append(f)
        clouds.append(c.points)
    # host output
    for threads in (1, 3, 0):
        got, offsets, dense = pcp.io.loadPCDFiles(files, ctx, threads=threads)
        assert list(np.diff(offsets)) == [len(c) for c in clouds] and not dense
        for i, f in enumerate(files):
            one = pcp.PointCloud()
            assert pcp.io.loadPCDFile(f, one) == 0
            assert fields_equal(got[offsets[i]:offsets[i + 1]], one.points), f
    # device output feeds the rest of the path without another copy
    dev = ctx.alloc(len(got) * 48)
    n, offsets2, dense = pcp.io.loadPCDFiles(files, ctx, out=dev)
    assert n == len(got) and list(offsets2) == list(offsets)
    assert fields_equal(dev.to_host(pcp.POINT_DTYPE, n), got)
    m = ctx.remove_duplicate((dev.ptr, n, 48), 0.05)
    ref = ctx.remove_duplicate(got[np.isfinite(got["x"])], 0.05)
    assert m == len(ref)
    dev.free()
    # dense files report dense
    _, _, dense = pcp.io.loadPCDFiles(files[:2], ctx)
    assert dense


def test_foreign_layouts_padding_and_errors(ctx, tmp_path, capsys):
    import pointcloudprocess_b200 as pcp
    # a float cloud with a padding field and an extra field: x y z do not match the double members (size 4), rgba does
    n = 100
    rng = np.random.default_rng(1)
    rec = np.zeros(n, dtype=[("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("_", "u1", (4,)), ("rgba", "<u4"), ("intensity", "<f4")])
    rec["x"], rec["rgba"], rec["intensity"] = rng.random(n), rng.integers(0, 1 << 24, n), rng.random(n)
    hdr = ("# .PCD v0.7\nVERSION 0.7\nFIELDS x y z _ rgba intensity\nSIZE 4 4 4 1 4 4\nTYPE F F F U U F\nCOUNT 1 1 1 4 1 1\n"
           "WIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA binary\n" % (n, n))
    f = str(tmp_path / "foreign.pcd")
    open(f, "wb").write(hdr.encode() + rec.tobytes())
    got, offsets, dense = pcp.io.loadPCDFiles([f], ctx)
    one = pcp.PointCloud()
    assert pcp.io.loadPCDFile(f, one) == 0
    assert fields_equal(got, one.points) and dense and (got["rgba"] == rec["rgba"]).all() and (got["x"] == 0).all()
    # a missing file and a truncated payload fail with the message printed, like the reader
    assert pcp.io.loadPCDFiles([f, str(tmp_path / "nope.pcd")], ctx) is None
    g = str(tmp_path / "short.pcd")
    open(g, "wb").write(hdr.encode() + rec.tobytes()[:-10])
    assert pcp.io.loadPCDFiles([f, g], ctx) is None
    assert "do not fit into the file" in capsys.readouterr().out  # caught at the header now: the payload is never read
    # capacity
    dev = ctx.alloc(10 * 48)
    assert pcp.io.loadPCDFiles([f], ctx, out=dev) is None
    dev.free()
