"""Host-side PCD v0.7 / LZF code of the library (csrc/pcd_io.cu) -- no GPU needed.
LZF is pinned byte for byte to the reference's own lzf.cpp (golden streams, and the compiled oracle/_ref when
present); the PCD layer to the reference's header text and payload layouts (pcd_helper.h:323-371, 518-735) and to
main_test.cpp:188-208's save / load round trip."""
import base64
import json
import os
import struct

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HEADER = ("# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z rgba stamp_id\nSIZE 8 8 8 4 4\n"
          "TYPE F F F U U\nCOUNT 1 1 1 1 1\nWIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA %s\n")


@pytest.fixture(scope="module")
def pcp():
    import pointcloudprocess_b200 as m
    return m


def make(pcp, n, seed=3):
    rng = np.random.default_rng(seed)
    xyz = rng.uniform(0, 50, (n, 3)) + [483189.100569, 3371741.645173, 430.094726]
    return pcp.make_cloud(xyz, rgba=rng.integers(0, 1 << 24, n, dtype=np.uint32), stamp_id=np.arange(n) % 10)


def fields_equal(a, b):
    return all(a[f].tobytes() == b[f].tobytes() for f in ("x", "y", "z", "w", "rgba", "stamp_id"))


def test_lzf_streams_equal_the_reference(pcp, orc):
    g = json.load(open(os.path.join(GOLD, "lzf_ref.json")))
    for c in g["cases"]:
        raw, lzf = base64.b64decode(c["raw"]), base64.b64decode(c["lzf"])
        assert pcp.io.lzfCompress(raw) == lzf
        assert pcp.io.lzfDecompress(lzf, len(raw)) == raw
    rng = np.random.default_rng(0)
    ref = orc.ref_lzf()
    for n in (1, 2, 3, 4, 17, 18, 19, 20, 21, 33, 263, 264, 265, 266, 267, 1000, 8191, 8192, 8193, 70000):
        for kind in range(4):
            if kind == 0:
                raw = bytes(rng.integers(0, 256, n, dtype=np.uint8))          # incompressible
            elif kind == 1:
                raw = bytes(rng.integers(0, 3, n, dtype=np.uint8))            # short matches
            elif kind == 2:
                raw = (b"abcdefghijklmnopqrstuvwxyz0123456789" * (n // 36 + 1))[:n]  # long matches up to the end
            else:
                raw = np.cumsum(rng.integers(0, 2, n // 8 + 1)).astype(np.float64).tobytes()[:n]  # field planes
            mine = pcp.io.lzfCompress(raw)
            assert mine == orc.lzf_compress(raw), (n, kind)
            if mine:
                assert pcp.io.lzfDecompress(mine, len(raw)) == raw
            if ref is not None and mine:
                import ctypes
                out = ctypes.create_string_buffer(int(len(raw) * 1.5) + 16)
                m = ref.ref_lzf_compress(raw, len(raw), out, len(out))
                assert out.raw[:m] == mine, (n, kind)
    assert pcp.io.lzfDecompress(b"\x05ab", 10) == b""   # truncated literal run
    assert pcp.io.lzfDecompress(b"\x20\x00", 10) == b""  # back reference before the start


def test_reference_double_io_round_trip(pcp, tmp_path):
    """main_test.cpp:188-208: ten points at pi*i saved binary and loaded back."""
    cloud = pcp.PointCloud()
    for i in range(10):
        cloud.push_back(np.pi * i, np.pi * i, np.pi * i)
    path = tmp_path / "1.pcd"
    assert pcp.io.savePCDFileBinary(path, cloud) == 0
    back = pcp.PointCloud()
    assert pcp.io.loadPCDFile(path, back) == 0
    assert back.size() == 10 and back.is_dense and back.height == 1 and back.width == 10
    assert fields_equal(back.points, cloud.points)
    data = open(path, "rb").read()
    head = (HEADER % (10, 10, "binary")).encode()
    assert data[:len(head)] == head and len(data) == len(head) + 10 * 32  # registered fields only, packed
    assert struct.unpack_from("<dddII", data, len(head) + 3 * 32) == (np.pi * 3, np.pi * 3, np.pi * 3, 0, 0)


def test_three_payloads_round_trip(pcp, orc, tmp_path):
    c = make(pcp, 5000)
    c["x"][7] = np.nan
    for mode, fn in (("ascii", pcp.io.savePCDFileASCII), ("binary", pcp.io.savePCDFileBinary),
                     ("binary_compressed", pcp.io.savePCDFileBinaryCompressed)):
        path = tmp_path / (mode + ".pcd")
        assert fn(path, c) == 0
        h = pcp.io.readHeader(path)
        assert h["data_type"] == mode and h["points"] == 5000 and h["fields"] == ["x", "y", "z", "rgba", "stamp_id"]
        assert h["sizes"] == [8, 8, 8, 4, 4] and h["types"] == ["F", "F", "F", "U", "U"] and h["data_offset"] == len(HEADER % (5000, 5000, mode))
        back = pcp.PointCloud()
        assert pcp.io.loadPCDFile(path, back) == 0
        if mode == "ascii":  # 8 significant digits (writeASCII's default precision); dense is not checked for ascii
            assert np.isnan(back.points["x"][7]) and back.is_dense
            ok = np.arange(5000) != 7
            assert np.allclose(back.points["x"][ok], c["x"][ok], rtol=1e-7, atol=0)
            assert np.array_equal(back.points["rgba"], c["rgba"]) and np.array_equal(back.points["stamp_id"], c["stamp_id"])
        else:
            assert fields_equal(back.points, c) and not back.is_dense
    # the compressed payload: two size words, then the reference's LZF over the fields as planes
    data = open(tmp_path / "binary_compressed.pcd", "rb").read()
    off = len(HEADER % (5000, 5000, "binary_compressed"))
    csize, usize = struct.unpack_from("<II", data, off)
    planes = b"".join(np.ascontiguousarray(c[f]).tobytes() for f in ("x", "y", "z", "rgba", "stamp_id"))
    assert usize == len(planes) == 5000 * 32 and len(data) == off + 8 + csize
    assert data[off + 8:] == orc.lzf_compress(planes)
    assert open(tmp_path / "ascii.pcd").read().split("\n")[11 + 7].startswith("nan ")


def test_foreign_files_fields_are_matched_by_name_type_and_size(pcp, tmp_path):
    # binary file with a padding field, an extra field and float x (not the record's type): only y, z, rgba match
    n = 4
    head = ("VERSION .7\nFIELDS x y _ z intensity rgba\nSIZE 4 8 1 8 4 4\nTYPE F F U F F U\nCOUNT 1 1 4 1 1 1\n"
            "WIDTH 2\nHEIGHT 2\nPOINTS 4\nDATA binary\n")
    body = b"".join(struct.pack("<fd4sdfI", 1.5 + i, 10.0 + i, b"\xff" * 4, 20.0 + i, 0.5, 7 + i) for i in range(n))
    path = tmp_path / "f.pcd"
    open(path, "wb").write(head.encode() + body)
    cl = pcp.PointCloud()
    assert pcp.io.loadPCDFile(path, cl) == 0
    assert cl.height == 2 and cl.width == 2 and cl.is_dense
    assert list(cl.points["x"]) == [0.0] * 4 and list(cl.points["y"]) == [10.0, 11.0, 12.0, 13.0]
    assert list(cl.points["z"]) == [20.0, 21.0, 22.0, 23.0] and list(cl.points["rgba"]) == [7, 8, 9, 10]
    assert list(cl.points["w"]) == [1.0] * 4 and list(cl.points["stamp_id"]) == [0] * 4
    # ascii with a comment, blank lines and tabs
    txt = "# comment\nVERSION 0.7\nFIELDS x y z\nSIZE 8 8 8\nTYPE F F F\nCOUNT 1 1 1\nWIDTH 2\nHEIGHT 1\nPOINTS 2\nDATA ascii\n1 2 3\n\n4\t5 nan\n"
    open(path, "w").write(txt)
    assert pcp.io.loadPCDFile(path, cl) == 0
    assert list(cl.points["x"]) == [1.0, 4.0] and np.isnan(cl.points["z"][1]) and cl.points["z"][0] == 3.0


def test_errors_follow_the_reference(pcp, tmp_path, capsys):
    cl = pcp.PointCloud()
    assert pcp.io.loadPCDFile(tmp_path / "missing.pcd", cl) == -1
    assert "Could not open file" in capsys.readouterr().out
    bad = tmp_path / "bad.pcd"
    open(bad, "w").write("VERSION 0.7\nFIELDS x y z\nSIZE 8 8 8\nTYPE F F F\nCOUNT 1 1 1\nWIDTH 3\nHEIGHT 2\nPOINTS 5\nDATA ascii\n")
    assert pcp.io.loadPCDFile(bad, cl) == -1 and "HEIGHT x WIDTH" in capsys.readouterr().out
    open(bad, "w").write((HEADER % (3, 3, "ascii")) + "1 2 3 4 5\n")
    assert pcp.io.loadPCDFile(bad, cl) == -1 and "different than expected" in capsys.readouterr().out
    open(bad, "wb").write((HEADER % (3, 3, "binary")).encode() + b"\0" * 40)
    assert pcp.io.loadPCDFile(bad, cl) == -1
    open(bad, "wb").write((HEADER % (3, 3, "binary_compressed")).encode() + struct.pack("<II", 3, 96) + b"\x05ab")
    assert pcp.io.loadPCDFile(bad, cl) == -1 and "decompressed lzf data" in capsys.readouterr().out
    assert pcp.io.savePCDFileBinary(tmp_path / "e.pcd", pcp.PointCloud()) == -1  # "Input point cloud has no data!"
    assert "no data" in capsys.readouterr().out


def test_malformed_headers_are_rejected_not_trusted(tmp_path):
    """A PCD header is untrusted input: sizes and counts that would wrap an allocation or a stride, unknown
    types, and binary payloads that the file does not hold all fail with a status and a message -- no crash, no
    over-read, no exception through the C ABI."""
    import ctypes as C
    from pointcloudprocess_b200 import _lib
    L = _lib.load()
    good = ("# .PCD v0.7\nVERSION 0.7\nFIELDS x y z rgba stamp_id\nSIZE 8 8 8 4 4\nTYPE F F F U U\nCOUNT 1 1 1 1 1\n"
            "WIDTH 3\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS 3\nDATA binary\n")
    cases = {
        "size0": good.replace("SIZE 8 8 8 4 4", "SIZE 0 8 8 4 4"),
        "size_neg": good.replace("SIZE 8 8 8 4 4", "SIZE -8 8 8 4 4"),
        "size3": good.replace("SIZE 8 8 8 4 4", "SIZE 3 8 8 4 4"),
        "count0": good.replace("COUNT 1 1 1 1 1", "COUNT 0 1 1 1 1"),
        "count_huge": good.replace("COUNT 1 1 1 1 1", "COUNT 2000000000 1 1 1 1"),
        "type": good.replace("TYPE F F F U U", "TYPE F Q F U U"),
        "float2": good.replace("SIZE 8 8 8 4 4", "SIZE 2 8 8 4 4"),
        "points_huge": good.replace("WIDTH 3", "WIDTH 4000000000").replace("POINTS 3", "POINTS 4000000000"),
        "short_payload": good,  # 3 x 32 bytes promised, 10 delivered
    }
    pts = (C.c_char * (48 * 16))()
    n, dense = C.c_size_t(0), C.c_int(1)
    hdr = _lib.PcdHeader()
    for name, text in cases.items():
        path = tmp_path / (name + ".pcd")
        path.write_bytes(text.encode() + b"\0" * 10)
        st = L.pcp_pcd_load(str(path).encode(), 0, pts, 16, 48, C.byref(n), C.byref(dense))
        assert st < 0, name
        assert len(L.pcp_last_error(None)) > 0
        assert L.pcp_pcd_read_header(str(path).encode(), C.byref(hdr)) < 0, name
    ok = tmp_path / "ok.pcd"
    ok.write_bytes(good.encode() + b"\0" * 96)
    assert L.pcp_pcd_load(str(ok).encode(), 0, pts, 16, 48, C.byref(n), C.byref(dense)) == 0 and n.value == 3
    # binary_compressed: size words that promise more than the file holds / planes beyond the uncompressed size
    comp = good.replace("DATA binary", "DATA binary_compressed")
    bad = tmp_path / "comp.pcd"
    bad.write_bytes(comp.encode() + (1 << 30).to_bytes(4, "little") + (96).to_bytes(4, "little") + b"\0" * 8)
    assert L.pcp_pcd_load(str(bad).encode(), 0, pts, 16, 48, C.byref(n), C.byref(dense)) < 0
