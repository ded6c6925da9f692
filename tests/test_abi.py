"""CPU tests of the boundary: libpcp_cuda.so loads and exports every symbol include/pcp_cuda.h declares;
no compute call is made (there is no GPU here) and pcp_create fails loudly instead of falling back."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "pcp_cuda.h")).read()
    return re.findall(r"PCP_API\s+[\w\s\*]+?\b(pcp_\w+)\s*\(", text)


def test_header_and_binding_agree():
    from pointcloudprocess_b200 import _lib
    assert sorted(header_symbols()) == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol():
    from pointcloudprocess_b200 import _lib
    L = _lib.load()
    for name in header_symbols():
        assert hasattr(L, name), name
    assert b"sm_100a" in L.pcp_version()


def test_struct_layouts_match_reference():
    import pointcloudprocess_b200 as pcp
    # PointXYZRGBA point_type.h:9-82 ; LAS_POINT_PROPERTY data_struct.h:161-173
    assert pcp.POINT_DTYPE.itemsize == 48 and pcp.POINT_DTYPE.fields["rgba"][1] == 32
    assert pcp.POINT_DTYPE.fields["stamp_id"][1] == 36
    assert pcp.PROP_DTYPE.itemsize == 48
    assert [pcp.PROP_DTYPE.fields[f][1] for f in ("normal_x", "Distance", "curvature", "PointID", "SegmentID",
                                                  "dis_from_point_plane")] == [0, 16, 24, 32, 36, 40]


def test_no_cpu_fallback_without_gpu():
    import pointcloudprocess_b200 as pcp
    L = pcp.load_library()
    if L.pcp_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pcp.PcpError) as e:
        pcp.Context(0)
    assert "no CPU fallback" in str(e.value)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "pointcloudprocess_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"(from|import)\s+oracle|liboracle|oracle\.h|oracle/|orc_\w+\(", text), \
                    os.path.join(dirpath, f)
    for f in os.listdir(os.path.join(ROOT, "include")):
        p = os.path.join(ROOT, "include", f)
        if os.path.isfile(p):
            assert "liboracle" not in open(p).read()
