"""The C++ facade (include/pcp/pcp_facade.hpp) compiled into the reference's own smoke functions
(tests/cpp/main_test_port.cpp = main_test.cpp:126-188 against our headers)."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    cxx = shutil.which("g++") or "/usr/bin/g++"
    exe = str(tmp_path / "main_test_port")
    libdir = os.path.join(ROOT, "pointcloudprocess_b200")
    r = subprocess.run([cxx, "-std=c++17", "-O1", "-Wall", "-fopenmp", os.path.join(ROOT, "tests", "cpp", "main_test_port.cpp"), "-o", exe,
                        "-L" + libdir, "-lpcp_cuda", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_facade_compiles_and_fails_loudly_without_gpu(tmp_path):
    import pointcloudprocess_b200 as pcp
    exe = _build(tmp_path)
    if pcp.load_library().pcp_device_count() > 0:
        pytest.skip("a GPU is present")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout


def test_reference_double_io_through_the_facade(tmp_path):
    """main_test.cpp:188-208 against our headers; PCD I/O is host code, so this runs without a GPU."""
    r = subprocess.run([_build(tmp_path), "io", str(tmp_path / "1.pcd")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().split("\n")
    assert lines[0] == "io 10"
    for i in range(10):
        x, y, z, rgba = lines[1 + i].split()
        assert float(x) == float(y) == float(z) == np.pi * i and rgba == "0"
    assert "Could not open file" in lines[11] and lines[12] == "missing -1"


@pytest.mark.gpu
def test_reference_smoke_functions_known_answers(tmp_path):
    r = subprocess.run([_build(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().split("\n")
    assert lines[0] == "voxel 10"  # 110 points -> 10 cells of 11 (main_test.cpp:126-154)
    for c in range(10):
        vals = [np.float32(c)] + [np.float32(0.1 * i) for i in range(10 * c, 10 * c + 10)]
        exp = sum(float(v) for v in vals) / 11.0
        assert [float(t) for t in lines[1 + c].split()] == [exp, exp, exp]
    assert lines[11] == "layout 4 -1 7 1000"  # voxel_grid.h:595-706 accessors on the saved leaf layout
    lines = lines[:11] + lines[12:]
    assert lines[11] == "knn 10"   # main_test.cpp:156-188: neighbours of the origin are 0..9 in order
    for i in range(10):
        idx, d2 = lines[12 + i].split()
        d = 0.1 * i
        assert int(idx) == i and float(d2) == (d * d + d * d) + d * d
    assert lines[22] == "batch 100 of 100; by index 3: 5 4 6"  # ties by index: 4 before 6
    lines = lines[:22] + lines[23:]
    # remove_duplicate centres on the centroid first (point_cloud_helper.cpp:42-63); the cell index truncates towards
    # zero (voxel_grid.h:74-86), so the 100 points on [-4.95, 4.95] fall into cells -4..4
    xs = np.array([0.1 * i for i in range(100)])
    cells = np.trunc((xs - xs.mean()) / 1.0)
    assert lines[22] == "dedup %d" % len(np.unique(cells)) == "dedup 9"
    assert lines[23].startswith("props 0 ")
    # calculate_plan_parameter_rpca through the C++ mirror == the oracle, record bytes hashed (FNV-1a)
    from oracle import oracle as orc
    i = np.arange(2000)
    P = np.c_[0.01 * (i % 50), 0.01 * (i // 50), 0.001 * ((i * 7919) % 13)]
    ref = orc.features_rpca(P, 20, 0.99, 0.5, 4).tobytes()
    h = 1469598103934665603
    for b in ref:
        h = ((h ^ b) * 1099511628211) & (2**64 - 1)
    assert lines[24] == "rpca 34 48 %d" % h
    # point_segment + tree_filter (static.cpp:8-36) through the C++ mirror == the oracle's region growing
    a = np.arange(6000) % 3000
    x = 0.02 * (a % 60) + 1e-4 * ((a * 7919) % 17)
    Q = np.c_[x, 0.02 * (a // 60), np.where(np.arange(6000) < 3000, 1e-4 * ((a * 31) % 11), 2.0 + 0.5 * x)]
    rg = orc.region_growing(Q, orc.features_rpca(Q, 20, 0.99, 0.5, 4), 50, 0.5, 0.940)
    h = 1469598103934665603
    for v in rg["points"]:
        h = ((h ^ int(v)) * 1099511628211) & (2**64 - 1)
    sizes = np.diff(rg["offsets"])
    assert lines[25] == "segment %d %d %d %d" % (len(sizes), len(rg["points"]), h, sizes[sizes > 50].sum())
    assert len(sizes) >= 2
    # the reference's OpenMP loop (calculate_feature.cpp:222-237) over one shared tree, the kd_tree_lod call shape and
    # the header / sensor pose copied by filter (voxel_grid.h:440-442)
    i = np.arange(400)
    G = np.c_[0.1 * (i % 20), 0.1 * (i // 20), 0.01 * ((i * 7919) % 13)]
    idx, d2, found = orc.knn_brute(G, G, 20)
    assert (idx[:, 0] == i).all() and (d2[:, 0] == 0).all()
    assert lines[26] == "omp %d 1 7 0 42 velodyne 1.5 0.5" % int(idx[:, 19].sum())
    # pcp_mgpu_* through the C ABI: same points, same records as the single-context calls
    t = lines[27].split()
    assert t[0] == "mgpu" and t[1] == t[2] and int(t[1]) > 1000 and t[3:] == ["1", "1"]
