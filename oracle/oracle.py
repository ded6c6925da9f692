"""ctypes doorway onto oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module.  The product package (pointcloudprocess_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

# PointXYZRGBA, /root/reference/point_type.h:9-82 (48 bytes, 16-aligned)
POINT_DTYPE = np.dtype(
    {"names": ["x", "y", "z", "w", "rgba", "stamp_id"],
     "formats": ["<f8", "<f8", "<f8", "<f8", "<u4", "<u4"],
     "offsets": [0, 8, 16, 24, 32, 36], "itemsize": 48})
# LAS_POINT_PROPERTY, /root/reference/data_struct.h:161-173 (48 bytes)
PROP_DTYPE = np.dtype(
    {"names": ["normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID", "SegmentID",
               "dis_from_point_plane"],
     "formats": ["<f4", "<f4", "<f4", "<f8", "<f8", "<i4", "<i4", "<f4"],
     "offsets": [0, 4, 8, 16, 24, 32, 36, 40], "itemsize": 48})
PLAN_DTYPE = np.dtype(
    {"names": ["normal_x", "normal_y", "normal_z", "min_value", "curvature", "Distance", "eig"],
     "formats": ["<f4", "<f4", "<f4", "<f4", "<f4", "<f4", ("<f8", (3,))],
     "offsets": [0, 4, 8, 12, 16, 20, 24], "itemsize": 48})


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, capture_output=True)
    ref = os.path.join(_HERE, "_ref", "liblzf_ref.so")
    if (force or not os.path.exists(ref)) and os.path.exists("/root/reference/lzf.cpp"):
        subprocess.run(["make", "-C", _HERE, "ref"], check=True, capture_output=True)
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _declare(_lib)
    return _lib


def ref_lzf():
    """The reference's own lzf.cpp (compiled, oracle/_ref) or None when it was never built."""
    p = os.path.join(_HERE, "_ref", "liblzf_ref.so")
    if not os.path.exists(p):
        return None
    L = C.CDLL(p)
    for f in (L.ref_lzf_compress, L.ref_lzf_decompress):
        f.restype = C.c_uint
        f.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_uint]
    return L


def _declare(L):
    vp, sz, i32, i64, dbl = C.c_void_p, C.c_size_t, C.c_int, C.c_int64, C.c_double
    L.orc_num_threads.restype = i32
    L.orc_jacobi.argtypes = [vp, i32, vp, vp]
    L.orc_plane_fit.argtypes = [vp, i32, vp]
    L.orc_valid_points.restype = sz
    L.orc_valid_points.argtypes = [vp, sz, sz, vp]
    L.orc_knn_brute.argtypes = [vp, sz, sz, vp, sz, sz, i32, vp, vp, vp]
    L.orc_kdtree_build.restype = vp
    L.orc_kdtree_build.argtypes = [vp, sz, sz]
    L.orc_kdtree_free.argtypes = [vp]
    L.orc_kdtree_size.restype = sz
    L.orc_kdtree_size.argtypes = [vp]
    L.orc_kdtree_knn.argtypes = [vp, vp, sz, sz, i32, vp, vp, vp, i32]
    L.orc_kdtree_radius.argtypes = [vp, vp, sz, sz, dbl, C.c_uint32, vp, vp, vp, i32]
    L.orc_count_boundary_ties.restype = i64
    L.orc_count_boundary_ties.argtypes = [vp, vp, sz, sz, i32, i32]
    L.orc_features_knn.argtypes = [vp, sz, sz, i32, vp, vp, vp, vp, i32]
    L.orc_features_radius.argtypes = [vp, sz, sz, dbl, C.c_uint32, vp, vp, vp, vp, i32]
    L.orc_point_plane_distance.restype = C.c_float
    L.orc_point_plane_distance.argtypes = [vp, C.c_float, C.c_float, C.c_float, C.c_float]
    L.orc_region_growing.restype = i64
    L.orc_region_growing.argtypes = [vp, sz, sz, vp, i32, dbl, dbl, dbl, i32, vp, vp, vp, i32]
    L.orc_voxel_grid.restype = i64
    L.orc_voxel_grid.argtypes = [vp, sz, sz, vp, i32, i32, vp, vp, vp, vp, vp, vp]
    L.orc_remove_duplicate.restype = i64
    L.orc_remove_duplicate.argtypes = [vp, sz, sz, C.c_float, vp]
    L.orc_centroid.argtypes = [vp, sz, sz, i32, vp, vp]
    L.orc_minmax.argtypes = [vp, sz, sz, i32, vp, vp]
    for f in (L.orc_lzf_compress, L.orc_lzf_decompress):
        f.restype = C.c_uint
        f.argtypes = [vp, C.c_uint, vp, C.c_uint]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _pts(a):
    """Accept a POINT_DTYPE array or an (n,3) float64 array; return (array, n, stride)."""
    a = np.ascontiguousarray(a)
    if a.dtype.names:
        return a, a.shape[0], a.dtype.itemsize
    a = np.ascontiguousarray(a, dtype=np.float64)
    assert a.ndim == 2 and a.shape[1] >= 3
    return a, a.shape[0], a.strides[0]


def make_cloud(xyz, rgba=None, stamp_id=None):
    xyz = np.asarray(xyz, dtype=np.float64)
    c = np.zeros(xyz.shape[0], dtype=POINT_DTYPE)
    c["x"], c["y"], c["z"], c["w"] = xyz[:, 0], xyz[:, 1], xyz[:, 2], 1.0
    if rgba is not None:
        c["rgba"] = rgba
    if stamp_id is not None:
        c["stamp_id"] = stamp_id
    return c


def jacobi(A):
    A = np.array(A, dtype=np.float64, order="C")
    n = A.shape[0]
    W = np.zeros(n)
    V = np.zeros((n, n))
    lib().orc_jacobi(_p(A), n, _p(W), _p(V))
    return W, V


def plane_fit(xyz):
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    out = np.zeros(1, dtype=PLAN_DTYPE)
    lib().orc_plane_fit(_p(xyz), xyz.shape[0], _p(out))
    return out[0]


def knn_brute(pts, queries, k, want_d2=True):
    a, n, s = _pts(pts)
    q, nq, qs = _pts(queries)
    idx = np.empty((nq, k), dtype=np.int32)
    d2 = np.empty((nq, k), dtype=np.float64) if want_d2 else None
    found = np.empty(nq, dtype=np.int32)
    lib().orc_knn_brute(_p(a), n, s, _p(q), nq, qs, k, _p(idx), _p(d2), _p(found))
    return idx, d2, found


class KdTree:
    def __init__(self, pts):
        self._a, n, s = _pts(pts)
        self._h = lib().orc_kdtree_build(_p(self._a), n, s)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_kdtree_free(self._h)
            self._h = None

    def size(self):
        return lib().orc_kdtree_size(self._h)

    def knn(self, queries, k, threads=0, want_d2=True):
        q, nq, qs = _pts(queries)
        idx = np.empty((nq, k), dtype=np.int32)
        d2 = np.empty((nq, k), dtype=np.float64) if want_d2 else None
        found = np.empty(nq, dtype=np.int32)
        lib().orc_kdtree_knn(self._h, _p(q), nq, qs, k, _p(idx), _p(d2), _p(found), threads)
        return idx, d2, found

    def radius(self, queries, r, max_nn=0, threads=0):
        q, nq, qs = _pts(queries)
        off = np.zeros(nq + 1, dtype=np.int64)
        lib().orc_kdtree_radius(self._h, _p(q), nq, qs, float(r), max_nn, _p(off), None, None, threads)
        idx = np.empty(off[-1], dtype=np.int32)
        d2 = np.empty(off[-1], dtype=np.float64)
        lib().orc_kdtree_radius(self._h, _p(q), nq, qs, float(r), max_nn, _p(off), _p(idx), _p(d2), threads)
        return off, idx, d2

    def boundary_ties(self, queries, k, threads=0):
        q, nq, qs = _pts(queries)
        return int(lib().orc_count_boundary_ties(self._h, _p(q), nq, qs, k, threads))


def features_knn(pts, k, want_eig=True, want_lps=True, want_idx=False, threads=0):
    a, n, s = _pts(pts)
    props = np.zeros(n, dtype=PROP_DTYPE)
    eig = np.zeros((n, 3)) if want_eig else None
    lps = np.zeros((n, 3), dtype=np.float32) if want_lps else None
    idx = np.empty((n, k), dtype=np.int32) if want_idx else None
    lib().orc_features_knn(_p(a), n, s, k, _p(props), _p(eig), _p(lps), _p(idx), threads)
    return props, eig, lps, idx


def features_rpca(pts, k=20, Pr=0.99, epi=0.5, seed=1, threads=0):
    """calculate_plan_parameter_rpca with the counter-based generator (oracle_knn.cpp); returns props."""
    a, n, s = _pts(pts)
    props = np.zeros(n, dtype=PROP_DTYPE)
    f = lib().orc_features_rpca
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_float, C.c_float, C.c_uint64, C.c_void_p, C.c_int]
    f(_p(a), n, s, k, Pr, epi, seed, _p(props), threads)
    return props


def rpca_iterations(Pr=0.99, epi=0.5):
    f = lib().orc_rpca_iterations
    f.restype = C.c_int
    f.argtypes = [C.c_float, C.c_float]
    return f(Pr, epi)


def features_radius(pts, r, max_nn=0, threads=0):
    a, n, s = _pts(pts)
    props = np.zeros(n, dtype=PROP_DTYPE)
    eig = np.zeros((n, 3))
    lps = np.zeros((n, 3), dtype=np.float32)
    counts = np.zeros(n, dtype=np.int32)
    lib().orc_features_radius(_p(a), n, s, float(r), max_nn, _p(props), _p(eig), _p(lps), _p(counts), threads)
    return props, eig, lps, counts


def region_growing(pts, props, k=50, distanceT=0.5, cosfaT=0.94, curv_t=0.005, min_points=5, threads=0):
    """TreeExtration::region_growning (extraction_tree.cpp:170-272).  Returns dict(props (copy with SegmentID),
    offsets (nseg+1), points (members in PlanSegment::PointID order), planes (nseg x 4: seed normal, Distance))."""
    a, n, stride = _pts(pts)
    pr = np.ascontiguousarray(props, dtype=PROP_DTYPE).copy()
    assert len(pr) == n
    offsets = np.zeros(n // (max(min_points, 0) + 1) + 2, dtype=np.int64)
    points = np.zeros(max(n, 1), dtype=np.int32)
    planes = np.zeros((len(offsets), 4), dtype=np.float32)
    nseg = lib().orc_region_growing(_p(a), n, stride, _p(pr), k, distanceT, cosfaT, curv_t, min_points, _p(offsets),
                                    _p(points), _p(planes), threads)
    return {"props": pr, "offsets": offsets[:nseg + 1].copy(), "points": points[:offsets[nseg]].copy(),
            "planes": planes[:nseg].copy()}


def voxel_grid(cloud, leaf, downsample_all=True, is_dense=True):
    a = np.ascontiguousarray(cloud)
    assert a.dtype == POINT_DTYPE
    n = a.shape[0]
    leaf3 = np.array([leaf] * 3 if np.isscalar(leaf) else leaf, dtype=np.float64)
    out = np.zeros(max(n, 1), dtype=POINT_DTYPE)
    min_b = np.zeros(3, dtype=np.int32)
    div_b = np.zeros(3, dtype=np.int32)
    cell = np.zeros((max(n, 1), 3), dtype=np.int32)
    cnt = np.zeros(max(n, 1), dtype=np.int32)
    ovf = C.c_int(0)
    m = lib().orc_voxel_grid(_p(a), n, 48, _p(leaf3), int(downsample_all), int(is_dense), _p(out), _p(min_b),
                             _p(div_b), _p(cell), _p(cnt), C.byref(ovf))
    return {"points": out[:m].copy(), "min_b": min_b, "div_b": div_b, "point_cell": cell[:n],
            "cell_count": cnt[:m].copy(), "overflow32": bool(ovf.value)}


def remove_duplicate(cloud, leaf):
    a = np.array(cloud, copy=True)
    assert a.dtype == POINT_DTYPE
    c = np.zeros(3)
    m = lib().orc_remove_duplicate(_p(a), a.shape[0], 48, float(np.float32(leaf)), _p(c))
    return a[:m].copy(), c


def centroid(cloud, is_dense=True):
    a, n, s = _pts(cloud)
    c = np.zeros(3)
    cnt = C.c_uint32(0)
    lib().orc_centroid(_p(a), n, s, int(is_dense), _p(c), C.byref(cnt))
    return c, cnt.value


def minmax(cloud, is_dense=True):
    a, n, s = _pts(cloud)
    mn, mx = np.zeros(3), np.zeros(3)
    lib().orc_minmax(_p(a), n, s, int(is_dense), _p(mn), _p(mx))
    return mn, mx


def lzf_compress(data: bytes):
    src = np.frombuffer(data, dtype=np.uint8)
    out = np.zeros(len(data) + len(data) // 16 + 64, dtype=np.uint8)
    n = lib().orc_lzf_compress(_p(src), len(data), _p(out), out.size)
    return out[:n].tobytes()


def lzf_decompress(data: bytes, out_len: int):
    src = np.frombuffer(data, dtype=np.uint8)
    out = np.zeros(max(out_len, 1), dtype=np.uint8)
    n = lib().orc_lzf_decompress(_p(src), len(data), _p(out), out_len)
    return out[:n].tobytes()
