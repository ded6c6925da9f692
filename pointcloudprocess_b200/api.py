"""Host-side mirror of the reference's interface for the hot path, over the C ABI (include/pcp_cuda.h).

Names, argument meaning and return conventions follow the reference classes:
  KdTreeFLANN          kd_tree.h:772-903       setInputCloud / nearestKSearch / radiusSearch
  VoxelGrid            voxel_grid.h:482-773    setInputCloud / setLeafSize / setDownsampleAllData / filter
  CalculateFeature     calculate_feature.h:11-16
  PointCloudHelper     point_cloud_helper.h:16-90,192-230   remove_duplicate / getMinMax3D / compute3DCentroid
Clouds are numpy structured arrays of POINT_DTYPE (PointXYZRGBA, point_type.h:9-82) on the host, or
DeviceArray handles for data that stays in HBM.  All compute happens in libpcp_cuda.so on the GPU.
"""
import ctypes as C

import numpy as np

from . import _lib

POINT_DTYPE = np.dtype(
    {"names": ["x", "y", "z", "w", "rgba", "stamp_id"],
     "formats": ["<f8", "<f8", "<f8", "<f8", "<u4", "<u4"],
     "offsets": [0, 8, 16, 24, 32, 36], "itemsize": 48})
PROP_DTYPE = np.dtype(
    {"names": ["normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID", "SegmentID",
               "dis_from_point_plane"],
     "formats": ["<f4", "<f4", "<f4", "<f8", "<f8", "<i4", "<i4", "<f4"],
     "offsets": [0, 4, 8, 16, 24, 32, 36, 40], "itemsize": 48})
PACKED_POINT_DTYPE = np.dtype(  # stride-32 cloud record = binary PCD payload record
    {"names": ["x", "y", "z", "rgba", "stamp_id"], "formats": ["<f8", "<f8", "<f8", "<u4", "<u4"],
     "offsets": [0, 8, 16, 24, 28], "itemsize": 32})
PACKED_PROP_DTYPE = np.dtype(  # pcp_point_property_packed
    {"names": ["normal_x", "normal_y", "normal_z", "Distance", "curvature", "PointID"],
     "formats": ["<f4", "<f4", "<f4", "<f4", "<f4", "<i4"], "offsets": [0, 4, 8, 12, 16, 20], "itemsize": 24})
PLAN_DTYPE = np.dtype(
    {"names": ["normal_x", "normal_y", "normal_z", "min_value", "curvature", "Distance", "eig"],
     "formats": ["<f4", "<f4", "<f4", "<f4", "<f4", "<f4", ("<f8", (3,))],
     "offsets": [0, 4, 8, 12, 16, 20, 24], "itemsize": 48})

SYNTH_KINDS = {"uniform_cube": 0, "strip": 1, "forest": 2, "city": 3, "lattice": 4}


def make_cloud(xyz, rgba=None, stamp_id=None):
    xyz = np.asarray(xyz, dtype=np.float64)
    c = np.zeros(xyz.shape[0], dtype=POINT_DTYPE)
    c["x"], c["y"], c["z"], c["w"] = xyz[:, 0], xyz[:, 1], xyz[:, 2], 1.0
    if rgba is not None:
        c["rgba"] = rgba
    if stamp_id is not None:
        c["stamp_id"] = stamp_id
    return c


class DeviceArray:
    """A device allocation owned by the library (pcp_dev_alloc)."""

    def __init__(self, ctx, nbytes):
        self.ctx, self.nbytes = ctx, int(nbytes)
        p = C.c_void_p()
        ctx._check(ctx.L.pcp_dev_alloc(ctx.h, max(self.nbytes, 1), C.byref(p)))
        self.ptr = p.value

    def free(self):
        if self.ptr:
            self.ctx.L.pcp_dev_free(self.ctx.h, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def to_host(self, dtype, count=None):
        dtype = np.dtype(dtype)
        n = self.nbytes // dtype.itemsize if count is None else count
        out = np.empty(n, dtype=dtype)
        self.ctx.memcpy(out.ctypes.data, self.ptr, n * dtype.itemsize)
        return out

    def from_host(self, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        self.ctx.memcpy(self.ptr, arr.ctypes.data, arr.nbytes)


def _ptr(a):
    """numpy array / DeviceArray / int address / None -> c_void_p value"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if isinstance(a, DeviceArray):
        return a.ptr
    return int(a)


def _pts(a):
    """(pointer, count, stride) of a cloud given as POINT_DTYPE array, (n,>=3) float64 array or
    (ptr, n, stride) tuple of a device buffer."""
    if isinstance(a, tuple):
        p, n, s = a
        return _ptr(p), int(n), int(s), None
    a = np.ascontiguousarray(a)
    if a.dtype.names:
        return a.ctypes.data, a.shape[0], a.dtype.itemsize, a
    a = np.ascontiguousarray(a, dtype=np.float64)
    if a.ndim == 1:
        a = a.reshape(1, -1)
    assert a.shape[1] >= 3
    return a.ctypes.data, a.shape[0], a.strides[0], a


class Context:
    """pcp_ctx: one per GPU."""

    def __init__(self, device=0):
        self.L = _lib.load()
        h = C.c_void_p()
        st = self.L.pcp_create(C.byref(h), int(device))
        if st != 0:
            msg = self.L.pcp_last_error(None)
            raise _lib.PcpError(st, msg.decode() if msg else "")
        self.h = h
        self.device = device

    def _check(self, st):
        _lib.check(self.h, st)

    def close(self):
        if getattr(self, "h", None):
            self.L.pcp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        """cuda_stream: a cudaStream_t handle (int).  None restores the context's own stream.  0 -- the handle
        torch reports for its default stream -- selects the legacy default stream explicitly (cudaStreamLegacy,
        (cudaStream_t)1): a NULL handle would mean "own stream" to pcp_set_stream, and the library's kernels would
        then race with the caller's work on stream 0."""
        handle = None if cuda_stream is None else (int(cuda_stream) or 1)
        self._check(self.L.pcp_set_stream(self.h, handle))

    def synchronize(self):
        self._check(self.L.pcp_synchronize(self.h))

    def memcpy(self, dst, src, nbytes):
        self._check(self.L.pcp_memcpy(self.h, _ptr(dst), _ptr(src), int(nbytes)))

    def memcpy_async(self, dst, src, nbytes):
        self._check(self.L.pcp_memcpy_async(self.h, _ptr(dst), _ptr(src), int(nbytes)))

    # copy lanes (pcp_copy_async ...): host<->device copies that overlap the compute stream
    def copy_async(self, lane, dst, src, nbytes):
        self._check(self.L.pcp_copy_async(self.h, int(lane), _ptr(dst), _ptr(src), int(nbytes)))

    def copy_fence(self, lane):
        self._check(self.L.pcp_copy_fence(self.h, int(lane)))

    def copy_wait(self, lane):
        self._check(self.L.pcp_copy_wait(self.h, int(lane)))

    def alloc(self, nbytes):
        return DeviceArray(self, nbytes)

    def pinned(self, nbytes, dtype=np.uint8):
        """numpy view over cudaHostAlloc memory (freed with free_pinned)."""
        p = C.c_void_p()
        self._check(self.L.pcp_host_alloc_pinned(self.h, int(nbytes), C.byref(p)))
        buf = (C.c_uint8 * int(nbytes)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=np.uint8).view(dtype)
        arr_ptr = p.value
        return arr, arr_ptr

    def free_pinned(self, ptr):
        self._check(self.L.pcp_host_free_pinned(self.h, ptr))

    def profile(self, on=True):
        self._check(self.L.pcp_profile_enable(self.h, int(on)))

    def profile_reset(self):
        self._check(self.L.pcp_profile_reset(self.h))

    def profile_get(self):
        arr = (_lib.StageTime * 64)()
        n = C.c_int(0)
        self._check(self.L.pcp_profile_get(self.h, arr, 64, C.byref(n)))
        return [{"name": arr[i].name.decode(), "ms": arr[i].ms, "launches": arr[i].launches, "bytes": arr[i].bytes}
                for i in range(min(n.value, 64))]

    def launch_count(self):
        return int(self.L.pcp_launch_count(self.h))

    def trim(self):
        self._check(self.L.pcp_trim(self.h))

    # ---- raw entry points (host or device pointers) ----
    def synth_cloud(self, kind, n_total, first=0, n=None, seed=1, centred=False, out=None):
        n = n_total - first if n is None else n
        kind = SYNTH_KINDS[kind] if isinstance(kind, str) else kind
        ret = None
        if out is None:
            ret = np.zeros(n, dtype=POINT_DTYPE)
            out = ret
        self._check(self.L.pcp_synth_cloud(self.h, kind, n_total, first, n, seed, int(centred), _ptr(out)))
        return ret

    def voxel_grid(self, cloud, leaf, downsample_all=True, out=None, want_cells=False):
        p, n, s, keep = _pts(cloud)
        leaf3 = np.array([leaf] * 3 if np.isscalar(leaf) else leaf, dtype=np.float64)
        host_out = out is None
        if host_out:
            out = np.zeros(max(n, 1), dtype=POINT_DTYPE if s == 48 else np.dtype((np.void, s)))
        m = C.c_size_t(0)
        min_b = np.zeros(3, dtype=np.int32)
        div_b = np.zeros(3, dtype=np.int32)
        cell = np.zeros((max(n, 1), 3), dtype=np.int32) if want_cells else None
        cnt = np.zeros(max(n, 1), dtype=np.int32) if want_cells else None
        self._check(self.L.pcp_voxel_grid(self.h, p, n, s, leaf3.ctypes.data, int(downsample_all), _ptr(out),
                                          C.byref(m), min_b.ctypes.data, div_b.ctypes.data, _ptr(cell), _ptr(cnt)))
        res = {"m": m.value, "min_b": min_b, "div_b": div_b}
        if host_out:
            res["points"] = out[:m.value]
        if want_cells:
            res["point_cell"] = cell[:n]
            res["cell_count"] = cnt[:m.value]
        return res

    def remove_duplicate(self, cloud, leaf):
        """In place on a (ptr, n, stride) device tuple (returns m) or on a copy of a host array (returns array)."""
        if isinstance(cloud, tuple):
            p, n, s, _ = _pts(cloud)
            nn = C.c_size_t(n)
            self._check(self.L.pcp_remove_duplicate(self.h, p, C.byref(nn), s, C.c_float(leaf)))
            return nn.value
        a = np.array(cloud, copy=True)
        nn = C.c_size_t(a.shape[0])
        self._check(self.L.pcp_remove_duplicate(self.h, a.ctypes.data, C.byref(nn), a.dtype.itemsize, C.c_float(leaf)))
        return a[:nn.value].copy()

    def remove_duplicate_to(self, cloud, leaf, out, centre=None, min_b=None, div_b=None):
        """Out of place (pcp_remove_duplicate_to): cloud is a (ptr, n, stride) tuple or host array, `out` a device
        pointer / host array with room for n records; returns m."""
        p, n, s, keep = _pts(cloud)
        m = C.c_size_t(0)
        c = None if centre is None else np.ascontiguousarray(centre, dtype=np.float64)
        mb = None if min_b is None else np.ascontiguousarray(min_b, dtype=np.int32)
        db = None if div_b is None else np.ascontiguousarray(div_b, dtype=np.int32)
        self._check(self.L.pcp_remove_duplicate_to(self.h, p, n, s, C.c_float(leaf), _ptr(c), _ptr(mb), _ptr(db), _ptr(out),
                                                   C.byref(m)))
        return m.value

    def cloud_pack(self, pts48, n, pts32):
        self._check(self.L.pcp_cloud_pack(self.h, _ptr(pts48), int(n), _ptr(pts32)))

    def cloud_unpack(self, pts32, n, pts48):
        self._check(self.L.pcp_cloud_unpack(self.h, _ptr(pts32), int(n), _ptr(pts48)))

    def props_pack(self, props, n, packed):
        self._check(self.L.pcp_props_pack(self.h, _ptr(props), int(n), _ptr(packed)))

    def props_unpack(self, packed, n=None, threads=0):
        packed = np.ascontiguousarray(packed)
        n = len(packed) if n is None else n
        out = np.zeros(n, dtype=PROP_DTYPE)
        self._check(self.L.pcp_props_unpack(packed.ctypes.data, n, out.ctypes.data, int(threads)))
        return out

    # ---- sharding primitives ----
    def remove_duplicate_centred(self, cloud, leaf, centre, min_b, div_b):
        """In place on a (ptr, n, stride) device tuple; returns m."""
        p, n, s, _ = _pts(cloud)
        nn = C.c_size_t(n)
        c = np.ascontiguousarray(centre, dtype=np.float64)
        mb = np.ascontiguousarray(min_b, dtype=np.int32)
        db = np.ascontiguousarray(div_b, dtype=np.int32)
        self._check(self.L.pcp_remove_duplicate_centred(self.h, p, C.byref(nn), s, C.c_float(leaf), c.ctypes.data,
                                                        mb.ctypes.data, db.ctypes.data))
        return nn.value

    def serial_sum(self, cloud, sum3, count):
        """Continues compute3DCentroid's serial sums over a chunk; returns (sum3, count)."""
        p, n, s, keep = _pts(cloud)
        acc = np.array(sum3, dtype=np.float64)
        cnt = C.c_uint64(int(count))
        self._check(self.L.pcp_serial_sum(self.h, p, n, s, acc.ctypes.data, C.byref(cnt)))
        return acc, cnt.value

    def cloud_stats(self, cloud, out=None):
        """-> 10 doubles: min xyz, max xyz, tree-order sums xyz, finite count (host array unless `out` is a device pointer)."""
        p, n, s, keep = _pts(cloud)
        ret = None
        if out is None:
            ret = np.zeros(10)
            out = ret
        self._check(self.L.pcp_cloud_stats(self.h, p, n, s, _ptr(out)))
        return ret

    def serial_sum_begin(self, cloud, predicted3=None):
        p, n, s, keep = _pts(cloud)
        h = C.c_void_p()
        pred = None if predicted3 is None else np.ascontiguousarray(predicted3, dtype=np.float64)
        self._check(self.L.pcp_serial_sum_begin(self.h, p, n, s, _ptr(pred), C.byref(h)))
        return h

    def serial_sum_end(self, handle, carry3):
        """carry3: host numpy array of 3 doubles (updated in place) or a device pointer."""
        self._check(self.L.pcp_serial_sum_end(handle, _ptr(carry3)))

    def partition_columns(self, cloud, axis, leaf, centre, min_b, bounds, out):
        p, n, s, keep = _pts(cloud)
        b = np.ascontiguousarray(bounds, dtype=np.int32)
        counts = np.zeros(len(b) - 1, dtype=np.uint64)
        self._check(self.L.pcp_partition_columns(self.h, p, n, s, int(axis), C.c_float(leaf), float(centre), int(min_b),
                                                 b.ctypes.data, len(b) - 1, _ptr(out), counts.ctypes.data))
        return counts

    def peek(self, dev_ptr, dtype, count=1):
        out = np.zeros(count, dtype=dtype)
        self._check(self.L.pcp_peek(self.h, _ptr(dev_ptr), out.nbytes, out.ctypes.data))
        return out

    def poke(self, dev_ptr, arr):
        """A small host array into device memory (pcp_poke: no place in the copy engine's queue)."""
        a = np.ascontiguousarray(arr)
        self._check(self.L.pcp_poke(self.h, _ptr(dev_ptr), a.nbytes, a.ctypes.data))

    def column_histogram(self, cloud, axis, leaf, centre, min_b, col0, cols_per_bin, hist):
        """hist: uint64 numpy array, or a (device pointer, nbins) tuple; counts are ADDED."""
        p, n, s, keep = _pts(cloud)
        if isinstance(hist, tuple):
            hp, nb = hist
        else:
            assert hist.dtype == np.uint64 and hist.flags.c_contiguous
            hp, nb = hist.ctypes.data, hist.shape[0]
        self._check(self.L.pcp_column_histogram(self.h, p, n, s, axis, C.c_float(leaf), float(centre), int(min_b),
                                                int(col0), int(cols_per_bin), int(nb), _ptr(hp)))

    def select_range(self, cloud, axis, lo, hi, out, cap=None):
        """cap: records `out` has room for, when that is less than the input's size (checked after the fact: the
        selections of this package are a small fraction of their input)."""
        p, n, s, keep = _pts(cloud)
        m = C.c_size_t(0)
        self._check(self.L.pcp_select_range(self.h, p, n, s, axis, float(lo), float(hi), _ptr(out), C.byref(m)))
        if cap is not None and m.value > cap:
            raise _lib.PcpError(-4, "select_range: %d records selected, room for %d" % (m.value, cap))
        return m.value

    def select_boundary(self, cloud, axis, kth_d2, lo, hi, cap, out_idx, out_xyz):
        """lo / hi: outer halo faces or None.  Returns the number selected, or minus that number when it exceeds
        cap (nothing usable was written: call again with more room)."""
        p, n, s, keep = _pts(cloud)
        c = C.c_size_t(0)
        st = self.L.pcp_select_boundary(self.h, p, n, s, axis, _ptr(kth_d2), int(lo is not None), float(lo or 0.0),
                                        int(hi is not None), float(hi or 0.0), int(cap), _ptr(out_idx), _ptr(out_xyz),
                                        C.byref(c))
        if st == -4 and c.value > cap:
            return -int(c.value)  # not enough room: the caller grows its buffers and calls again
        self._check(st)
        return c.value

    def select_columns(self, cloud, axis, leaf, centre, min_b, col_lo, col_hi, out):
        p, n, s, keep = _pts(cloud)
        m = C.c_size_t(0)
        self._check(self.L.pcp_select_columns(self.h, p, n, s, axis, C.c_float(leaf), float(centre), int(min_b),
                                              int(col_lo), int(col_hi), _ptr(out), C.byref(m)))
        return m.value

    def features_merge(self, nq, k, nlists, xyz_lists, d2_lists, point_id, props, eig=None, lps=None):
        self._check(self.L.pcp_features_merge(self.h, int(nq), int(k), int(nlists), _ptr(xyz_lists), _ptr(d2_lists),
                                              _ptr(point_id), _ptr(props), _ptr(eig), _ptr(lps)))

    def sort_pairs(self, keys, vals=None, bits=None):
        """In-place stable sort of (key, value) pairs by the low `bits` bits of the key (pcp_sort_pairs).
        keys: uint32 / uint64 numpy array (or device pointer tuple (ptr, n, key_bytes)); vals: uint32 or None."""
        if isinstance(keys, tuple):
            kp, n, kb = keys
        else:
            assert keys.dtype in (np.uint32, np.uint64) and keys.flags.c_contiguous
            kp, n, kb = keys.ctypes.data, keys.shape[0], keys.dtype.itemsize
        self._check(self.L.pcp_sort_pairs(self.h, _ptr(kp), _ptr(vals), int(n), int(kb), int(bits or 8 * kb)))

    def exclusive_sum(self, a, out=None):
        a = np.ascontiguousarray(a, dtype=np.uint32)
        out = np.empty_like(a) if out is None else out
        self._check(self.L.pcp_exclusive_sum_u32(self.h, a.ctypes.data, out.ctypes.data, a.shape[0]))
        return out

    def centroid(self, cloud):
        p, n, s, keep = _pts(cloud)
        c = np.zeros(3)
        cnt = C.c_uint64(0)
        self._check(self.L.pcp_centroid(self.h, p, n, s, c.ctypes.data, C.byref(cnt)))
        return c, cnt.value

    def minmax(self, cloud):
        p, n, s, keep = _pts(cloud)
        mn, mx = np.zeros(3), np.zeros(3)
        self._check(self.L.pcp_minmax(self.h, p, n, s, mn.ctypes.data, mx.ctypes.data))
        return mn, mx

    def plane_fit(self, pts, seg_offsets):
        p, n, s, keep = _pts(pts)
        off = np.ascontiguousarray(seg_offsets, dtype=np.int64)
        out = np.zeros(off.shape[0] - 1, dtype=PLAN_DTYPE)
        self._check(self.L.pcp_plane_fit(self.h, p, s, off.ctypes.data, off.shape[0] - 1, out.ctypes.data))
        return out


class MultiGpu:
    """pcp_mgpu: the remove_duplicate -> kNN -> plane-fit pipeline over several GPUs of this process (C ABI, no torch)."""

    def __init__(self, ndev, devices=None):
        self.L = _lib.load()
        h = C.c_void_p()
        dv = None if devices is None else np.ascontiguousarray(devices, dtype=np.int32)
        st = self.L.pcp_mgpu_create(int(ndev), _ptr(dv), C.byref(h))
        if st != 0:
            raise _lib.PcpError(st, (self.L.pcp_last_error(None) or b"").decode())
        self.h, self.ndev = h, int(ndev)

    def close(self):
        if getattr(self, "h", None):
            self.L.pcp_mgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def features_knn(self, cloud, leaf=0.02, k=20, halo=2.0):
        """cloud: host array of POINT_DTYPE or PACKED_POINT_DTYPE.  -> (voxelised cloud, props), slab after slab."""
        a = np.ascontiguousarray(cloud)
        n, s = a.shape[0], a.dtype.itemsize
        out = np.zeros(max(n, 1), dtype=a.dtype)
        props = np.zeros(max(n, 1), dtype=PROP_DTYPE)
        m = C.c_size_t(0)
        st = self.L.pcp_mgpu_features_knn(self.h, a.ctypes.data, n, s, C.c_float(leaf if leaf else 0.0), int(k), float(halo),
                                          out.ctypes.data, props.ctypes.data, C.byref(m))
        if st != 0:
            raise _lib.PcpError(st, (self.L.pcp_last_error(None) or b"").decode())
        return out[:m.value], props[:m.value]


class Index:
    """pcp_index: the search structure built by KdTreeFLANN.setInputCloud."""

    def __init__(self, ctx, cloud, cell_size=0.0):
        self.ctx = ctx
        p, n, s, keep = _pts(cloud)
        self.n_orig = n
        h = C.c_void_p()
        ctx._check(ctx.L.pcp_index_build(ctx.h, p, n, s, float(cell_size), C.byref(h)))
        self.h = h

    def free(self):
        if getattr(self, "h", None) and self.ctx.h:
            self.ctx.L.pcp_index_free(self.h)
        self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass

    def size(self):
        return int(self.ctx.L.pcp_index_size(self.h))

    def cell_size(self):
        return float(self.ctx.L.pcp_index_cell_size(self.h))

    def knn(self, queries, k, want_d2=True):
        p, nq, s, keep = _pts(queries)
        idx = np.empty((nq, k), dtype=np.int32)
        d2 = np.empty((nq, k), dtype=np.float64) if want_d2 else None
        found = np.empty(nq, dtype=np.int32)
        self.ctx._check(self.ctx.L.pcp_knn(self.h, p, nq, s, k, idx.ctypes.data, _ptr(d2), found.ctypes.data))
        return idx, d2, found

    def radius(self, queries, r, max_nn=0):
        p, nq, s, keep = _pts(queries)
        off = np.zeros(nq + 1, dtype=np.int64)
        self.ctx._check(self.ctx.L.pcp_radius_count(self.h, p, nq, s, float(r), max_nn, off.ctypes.data))
        idx = np.empty(off[-1], dtype=np.int32)
        d2 = np.empty(off[-1], dtype=np.float64)
        self.ctx._check(self.ctx.L.pcp_radius_fill(self.h, p, nq, s, float(r), max_nn, off.ctypes.data,
                                                   idx.ctypes.data, d2.ctypes.data))
        return off, idx, d2

    def knn_device(self, q_ptr, nq, qstride, k, idx_ptr, d2_ptr):
        """pcp_knn with caller-owned (device) buffers."""
        self.ctx._check(self.ctx.L.pcp_knn(self.h, _ptr(q_ptr), int(nq), int(qstride), int(k), _ptr(idx_ptr),
                                           _ptr(d2_ptr), None))

    def features_knn(self, k, props=None, eig=None, lps=None, knn_idx=None, want_eig=True, want_lps=True,
                     want_idx=False, query_limit=-1, kth_d2=None):
        """Host outputs are allocated unless device pointers are passed in."""
        n = self.n_orig
        ret = {}
        if props is None:
            props = ret["props"] = np.zeros(n, dtype=PROP_DTYPE)
        if eig is None and want_eig:
            eig = ret["eig"] = np.zeros((n, 3))
        if lps is None and want_lps:
            lps = ret["lps"] = np.zeros((n, 3), dtype=np.float32)
        if knn_idx is None and want_idx:
            knn_idx = ret["knn_idx"] = np.empty((n, k), dtype=np.int32)
        self.ctx._check(self.ctx.L.pcp_features_knn_ex(self.h, k, int(query_limit), _ptr(props), _ptr(eig), _ptr(lps),
                                                       _ptr(knn_idx), _ptr(kth_d2)))
        return ret

    def features_radius(self, r, max_nn=0, props=None, eig=None, lps=None, counts=None, want_extra=True):
        n = self.n_orig
        ret = {}
        if props is None:
            props = ret["props"] = np.zeros(n, dtype=PROP_DTYPE)
        if want_extra:
            if eig is None:
                eig = ret["eig"] = np.zeros((n, 3))
            if lps is None:
                lps = ret["lps"] = np.zeros((n, 3), dtype=np.float32)
            if counts is None:
                counts = ret["counts"] = np.zeros(n, dtype=np.int32)
        self.ctx._check(self.ctx.L.pcp_features_radius(self.h, float(r), max_nn, _ptr(props), _ptr(eig), _ptr(lps),
                                                       _ptr(counts)))
        return ret

    def features_rpca(self, k=20, pr=0.99, epi=0.5, seed=1, props=None):
        """calculate_plan_parameter_rpca (calculate_feature.cpp:208-368) with a counter-based hypothesis
        generator (include/pcp_cuda.h: pcp_features_rpca).  Returns the props records (host array unless a
        device array is passed)."""
        n = self.n_orig
        ret = props
        if props is None:
            props = ret = np.zeros(n, dtype=PROP_DTYPE)
        self.ctx._check(self.ctx.L.pcp_features_rpca(self.h, int(k), float(pr), float(epi), int(seed) & (2**64 - 1),
                                                     _ptr(props)))
        return ret

    def region_growing(self, props, k=50, distanceT=0.5, cosfaT=0.94, curv_t=0.005, min_points=5, want_stats=False):
        """TreeExtration::region_growning (extraction_tree.cpp:170-272) over this index's cloud (include/pcp_cuda.h:
        pcp_region_growing).  `props` (host records) gets its SegmentID rewritten.  Returns dict(offsets, points,
        planes[, stats]): the kept regions in the order found, members in PlanSegment::PointID order."""
        n = self.n_orig
        assert props.dtype == PROP_DTYPE and len(props) == n and props.flags.c_contiguous
        cap = n // (max(int(min_points), 0) + 1) + 1
        offsets = np.zeros(cap + 1, dtype=np.int64)
        points = np.zeros(max(n, 1), dtype=np.int32)
        planes = np.zeros((cap, 4), dtype=np.float32)
        nseg = C.c_size_t(0)
        stats = np.zeros(6, dtype=np.uint64)
        self.ctx._check(self.ctx.L.pcp_region_growing(self.h, _ptr(props), int(k), float(distanceT), float(cosfaT),
                                                      float(curv_t), int(min_points), _ptr(offsets), cap, _ptr(points),
                                                      _ptr(planes), C.byref(nseg), _ptr(stats)))
        m = nseg.value
        ret = {"offsets": offsets[:m + 1].copy(), "points": points[:offsets[m]].copy(), "planes": planes[:m].copy()}
        if want_stats:
            ret["stats"] = dict(zip(("seeds", "solo_passes", "wide_passes", "scanned", "members", "candidates"),
                                    (int(v) for v in stats)))
        return ret


# ---- reference-shaped classes ------------------------------------------------------------------------
_default_ctx = None


def default_context():
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context(0)
    return _default_ctx


class PointCloud:
    """point_cloud.h:54-535: points + width/height/is_dense."""

    def __init__(self, points=None):
        self.points = np.zeros(0, dtype=POINT_DTYPE) if points is None else points
        self.height = 1
        self.is_dense = True

    @property
    def width(self):
        return len(self.points) // max(int(self.height), 1)

    def size(self):
        return len(self.points)

    def push_back(self, x, y, z, rgba=0, stamp_id=0):
        pts = np.zeros(len(self.points) + 1, dtype=POINT_DTYPE)  # (np.concatenate would drop the record's padding)
        for f in POINT_DTYPE.names:
            pts[f][:-1] = self.points[f]
        last = pts[-1:]
        last["x"], last["y"], last["z"], last["w"], last["rgba"], last["stamp_id"] = x, y, z, 1.0, rgba, stamp_id
        self.points = pts


class io:
    """io::loadPCDFile / savePCDFile* (pcd_helper.h:1343-1497) and lzfCompress / lzfDecompress (lzf.h) over the
    library's host-side PCD code (csrc/pcd_io.cu).  Return conventions of the reference: 0 = ok, -1 = error (the
    message is printed, as PCDReader does)."""

    @staticmethod
    def _fail(L):
        print(L.pcp_last_error(None).decode())
        return -1

    @staticmethod
    def readHeader(file_name):
        """-> dict (fields, sizes, types, counts, width, height, points, data_type, data_offset, viewpoint) or None."""
        L = _lib.load()
        h = _lib.PcdHeader()
        if L.pcp_pcd_read_header(str(file_name).encode(), C.byref(h)) != 0:
            io._fail(L)
            return None
        n = h.nfields
        return {"fields": [h.names[i].value.decode() for i in range(n)], "sizes": list(h.size[:n]),
                "types": [h.type[i:i + 1].decode() for i in range(n)], "counts": list(h.count[:n]), "width": h.width,
                "height": h.height, "points": h.points, "data_type": ("ascii", "binary", "binary_compressed")[h.data_type],
                "data_offset": h.data_offset, "viewpoint": list(h.viewpoint)}

    @staticmethod
    def loadPCDFiles(file_names, ctx=None, out=None, threads=0):
        """The pipelined loader (include/pcp_cuda.h: pcp_pcd_load_many): several files into one cloud; payloads are
        decoded by host threads, uploaded as stored and assembled into records on the device.  `out`: a DeviceArray
        to keep the cloud on the GPU (returns (n, offsets, is_dense)), else a host POINT_DTYPE array is returned with
        (offsets, is_dense).  Returns None after printing the message on error, like the single-file reader."""
        ctx = ctx or default_context()
        names = [str(f).encode() for f in file_names]
        total = 0
        for f in file_names:
            h = io.readHeader(f)
            if h is None:
                return None
            total += int(h["points"])
        arr = (C.c_char_p * len(names))(*names)
        offsets = np.zeros(len(names) + 1, dtype=np.uintp)
        dense = C.c_int(1)
        host = None
        if out is None:
            host = np.zeros(total, dtype=POINT_DTYPE)
            ptr, cap = host.ctypes.data, total
        else:
            ptr, cap = _ptr(out), out.nbytes // 48
        st = ctx.L.pcp_pcd_load_many(ctx.h, arr, len(names), 0, ptr, cap, 48, int(threads), offsets.ctypes.data,
                                     C.byref(dense))
        if st != 0:
            print(ctx.L.pcp_last_error(ctx.h).decode())
            return None
        offsets = offsets.astype(np.int64)
        if out is None:
            return host, offsets, bool(dense.value)
        return int(offsets[-1]), offsets, bool(dense.value)

    @staticmethod
    def loadPCDFile(file_name, cloud):
        L = _lib.load()
        h = _lib.PcdHeader()
        if L.pcp_pcd_read_header(str(file_name).encode(), C.byref(h)) != 0:
            return io._fail(L)
        pts = np.zeros(int(h.points), dtype=POINT_DTYPE)
        n, dense = C.c_size_t(0), C.c_int(1)
        if L.pcp_pcd_load(str(file_name).encode(), 0, pts.ctypes.data, len(pts), 48, C.byref(n), C.byref(dense)) != 0:
            return io._fail(L)
        cloud.points, cloud.height, cloud.is_dense = pts, int(h.height), bool(dense.value)
        return 0

    @staticmethod
    def _save(file_name, cloud, mode, precision=8):
        L = _lib.load()
        pts = np.ascontiguousarray(cloud.points if isinstance(cloud, PointCloud) else cloud)
        height = int(cloud.height) if isinstance(cloud, PointCloud) else 1
        width = len(pts) // max(height, 1)
        st = L.pcp_pcd_save(str(file_name).encode(), 0, pts.ctypes.data, len(pts), pts.dtype.itemsize, width, height, mode,
                            precision)
        return 0 if st == 0 else io._fail(L)

    @staticmethod
    def savePCDFile(file_name, cloud, binary_mode=True):
        return io._save(file_name, cloud, 1 if binary_mode else 0)

    @staticmethod
    def savePCDFileASCII(file_name, cloud, precision=8):
        return io._save(file_name, cloud, 0, precision)

    @staticmethod
    def savePCDFileBinary(file_name, cloud):
        return io._save(file_name, cloud, 1)

    @staticmethod
    def savePCDFileBinaryCompressed(file_name, cloud):
        return io._save(file_name, cloud, 2)

    @staticmethod
    def lzfCompress(data):
        """-> the LZF stream (bytes), b"" when the data does not shrink into 1.5x its size (lzf.cpp returns 0)."""
        L = _lib.load()
        src = bytes(data)
        out = C.create_string_buffer(int(len(src) * 1.5) + 16)
        n = L.pcp_lzf_compress(src, len(src), out, len(out))
        return out.raw[:n]

    @staticmethod
    def lzfDecompress(data, out_len):
        L = _lib.load()
        src = bytes(data)
        out = C.create_string_buffer(max(int(out_len), 1))
        n = L.pcp_lzf_decompress(src, len(src), out, int(out_len))
        return out.raw[:n]


class KdTreeFLANN:
    def __init__(self, ctx=None, sorted=True):
        self.ctx = ctx or default_context()
        self.index = None

    def setInputCloud(self, cloud):
        pts = cloud.points if isinstance(cloud, PointCloud) else cloud
        self.index = Index(self.ctx, pts)

    def nearestKSearch(self, point, k):
        """-> (count, k_indices, k_sqr_distances); k is clamped to the number of valid points (kd_tree.h:820)."""
        q = np.array([[point["x"], point["y"], point["z"]]]) if hasattr(point, "dtype") and point.dtype.names \
            else np.asarray(point, dtype=np.float64).reshape(1, -1)
        kk = min(k, self.index.size())
        if kk <= 0:
            return 0, np.zeros(0, np.int32), np.zeros(0)
        idx, d2, found = self.index.knn(q, kk)
        return int(found[0]), idx[0, :found[0]].copy(), d2[0, :found[0]].copy()

    def radiusSearch(self, point, radius, max_nn=0):
        q = np.array([[point["x"], point["y"], point["z"]]]) if hasattr(point, "dtype") and point.dtype.names \
            else np.asarray(point, dtype=np.float64).reshape(1, -1)
        off, idx, d2 = self.index.radius(q, radius, max_nn)
        return int(off[1]), idx, d2


class VoxelGrid:
    """voxel_grid.h:455-810 (setters / accessors) over pcp_voxel_grid.  The leaf layout (voxel_grid.h:964-983,
    1031-1032) is kept sparse: cell index -> output index; getLeafLayout() expands it."""

    def __init__(self, ctx=None):
        self.ctx = ctx or default_context()
        self.input = None
        self.leaf = None
        self.downsample_all_data = True  # voxel_grid.h:499
        self.save_leaf_layout = False
        self.min_b = self.div_b = None
        self._layout = None

    def setInputCloud(self, cloud):
        self.input = cloud

    def setLeafSize(self, lx, ly, lz):
        self.leaf = (float(lx), float(ly), float(lz))

    def getLeafSize(self):
        return self.leaf

    def setDownsampleAllData(self, flag):
        self.downsample_all_data = bool(flag)

    def getDownsampleAllData(self):
        return self.downsample_all_data

    def setSaveLeafLayout(self, flag):
        self.save_leaf_layout = bool(flag)

    def getSaveLeafLayout(self):
        return self.save_leaf_layout

    def getMinBoxCoordinates(self):
        return self.min_b

    def getMaxBoxCoordinates(self):
        return None if self.min_b is None else self.min_b + self.div_b - 1

    def getNrDivisions(self):
        return self.div_b

    def getDivisionMultiplier(self):
        d = self.div_b
        return None if d is None else np.array([1, d[0], d[0] * d[1]], dtype=np.int64)

    def getGridCoordinates(self, x, y, z):
        """(int)(x * inverse_leaf) per axis, truncating (voxel_grid.h:672-678)."""
        inv = [1.0 / l for l in self.leaf]
        return np.array([int(x * inv[0]), int(y * inv[1]), int(z * inv[2])], dtype=np.int32)

    def getCentroidIndexAt(self, ijk):
        if self._layout is None:
            return -1
        rel = np.asarray(ijk, dtype=np.int64) - self.min_b
        idx = int(rel.dot(self.getDivisionMultiplier()))
        if idx < 0 or idx >= int(np.prod(self.div_b.astype(np.int64))):
            return -1
        return self._layout.get(idx, -1)

    @staticmethod
    def _xyz(p):
        if hasattr(p, "dtype") and p.dtype.names:
            return float(p["x"]), float(p["y"]), float(p["z"])
        if isinstance(p, np.void):  # a record of an unnamed dtype: xyz doubles come first
            return tuple(np.frombuffer(p.tobytes()[:24], dtype=np.float64))
        return tuple(float(v) for v in p[:3])

    def getCentroidIndex(self, p):
        return self.getCentroidIndexAt(self.getGridCoordinates(*self._xyz(p)))

    def getNeighborCentroidIndices(self, reference_point, relative_coordinates):
        """relative_coordinates: 3 x m integer offsets; -1 for cells outside the grid (voxel_grid.h:627-646)."""
        ijk = self.getGridCoordinates(*self._xyz(reference_point)).astype(np.int64)
        rel = np.asarray(relative_coordinates, dtype=np.int64).reshape(3, -1)
        lo, hi = self.min_b - ijk, self.getMaxBoxCoordinates() - ijk
        out = []
        for c in range(rel.shape[1]):
            d = rel[:, c]
            out.append(self.getCentroidIndexAt(ijk + d) if (lo <= d).all() and (hi >= d).all() else -1)
        return out

    def getLeafLayout(self):
        if self._layout is None:
            return np.zeros(0, dtype=np.int32)
        full = np.full(int(np.prod(self.div_b.astype(np.int64))), -1, dtype=np.int32)
        full[np.fromiter(self._layout.keys(), dtype=np.int64)] = np.fromiter(self._layout.values(), dtype=np.int32)
        return full

    def filter(self):
        out = PointCloud()
        if self.input is None:  # voxel_grid.h:815-821
            print("[applyFilter] No input dataset given")
            return out
        pts = self.input.points if isinstance(self.input, PointCloud) else self.input
        r = self.ctx.voxel_grid(pts, self.leaf, self.downsample_all_data, want_cells=self.save_leaf_layout)
        self.min_b, self.div_b = r["min_b"], r["div_b"]
        self._layout = None
        if self.save_leaf_layout:
            cell = r["point_cell"].astype(np.int64)
            cell = cell[cell[:, 0] >= 0]
            idx = np.unique(cell.dot(self.getDivisionMultiplier()))  # outputs are in ascending cell index
            self._layout = {int(v): i for i, v in enumerate(idx)}
        out.points = r["points"].copy()
        return out


class CalculateFeature:
    def __init__(self, ctx=None):
        self.ctx = ctx or default_context()

    def calculate_plan_parameter_h_points(self, cloud):
        pts = cloud.points if isinstance(cloud, PointCloud) else cloud
        n = len(pts)
        return self.ctx.plane_fit(pts, [0, n])[0]

    def calculate_plan_parameter(self, cloud, radius=None, k=20):
        """calculate_feature.h:15 (declared, never defined in the reference): kNN(k) plane fit per point,
        or radius neighbourhoods when `radius` is given."""
        pts = cloud.points if isinstance(cloud, PointCloud) else cloud
        ix = Index(self.ctx, pts)
        try:
            if radius is None:
                return ix.features_knn(k, want_eig=False, want_lps=False)["props"]
            return ix.features_radius(radius, want_extra=False)["props"]
        finally:
            ix.free()

    def calculate_plan_parameter_rpca(self, cloud, radius, pr=0.99, epi=0.5, seed=1):
        """calculate_feature.cpp:208-368: the robust fit (3-point hypotheses, h-subset refit, MAD inliers) on the
        20 nearest neighbours; like the reference it ignores `radius` (:233 uses k = 20).  The reference's rand()
        hypotheses are replaced by a counter-based generator keyed by `seed`, so the result is reproducible and
        is checked bit for bit by the parity tests."""
        pts = cloud.points if isinstance(cloud, PointCloud) else cloud
        ix = Index(self.ctx, pts)
        try:
            return ix.features_rpca(20, pr, epi, seed)
        finally:
            ix.free()

    @staticmethod
    def compute_iteration_number(pr, epi, ctx=None):
        """calculate_feature.cpp:28-33 for 3-point hypotheses."""
        return (ctx or default_context()).L.pcp_rpca_iterations(float(pr), float(epi))


class TreeExtration:
    """extraction_tree.h (the reference spells it this way): region growing over the plane-fit records."""

    def __init__(self, ctx=None):
        self.ctx = ctx or default_context()

    def region_growning(self, PointProperty, Cloud, distanceT, radius, cosfaT):
        """extraction_tree.cpp:170-272.  `radius` is not read by the reference either (the search is k = 50, :222).
        Returns the list of PlanSegment dicts (SegmentID, PointID array, normal_x/y/z, Distance) and rewrites
        PointProperty[...]['SegmentID'] in place."""
        pts = Cloud.points if isinstance(Cloud, PointCloud) else Cloud
        ix = Index(self.ctx, pts)
        try:
            r = ix.region_growing(PointProperty, 50, distanceT, cosfaT)
        finally:
            ix.free()
        segs = []
        for s in range(len(r["planes"])):
            pl = r["planes"][s]
            segs.append({"SegmentID": s, "PointID": r["points"][r["offsets"][s]:r["offsets"][s + 1]],
                         "normal_x": pl[0], "normal_y": pl[1], "normal_z": pl[2], "Distance": pl[3]})
        return segs


def point_segment(src_cloud, ctx=None, seed=1):
    """static.cpp:8-21: robust plane fit of every point, then region growing with the reference's constants."""
    ctx = ctx or default_context()
    props = CalculateFeature(ctx).calculate_plan_parameter_rpca(src_cloud, 0.15, 0.99, 0.5, seed)
    return TreeExtration(ctx).region_growning(props, src_cloud, 0.5, 0.2, 0.940), props


class PointCloudHelper:
    @staticmethod
    def remove_duplicate(cloud, voxel_grid_size, min_num=None, ctx=None):
        ctx = ctx or default_context()
        pts = cloud.points if isinstance(cloud, PointCloud) else cloud
        out = ctx.remove_duplicate(pts, np.float32(voxel_grid_size))
        if min_num is not None and len(out) < min_num:  # point_cloud_helper.cpp:65-73
            out = pts
        if isinstance(cloud, PointCloud):
            cloud.points = out
            return cloud
        return out

    @staticmethod
    def getMinMax3D(cloud, ctx=None):
        ctx = ctx or default_context()
        return ctx.minmax(cloud.points if isinstance(cloud, PointCloud) else cloud)

    @staticmethod
    def compute3DCentroid(cloud, ctx=None):
        ctx = ctx or default_context()
        return ctx.centroid(cloud.points if isinstance(cloud, PointCloud) else cloud)
