"""ctypes binding of libpcp_cuda.so (include/pcp_cuda.h).  There is no CPU fallback: if the shared library
is missing or no CUDA device is usable, every entry point raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpcp_cuda.so")

# every symbol include/pcp_cuda.h declares (checked by tests/test_abi.py against the header text)
SYMBOLS = [
    "pcp_create", "pcp_destroy", "pcp_last_error", "pcp_set_stream", "pcp_synchronize", "pcp_device_count",
    "pcp_version", "pcp_trim", "pcp_profile_enable", "pcp_profile_reset", "pcp_profile_get", "pcp_launch_count",
    "pcp_dev_alloc", "pcp_dev_free", "pcp_host_alloc_pinned", "pcp_host_free_pinned", "pcp_memcpy", "pcp_memcpy_async",
    "pcp_copy_async", "pcp_copy_fence", "pcp_copy_wait",
    "pcp_index_build", "pcp_index_free", "pcp_index_size", "pcp_index_cell_size", "pcp_knn", "pcp_radius_count",
    "pcp_radius_fill", "pcp_features_knn", "pcp_features_radius", "pcp_plane_fit", "pcp_voxel_grid",
    "pcp_remove_duplicate", "pcp_features_knn_ex", "pcp_remove_duplicate_centred", "pcp_serial_sum",
    "pcp_column_histogram", "pcp_select_range", "pcp_select_columns", "pcp_select_boundary", "pcp_features_merge", "pcp_centroid", "pcp_minmax", "pcp_synth_cloud",
    "pcp_features_rpca", "pcp_rpca_iterations", "pcp_region_growing",
    "pcp_remove_duplicate_to", "pcp_mgpu_create", "pcp_mgpu_destroy", "pcp_mgpu_device_count", "pcp_mgpu_ctx", "pcp_mgpu_last_error", "pcp_mgpu_features_knn",
    "pcp_serial_sum_begin", "pcp_serial_sum_end", "pcp_cloud_stats", "pcp_partition_columns", "pcp_peek", "pcp_poke",
    "pcp_props_pack", "pcp_props_unpack", "pcp_cloud_pack", "pcp_cloud_unpack", "pcp_sort_pairs", "pcp_exclusive_sum_u32", "pcp_exclusive_sum_u32_noalloc",
    "pcp_pcd_read_header", "pcp_pcd_load", "pcp_pcd_load_many", "pcp_pcd_save", "pcp_lzf_compress", "pcp_lzf_decompress",
]


class StageTime(C.Structure):
    _fields_ = [("name", C.c_char * 48), ("ms", C.c_double), ("launches", C.c_int64), ("bytes", C.c_double)]


class PcpError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("pcp status %d: %s" % (status, msg))
        self.status = status


_lib = None


class PcdHeader(C.Structure):
    """pcp_pcd_header (include/pcp_cuda.h)."""
    _fields_ = [("points", C.c_uint64), ("data_offset", C.c_uint64), ("width", C.c_uint32), ("height", C.c_uint32),
                ("data_type", C.c_int32), ("nfields", C.c_int32), ("names", (C.c_char * 32) * 16),
                ("size", C.c_int32 * 16), ("count", C.c_int32 * 16), ("type", C.c_char * 16), ("viewpoint", C.c_double * 7)]


def load():
    """Load the shared library (no CUDA call is made here)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PcpError(-2, "libpcp_cuda.so is not built (run `python -m pointcloudprocess_b200.build` or "
                           "__graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, sz, i32, i64, u32, u64, dbl = C.c_void_p, C.c_size_t, C.c_int, C.c_int64, C.c_uint32, C.c_uint64, C.c_double
    pp = C.POINTER(C.c_void_p)
    sig = {
        "pcp_create": (i32, [pp, i32]),
        "pcp_destroy": (None, [vp]),
        "pcp_last_error": (C.c_char_p, [vp]),
        "pcp_set_stream": (i32, [vp, vp]),
        "pcp_synchronize": (i32, [vp]),
        "pcp_device_count": (i32, []),
        "pcp_version": (C.c_char_p, []),
        "pcp_trim": (i32, [vp]),
        "pcp_profile_enable": (i32, [vp, i32]),
        "pcp_profile_reset": (i32, [vp]),
        "pcp_profile_get": (i32, [vp, C.POINTER(StageTime), i32, C.POINTER(i32)]),
        "pcp_launch_count": (i64, [vp]),
        "pcp_dev_alloc": (i32, [vp, sz, pp]),
        "pcp_dev_free": (i32, [vp, vp]),
        "pcp_host_alloc_pinned": (i32, [vp, sz, pp]),
        "pcp_host_free_pinned": (i32, [vp, vp]),
        "pcp_memcpy": (i32, [vp, vp, vp, sz]),
        "pcp_memcpy_async": (i32, [vp, vp, vp, sz]),
        "pcp_copy_async": (i32, [vp, i32, vp, vp, sz]),
        "pcp_copy_fence": (i32, [vp, i32]),
        "pcp_copy_wait": (i32, [vp, i32]),
        "pcp_index_build": (i32, [vp, vp, sz, sz, dbl, pp]),
        "pcp_index_free": (None, [vp]),
        "pcp_index_size": (i64, [vp]),
        "pcp_index_cell_size": (dbl, [vp]),
        "pcp_knn": (i32, [vp, vp, sz, sz, i32, vp, vp, vp]),
        "pcp_radius_count": (i32, [vp, vp, sz, sz, dbl, u32, vp]),
        "pcp_radius_fill": (i32, [vp, vp, sz, sz, dbl, u32, vp, vp, vp]),
        "pcp_features_knn": (i32, [vp, i32, vp, vp, vp, vp]),
        "pcp_features_radius": (i32, [vp, dbl, u32, vp, vp, vp, vp]),
        "pcp_features_rpca": (i32, [vp, i32, C.c_float, C.c_float, u64, vp]),
        "pcp_rpca_iterations": (i32, [C.c_float, C.c_float]),
        "pcp_region_growing": (i32, [vp, vp, i32, dbl, dbl, dbl, i32, vp, sz, vp, vp, C.POINTER(sz), vp]),
        "pcp_plane_fit": (i32, [vp, vp, sz, vp, sz, vp]),
        "pcp_voxel_grid": (i32, [vp, vp, sz, sz, vp, i32, vp, C.POINTER(sz), vp, vp, vp, vp]),
        "pcp_remove_duplicate": (i32, [vp, vp, C.POINTER(sz), sz, C.c_float]),
        "pcp_features_knn_ex": (i32, [vp, i32, i64, vp, vp, vp, vp, vp]),
        "pcp_remove_duplicate_centred": (i32, [vp, vp, C.POINTER(sz), sz, C.c_float, vp, vp, vp]),
        "pcp_serial_sum": (i32, [vp, vp, sz, sz, vp, C.POINTER(u64)]),
        "pcp_column_histogram": (i32, [vp, vp, sz, sz, i32, C.c_float, dbl, i32, i32, i32, i32, vp]),
        "pcp_select_range": (i32, [vp, vp, sz, sz, i32, dbl, dbl, vp, C.POINTER(sz)]),
        "pcp_select_columns": (i32, [vp, vp, sz, sz, i32, C.c_float, dbl, i32, i32, i32, vp, C.POINTER(sz)]),
        "pcp_select_boundary": (i32, [vp, vp, sz, sz, i32, vp, i32, dbl, i32, dbl, sz, vp, vp, C.POINTER(sz)]),
        "pcp_features_merge": (i32, [vp, sz, i32, i32, vp, vp, vp, vp, vp, vp]),
        "pcp_pcd_read_header": (i32, [C.c_char_p, C.POINTER(PcdHeader)]),
        "pcp_pcd_load": (i32, [C.c_char_p, i32, vp, sz, sz, C.POINTER(sz), C.POINTER(i32)]),
        "pcp_pcd_load_many": (i32, [vp, C.POINTER(C.c_char_p), sz, i32, vp, sz, sz, i32, vp, C.POINTER(i32)]),
        "pcp_pcd_save": (i32, [C.c_char_p, i32, vp, sz, sz, u32, u32, i32, i32]),
        "pcp_lzf_compress": (u32, [vp, u32, vp, u32]),
        "pcp_lzf_decompress": (u32, [vp, u32, vp, u32]),
        "pcp_centroid": (i32, [vp, vp, sz, sz, vp, C.POINTER(u64)]),
        "pcp_minmax": (i32, [vp, vp, sz, sz, vp, vp]),
        "pcp_synth_cloud": (i32, [vp, i32, u64, u64, u64, u64, i32, vp]),
        "pcp_sort_pairs": (i32, [vp, vp, vp, sz, i32, i32]),
        "pcp_remove_duplicate_to": (i32, [vp, vp, sz, sz, C.c_float, vp, vp, vp, vp, C.POINTER(sz)]),
        "pcp_props_pack": (i32, [vp, vp, sz, vp]),
        "pcp_serial_sum_begin": (i32, [vp, vp, sz, sz, vp, pp]),
        "pcp_serial_sum_end": (i32, [vp, vp]),
        "pcp_cloud_stats": (i32, [vp, vp, sz, sz, vp]),
        "pcp_partition_columns": (i32, [vp, vp, sz, sz, i32, C.c_float, dbl, i32, vp, i32, vp, vp]),
        "pcp_peek": (i32, [vp, vp, sz, vp]),
        "pcp_poke": (i32, [vp, vp, sz, vp]),
        "pcp_mgpu_create": (i32, [i32, vp, pp]),
        "pcp_mgpu_destroy": (None, [vp]),
        "pcp_mgpu_device_count": (i32, [vp]),
        "pcp_mgpu_ctx": (vp, [vp, i32]),
        "pcp_mgpu_last_error": (C.c_char_p, [vp]),
        "pcp_mgpu_features_knn": (i32, [vp, vp, sz, sz, C.c_float, i32, dbl, vp, vp, C.POINTER(sz)]),
        "pcp_cloud_pack": (i32, [vp, vp, sz, vp]),
        "pcp_cloud_unpack": (i32, [vp, vp, sz, vp]),
        "pcp_props_unpack": (i32, [vp, sz, vp, i32]),
        "pcp_exclusive_sum_u32": (i32, [vp, vp, vp, sz]),
        "pcp_exclusive_sum_u32_noalloc": (i32, [vp, vp, vp, sz]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)  # AttributeError here = the .so does not export what the header declares
        f.restype = res
        f.argtypes = args
    _lib = L
    return L


def check(ctx_handle, status):
    if status != 0:
        msg = load().pcp_last_error(ctx_handle)
        raise PcpError(status, msg.decode("utf-8", "replace") if msg else "")
