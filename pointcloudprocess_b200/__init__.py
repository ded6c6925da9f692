"""pointcloudprocess_b200 -- B200-native kNN / radius search, eigen-features and voxel grid for
RioWong/PointCloudProcess's hot path.  Everything computes in libpcp_cuda.so (hand-written sm_100a
kernels behind include/pcp_cuda.h); this package is the thin host-side mirror of the reference interface."""
from .api import (POINT_DTYPE, PROP_DTYPE, PLAN_DTYPE, PACKED_POINT_DTYPE, PACKED_PROP_DTYPE, SYNTH_KINDS, CalculateFeature, Context, DeviceArray, Index, MultiGpu,
                  KdTreeFLANN, PointCloud, PointCloudHelper, TreeExtration, VoxelGrid, default_context, io, make_cloud,
                  point_segment)
from ._lib import PcpError, load as load_library

__all__ = ["POINT_DTYPE", "PROP_DTYPE", "PLAN_DTYPE", "PACKED_POINT_DTYPE", "PACKED_PROP_DTYPE", "SYNTH_KINDS", "CalculateFeature", "Context", "DeviceArray", "io",
           "Index", "MultiGpu", "KdTreeFLANN", "PointCloud", "PointCloudHelper", "TreeExtration", "VoxelGrid", "default_context",
           "make_cloud", "point_segment",
           "PcpError", "load_library"]
