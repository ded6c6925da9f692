/*
 * pcp_cuda.h -- C ABI of libpcp_cuda.so: the B200 (sm_100a) replacement for the data-parallel hot path of
 * RioWong/PointCloudProcess.  Plain pointers and sizes only; every entry point names the reference
 * interface it replaces (file:line relative to the reference checkout).
 *
 * Conventions
 *   - Every function returns 0 on success, a negative pcp_status on failure; text via pcp_last_error().
 *   - Points are (base pointer, count, byte stride); the first 24 bytes of a record are x,y,z doubles.
 *     PointXYZRGBA (point_type.h:9-82) has stride 48 with rgba at byte 32 and stamp_id at byte 36,
 *     PointC (point_type.h:300-323) stride 64, a packed double[3] array stride 24.
 *     Stride 32 is the PACKED PointXYZRGBA: x, y, z doubles, rgba at byte 24, stamp_id at byte 28 -- exactly one
 *     record of the binary PCD payload the reference writes for its clouds (FIELDS x y z rgba stamp_id, SIZE 8 8 8
 *     4 4; pcd_helper.cpp:1074-1089), so a file's payload can be uploaded as stored (32 instead of 48 bytes per
 *     point); the voxel grid / remove_duplicate entry points read and write that layout when stride == 32.
 *   - Every data pointer may be a HOST pointer (pageable or pinned) or a DEVICE pointer on the context's
 *     GPU; the library finds out with cudaPointerGetAttributes.  Host buffers are staged through the
 *     context's own device arena; device buffers are used in place.  All pointers of one call that
 *     describe outputs may be mixed freely.
 *   - The caller owns every buffer it passes; the library owns its device arena (freed by pcp_destroy).
 *   - There is no CPU fallback: without a usable CUDA device pcp_create fails and nothing else works.
 *   - A pcp_ctx is not thread-safe; a built pcp_index is immutable and may be queried from several host
 *     threads (calls are serialised on the context's stream by an internal mutex).
 */
#ifndef PCP_CUDA_H
#define PCP_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PCP_API __attribute__((visibility("default")))
#else
#define PCP_API
#endif

typedef struct pcp_ctx pcp_ctx;
typedef struct pcp_index pcp_index;

typedef enum pcp_status {
    PCP_OK = 0,
    PCP_ERR_INVALID = -1, /* bad argument */
    PCP_ERR_CUDA = -2,    /* CUDA runtime error, see pcp_last_error */
    PCP_ERR_NOMEM = -3,
    PCP_ERR_RANGE = -4,   /* value not representable (e.g. > 2^31-1 points, voxel grid wider than 2^21) */
    PCP_ERR_NCCL = -5,    /* reserved: the library itself uses peer copies, not NCCL (pcp_mgpu_*) */
    PCP_ERR_IO = -6       /* file could not be opened / read / written */
} pcp_status;

/* LAS_POINT_PROPERTY, data_struct.h:161-173.  48 bytes. */
typedef struct pcp_point_property {
    float normal_x, normal_y, normal_z;
    double Distance;
    double curvature;
    int32_t PointID;
    int32_t SegmentID;
    float dis_from_point_plane;
} pcp_point_property;

/* The same record without its dead weight, for the trip over PCIe (24 bytes, lossless): Distance and curvature are
 * PlanSegment's FLOAT members widened to double (calculate_feature.cpp:344-345), SegmentID and dis_from_point_plane
 * are not set by calculate_plan_parameter*.  pcp_props_pack / pcp_props_unpack convert. */
typedef struct pcp_point_property_packed {
    float normal_x, normal_y, normal_z;
    float Distance;
    float curvature;
    int32_t PointID;
} pcp_point_property_packed;

/* PlanSegment (data_struct.h:188-198) without its std::vector member, + the three eigenvalues. */
typedef struct pcp_plan_segment {
    float normal_x, normal_y, normal_z;
    float min_value; /* lambda3 */
    float curvature;
    float Distance;
    double eig[3]; /* lambda1 >= lambda2 >= lambda3 of the un-normalised scatter X*X^T */
} pcp_plan_segment;

/* ---- context ------------------------------------------------------------------------------------ */
PCP_API int pcp_create(pcp_ctx** out, int cuda_device);
PCP_API void pcp_destroy(pcp_ctx* ctx);
PCP_API const char* pcp_last_error(const pcp_ctx* ctx); /* ctx may be NULL: error of a failed pcp_create */
/* Run on a caller-owned stream (a cudaStream_t passed as void*); NULL restores the context's own (non-blocking)
 * stream.  To share the default stream with the caller pass cudaStreamLegacy ((cudaStream_t)0x1), not NULL. */
PCP_API int pcp_set_stream(pcp_ctx* ctx, void* cuda_stream);
PCP_API int pcp_synchronize(pcp_ctx* ctx);
PCP_API int pcp_device_count(void);
PCP_API const char* pcp_version(void);
/* Release cached arena blocks back to the driver. */
PCP_API int pcp_trim(pcp_ctx* ctx);

/* Stage timing: when enabled the library brackets each kernel stage with CUDA events on its stream.
 * pcp_profile_get copies up to cap entries (name, accumulated ms, launches) recorded since the last
 * pcp_profile_reset and returns the number of stages via *n. */
typedef struct pcp_stage_time {
    char name[48];
    double ms;
    int64_t launches;
    double bytes; /* algorithmic bytes attributed to the stage (0 if not applicable) */
} pcp_stage_time;
PCP_API int pcp_profile_enable(pcp_ctx* ctx, int on);
PCP_API int pcp_profile_reset(pcp_ctx* ctx);
PCP_API int pcp_profile_get(pcp_ctx* ctx, pcp_stage_time* out, int cap, int* n);
/* Number of kernels launched by this context since creation (bench.py's gpu_launches). */
PCP_API int64_t pcp_launch_count(const pcp_ctx* ctx);

/* Raw device / pinned-host memory for callers that keep clouds resident between calls. */
PCP_API int pcp_dev_alloc(pcp_ctx* ctx, size_t bytes, void** dptr);
PCP_API int pcp_dev_free(pcp_ctx* ctx, void* dptr);
PCP_API int pcp_host_alloc_pinned(pcp_ctx* ctx, size_t bytes, void** hptr);
PCP_API int pcp_host_free_pinned(pcp_ctx* ctx, void* hptr);
PCP_API int pcp_memcpy(pcp_ctx* ctx, void* dst, const void* src, size_t bytes); /* any direction, synchronous */
/* Same, ordered on the context's stream without waiting (host memory should be pinned to overlap). */
PCP_API int pcp_memcpy_async(pcp_ctx* ctx, void* dst, const void* src, size_t bytes);
/* Copy lanes: PCP_COPY_LANES extra streams for host<->device copies that overlap the context's compute stream
 * (upload the next cloud and download the previous results while the current one is processed).
 *   pcp_copy_async  the copy starts once the work submitted to the context SO FAR has finished (so it may read a
 *                   result buffer, or overwrite an input buffer, of an earlier call); returns at once.
 *   pcp_copy_fence  work submitted to the context from now on waits for the copies enqueued on the lane so far.
 *   pcp_copy_wait   the calling thread waits for the copies enqueued on the lane so far.
 * Host memory must be pinned (pcp_host_alloc_pinned) for the copy to overlap. */
#define PCP_COPY_LANES 4
PCP_API int pcp_copy_async(pcp_ctx* ctx, int lane, void* dst, const void* src, size_t bytes);
PCP_API int pcp_copy_fence(pcp_ctx* ctx, int lane);
PCP_API int pcp_copy_wait(pcp_ctx* ctx, int lane);

/* ---- search structure ------------------------------------------------------------------------------
 * Replaces KdTreeFLANN<PointT>::setInputCloud + convertCloudToArray (kd_tree.h:772-798, 928-956):
 * non-finite points are dropped and results carry ORIGINAL cloud indices.  The FLANN kd-tree is replaced
 * by Morton-sorted points with per-level cell tables.  cell_size <= 0 picks the base cell from the data. */
PCP_API int pcp_index_build(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double cell_size,
                            pcp_index** out);
PCP_API void pcp_index_free(pcp_index* index);
PCP_API int64_t pcp_index_size(const pcp_index* index);      /* valid (finite) points, total_nr_points_ */
PCP_API double pcp_index_cell_size(const pcp_index* index);

/* KdTreeFLANN::nearestKSearch (kd_tree.h:814-845) for nq query points at once.
 * idx: nq*k original indices ascending by (d2, index), padded with -1; d2: nq*k squared distances
 * (L2_Simple order x,y,z, no FMA) padded with +inf, or NULL; found: nq counts (= min(k, valid)), or NULL. */
PCP_API int pcp_knn(const pcp_index* index, const void* q, size_t nq, size_t qstride, int k, int32_t* idx,
                    double* d2, int32_t* found);

/* KdTreeFLANN::radiusSearch (kd_tree.h:863-903): strict d2 < r*r, sorted ascending by (d2, index),
 * capped at max_nn (0 or > valid = unlimited).  Two calls: count fills offsets[nq+1] (exclusive prefix
 * sums), fill writes idx / d2 (d2 may be NULL) at those offsets. */
PCP_API int pcp_radius_count(const pcp_index* index, const void* q, size_t nq, size_t qstride, double r,
                             uint32_t max_nn, int64_t* offsets);
PCP_API int pcp_radius_fill(const pcp_index* index, const void* q, size_t nq, size_t qstride, double r,
                            uint32_t max_nn, const int64_t* offsets, int32_t* idx, double* d2);

/* ---- features --------------------------------------------------------------------------------------
 * The deterministic per-point feature pass the reference declares but never defines,
 * CalculateFeature::calculate_plan_parameter(CloudPtr, double radius) (calculate_feature.h:15):
 * for every cloud point j (in original order) neighbours = nearestKSearch(point j, k) -- self included,
 * as at calculate_feature.cpp:230-233 -- then calculate_plan_parameter_h_points on them
 * (calculate_feature.cpp:119-206) and the N<=3 fallback of calculate_feature.cpp:356-363.
 * props: n records of pcp_point_property for the ORIGINAL n points (dropped points get the fallback).
 * eig3 (n*3 doubles lambda1>=lambda2>=lambda3), lps3 (n*3 floats linearity, planarity, scattering =
 * (l1-l2)/l1, (l2-l3)/l1, l3/l1), knn_idx (n*k) are optional (NULL). */
PCP_API int pcp_features_knn(const pcp_index* index, int k, void* props, double* eig3, float* lps3,
                             int32_t* knn_idx);
/* Same, for a shard of a larger cloud: only the points with original index < query_limit are
 * queries (the rest -- halo points received from neighbouring shards -- are candidates only; query_limit < 0 =
 * all), and kth_d2 (n floats, optional) receives an upper bound of every query's k-th squared distance, which
 * tells the caller whether the neighbourhood can reach beyond the halo. */
PCP_API int pcp_features_knn_ex(const pcp_index* index, int k, int64_t query_limit, void* props, double* eig3,
                                float* lps3, int32_t* knn_idx, float* kth_d2);
/* Same with radius neighbourhoods (strict d2 < r*r, max_nn as above); counts (n) optional. */
PCP_API int pcp_features_radius(const pcp_index* index, double r, uint32_t max_nn, void* props, double* eig3,
                                float* lps3, int32_t* counts);
/* CalculateFeature::calculate_plan_parameter_rpca (calculate_feature.cpp:208-368), the robust fit the reference
 * ships: for every point, its k nearest neighbours (:230-233); `iterations` = compute_iteration_number(Pr, epi, 3)
 * (:28-33) three-point hypotheses (:243-256); for each the h = (int)(2/3 N) neighbours nearest the hypothesis
 * plane are refitted through their FLOAT coordinates (LAS_POINT_PROPERTY_sim, :258-275) and the refit with the
 * smallest lambda3 wins (:277-289); the neighbours whose distance to it passes the MAD test Rz < 2.5 (:291-331)
 * are fitted once more (:338-355); N <= 3 or <= 3 inliers give the fallback record (:356-363).
 * The reference draws hypotheses with rand() under OpenMP (srand(time), :210,:249) and is not reproducible; here
 * draw n of hypothesis i of point j is (splitmix64(seed*0x9E3779B97F4A7C15 + j*0xD1B54A32D192ED03 + 3i+n) >> 33) % N,
 * so the result is a pure function of (cloud, k, Pr, epi, seed). Ties in the two sorts keep neighbour order and
 * the first hypothesis with the smallest lambda3 wins (the reference's std::sort leaves both unspecified).
 * k <= 32 (the reference uses 20). props: n records for the ORIGINAL n points. */
PCP_API int pcp_features_rpca(const pcp_index* index, int k, float Pr, float epi, uint64_t seed, void* props);
/* compute_iteration_number(Pr, epi, 3) (calculate_feature.cpp:28-33): 34 for the reference's 0.99 / 0.5. */
PCP_API int pcp_rpca_iterations(float Pr, float epi);

/* TreeExtration::region_growning (extraction_tree.cpp:170-272), the consumer of the records above in
 * point_segment (static.cpp:8-21: distanceT 0.5, cosfaT 0.94) -- SURVEY 8(f)-1.  Seeds are taken in ascending point
 * index among the points never touched by an earlier region (curvature < curv_t, 0.005 at :194); a FIFO pops members
 * and admits the still unsegmented ones among the k (50 at :222) nearest neighbours of each whose normal has
 * |cos| > cosfaT with the FIRST seed's normal (:25-33, float) and whose float-truncated coordinates lie within
 * distanceT of the popped member's plane (:47-64); a region of more than min_points (5 at :249) members is kept with
 * the next label, a smaller one is dissolved (its members can be absorbed later but never seed).  Results are those
 * of the sequential loop: labels, membership and member order.
 * index: built over the cloud the records describe.  props: n records, read (normal, Distance, curvature, PointID)
 * and SegmentID rewritten (-1 = UNSEGMENTATION).  seg_offsets (seg_cap + 1), seg_points (n: the members of kept
 * regions in PlanSegment::PointID order), seg_plane (seg_cap x 4 floats: PlanSegment normal_x/y/z, Distance = the
 * first seed's).  PCP_ERR_RANGE when more than seg_cap regions are found (n / (min_points + 1) + 1 always suffices).
 * stats (optional, 6 values): seeds tried, single-block passes, all-block passes, candidates scanned, members kept,
 * low-curvature points. */
PCP_API int pcp_region_growing(const pcp_index* index, void* props, int k, double distanceT, double cosfaT,
                               double curv_t, int min_points, int64_t* seg_offsets, size_t seg_cap,
                               int32_t* seg_points, float* seg_plane, size_t* nseg, uint64_t* stats);

/* CalculateFeature::calculate_plan_parameter_h_points (calculate_feature.cpp:119-206) for nseg segments:
 * segment s is the points [seg_offsets[s], seg_offsets[s+1]) of pts. */
PCP_API int pcp_plane_fit(pcp_ctx* ctx, const void* pts, size_t stride, const int64_t* seg_offsets,
                          size_t nseg, pcp_plan_segment* out);

/* ---- voxel grid ------------------------------------------------------------------------------------
 * VoxelGrid<PointXYZRGBA>::applyFilter (voxel_grid.h:811-1056, no field filter) incl. its numerics:
 * truncating cell index, DBL_MIN max seed, float-rounded field sums in double, rgb re-pack.
 * out_pts: capacity n records of `stride` bytes; *m receives the number of occupied cells; output is in
 * ascending (k,j,i) cell order.  min_b/div_b (3 ints each), point_cell (n*3: cell of every input point,
 * -1 for dropped points), cell_count (capacity n) are optional.  Non-finite points are always dropped. */
PCP_API int pcp_voxel_grid(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, const double leaf[3],
                           int downsample_all, void* out_pts, size_t* m, int32_t min_b[3], int32_t div_b[3],
                           int32_t* point_cell, int32_t* cell_count);

/* PointCloudHelper::remove_duplicate (point_cloud_helper.cpp:42-63): centre on the serial-order centroid
 * (point_cloud_helper.h:192-230, reproduced bit-exactly), voxel grid with a float leaf, un-centre.
 * In place: pts_inout holds *n_inout records on entry and the filtered cloud on return. */
PCP_API int pcp_remove_duplicate(pcp_ctx* ctx, void* pts_inout, size_t* n_inout, size_t stride, float leaf);

/* Out of place: pts (n records) is left alone, the filtered cloud goes to pts_out (capacity n records of the same
 * stride), *m receives its size -- no copy of the cloud is made at all (the in-place form needs two).  centre, min_b,
 * div_b: NULL for the reference behaviour (own centroid, own box); centre alone, or all three, as
 * pcp_remove_duplicate_centred below. */
PCP_API int pcp_remove_duplicate_to(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, float leaf, const double* centre,
                                    const int32_t* min_b, const int32_t* div_b, void* pts_out, size_t* m);

/* LAS_POINT_PROPERTY records <-> packed records (pcp_point_property_packed).  pack: host or device arrays, on the
 * context's stream (asynchronous when both are device arrays); unpack: plain host code on `threads` threads
 * (<= 0: one per core), for callers that want the reference's struct back. */
PCP_API int pcp_props_pack(pcp_ctx* ctx, const void* props, size_t n, pcp_point_property_packed* packed);
PCP_API int pcp_props_unpack(const pcp_point_property_packed* packed, size_t n, void* props, int threads);
/* PointXYZRGBA records (stride 48) <-> packed cloud records (stride 32, see the conventions above); host or device
 * arrays.  unpack sets w = 1 (point_type.h:23) and zeroes the padding. */
PCP_API int pcp_cloud_pack(pcp_ctx* ctx, const void* pts48, size_t n, void* pts32);
PCP_API int pcp_cloud_unpack(pcp_ctx* ctx, const void* pts32, size_t n, void* pts48);

/* ---- sharding (one shard per GPU; the cut follows cloud_grid.cpp's tiles) --------------------
 * remove_duplicate of a SHARD of a larger cloud: centred on the given centre (the whole cloud's serial-order
 * centroid) and voxelised in the whole cloud's grid box (min_b / div_b of the centred whole cloud,
 * voxel_grid.h:835-847), so that shards cut along voxel columns reproduce the single-device cells exactly. */
PCP_API int pcp_remove_duplicate_centred(pcp_ctx* ctx, void* pts_inout, size_t* n_inout, size_t stride, float leaf,
                                         const double centre[3], const int32_t min_b[3], const int32_t div_b[3]);
/* compute3DCentroid's serial-order sums continued over a chunk: sum3 (in/out) holds the running x,y,z sums of
 * the points that precede the chunk, count (in/out, optional) the running number of finite points. */
PCP_API int pcp_serial_sum(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double sum3[3], uint64_t* count);
/* The same in two steps, for a cloud whose point order runs over several chunks or GPUs (DEVICE memory): `begin`
 * does the parallel part for this chunk under a PREDICTED carry (an approximate sum of all points before the chunk,
 * e.g. the tree-order sums of pcp_cloud_stats: only its binades matter; NULL = 0), `end` continues the exact
 * running sums carry3_inout (host or device array: a device array is walked in place without synchronising, so the
 * value can travel on to the next GPU) through the chunk -- tens of microseconds, the only sequential step of a
 * chain of chunks -- and releases the handle (carry3_inout == NULL just releases). */
PCP_API int pcp_serial_sum_begin(pcp_ctx* ctx, const void* dev_pts, size_t n, size_t stride, const double predicted3[3],
                                 void** handle);
PCP_API int pcp_serial_sum_end(void* handle, double* carry3_inout);
/* One pass over a cloud: out10 (host or device array) = min x,y,z | max x,y,z (true extremes of the finite points;
 * getMinMax3D's DBL_MIN seed is the caller's) | x,y,z sums in TREE order (a prediction for pcp_serial_sum_begin,
 * never a result) | number of finite points. */
PCP_API int pcp_cloud_stats(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double* out10);
/* Stable partition of a cloud by destination shard: part p takes the finite points whose voxel column along `axis`
 * lies in [bounds[p], bounds[p+1]) (columns below bounds[1] go to part 0, columns from bounds[nparts-1] up to the last
 * part); out_pts receives the parts one after the other, input order kept inside each; counts[nparts] (host) their
 * sizes.  At most 64 parts, stride a multiple of 16.  The sender's half of the all-to-all that moves every point
 * to the GPU that owns its voxel. */
PCP_API int pcp_partition_columns(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, float leaf, double centre,
                                  int32_t min_b, const int32_t* bounds, int nparts, void* out_pts, uint64_t* counts);
/* At most 64 KiB of device memory read on the host (peek) / written from the host (poke), ordered behind the work
 * submitted to the context so far and never queued behind a bulk copy on a copy lane: a kernel moves them through
 * mapped host memory.  (A 4-byte cudaMemcpy waits for every piece of a gigabyte upload queued before it on the copy
 * engine: that alone serialised the pipelined passes with their own uploads.)  Both wait for the kernel. */
PCP_API int pcp_peek(pcp_ctx* ctx, const void* dev_src, size_t bytes, void* host_dst);
PCP_API int pcp_poke(pcp_ctx* ctx, void* dev_dst, size_t bytes, const void* host_src);
/* Histogram of voxel columns along `axis`: column = (int)((p + (-centre)) * (1/leaf) - min_b) as at
 * voxel_grid.h:936; bin = (column - col0) / cols_per_bin clamped to [0, nbins); counts are ADDED to hist_inout. */
PCP_API int pcp_column_histogram(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, float leaf,
                                 double centre, int32_t min_b, int32_t col0, int32_t cols_per_bin, int32_t nbins,
                                 uint64_t* hist_inout);
/* Stable selection (input order kept) of the finite points with lo <= coordinate < hi / col_lo <= column < col_hi. */
PCP_API int pcp_select_range(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, double lo, double hi,
                             void* out_pts, size_t* m);
PCP_API int pcp_select_columns(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, float leaf,
                               double centre, int32_t min_b, int32_t col_lo, int32_t col_hi, void* out_pts, size_t* m);
/* The queries of a shard whose k-th neighbour may lie beyond the halo the shard received: point i of the first m
 * is selected when kth_d2[i] (pcp_features_knn_ex) >= reach^2, reach = coordinate - lo (has_lo) or hi - coordinate
 * (has_hi), lo / hi being the outer faces of the halo along `axis`.  Writes the selected indices (ascending) and
 * their coordinates (3 doubles each), at most `cap`; *count receives the number selected (PCP_ERR_RANGE when it
 * exceeds cap: call again with more room). */
PCP_API int pcp_select_boundary(pcp_ctx* ctx, const void* pts, size_t m, size_t stride, int axis, const float* kth_d2,
                                int has_lo, double lo, int has_hi, double hi, size_t cap, int32_t* out_idx,
                                double* out_xyz, size_t* count);
/* Exact features of queries whose neighbourhood spans shards: list l (of nlists <= 4) holds, for each of the nq
 * queries, k candidate neighbours sorted by squared distance (xyz_lists [nlists][nq][k][3], d2_lists
 * [nlists][nq][k], +inf padded).  The k nearest distinct candidates are merged and fitted
 * (calculate_feature.cpp:119-206); results go to props[point_id[q]] (device arrays). */
PCP_API int pcp_features_merge(pcp_ctx* ctx, size_t nq, int k, int nlists, const double* xyz_lists,
                               const double* d2_lists, const int32_t* point_id, void* props, double* eig3, float* lps3);

/* ---- several GPUs of one process ------------------------------------------------------------------------------
 * The whole pipeline of point_cloud_spliter.cpp:127-128 -- remove_duplicate(leaf), then the per-point plane fit on the
 * k nearest neighbours -- over the GPUs of one process, for a caller shaped like the reference's main.cpp: one cloud
 * in host memory in, results in host memory out, no launcher, no torch, no NCCL (the exchanges are peer-to-peer
 * copies over NVLink between host-thread barriers; one host thread per device inside).  The reference has no
 * multi-device code; the cut follows its own chunking (5 m blocks, point_cloud_spliter.cpp:221-234) and CloudGrid's
 * 1 m tiles (cloud_grid.cpp:34-78): slabs of 1 m voxel columns along the longer horizontal side, with a halo.
 *   pts      n records of stride 32 (packed) or 48 (PointXYZRGBA) in HOST memory;
 *   leaf     remove_duplicate's leaf (float, point_cloud_helper.h:16); <= 0 skips the voxel grid;
 *   halo     metres of neighbouring slab each device also sees (>= the k-th neighbour distance of nearly all points;
 *            2 m for street-scene clouds); the result does not depend on it (queries that outreach it are answered
 *            by the neighbour exactly);
 *   out_pts  capacity n records of the same stride: the voxelised cloud; props: capacity n records of 48 bytes,
 *            props[i] describes out_pts[i] (PointID = i); *m = their number.
 * The records are those of the single-device calls, point for point; their ORDER is slab after slab along the cut
 * axis, each slab in voxel order (the single-device order is voxel order over the whole cloud). */
typedef struct pcp_mgpu pcp_mgpu;
PCP_API int pcp_mgpu_create(int ndev, const int* devices /* NULL: 0 .. ndev-1 */, pcp_mgpu** out);
PCP_API void pcp_mgpu_destroy(pcp_mgpu* mg);
PCP_API int pcp_mgpu_device_count(const pcp_mgpu* mg);
PCP_API pcp_ctx* pcp_mgpu_ctx(pcp_mgpu* mg, int i); /* device i's context (profiling, synchronisation) */
PCP_API const char* pcp_mgpu_last_error(const pcp_mgpu* mg);
PCP_API int pcp_mgpu_features_knn(pcp_mgpu* mg, const void* pts, size_t n, size_t stride, float leaf, int k, double halo,
                                  void* out_pts, void* props, size_t* m);

/* compute3DCentroid (point_cloud_helper.h:192-230) and getMinMax3D (point_cloud_helper.h:59-90). */
PCP_API int pcp_centroid(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double c[3], uint64_t* count);
PCP_API int pcp_minmax(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double mn[3], double mx[3]);

/* ---- the key sort ---------------------------------------------------------------------------------------
 * The stable LSD radix sort (csrc/sort.cu, hand-written for sm_100a) that stands where the reference sorts its cell
 * indices (std::sort, voxel_grid.h:948) and FLANN builds its tree (kd_tree.h:795-797); exported for callers that
 * bring their own keys and for the parity tests.  Sorts n (key, value) pairs in place by the low `bits` bits of the
 * key (key_bytes 4 or 8); equal keys keep their order.  vals == NULL: keys only.  n < 2^30. */
PCP_API int pcp_sort_pairs(pcp_ctx* ctx, void* keys, uint32_t* vals, size_t n, int key_bytes, int bits);
/* Exclusive prefix sums of n uint32 values (out may equal in). */
PCP_API int pcp_exclusive_sum_u32(pcp_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n);
/* The same prefix sums for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scan's scratch (one word per tile of 4096 items) lives in a fixed 64 KiB array in device memory that every
 * GPU loading the library holds until the process exits, so the call makes no cudaMalloc / cudaFree / pool
 * allocation.  Limits: n <= PCP_NOALLOC_SCAN_MAX_N (PCP_ERR_RANGE beyond it); in and out must be DEVICE pointers
 * on the context's GPU (PCP_ERR_INVALID otherwise; out may equal in).  The call is asynchronous on the context's
 * stream (pcp_synchronize waits).  All calls on a device share the one array: calls that overlap on the GPU (other
 * streams, other contexts on the same device) race; calls on one stream are ordered.  Results are bit-identical
 * to pcp_exclusive_sum_u32's. */
#define PCP_NOALLOC_SCAN_MAX_N (16384 * 4096) /* 67 108 864 items */
PCP_API int pcp_exclusive_sum_u32_noalloc(pcp_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n);

/* ---- synthetic clouds (bench / tests) ----------------------------------------------------
 * Fills n PointXYZRGBA records (stride 48) of the named kind; point i is a pure function of
 * (kind, n_total, seed, first + i).  kind: 0 uniform_cube, 1 strip, 2 forest, 3 city, 4 lattice.
 * centred != 0 omits the UTM-scale world offset. */
PCP_API int pcp_synth_cloud(pcp_ctx* ctx, int kind, uint64_t n_total, uint64_t first, uint64_t n, uint64_t seed,
                            int centred, void* out_pts);

/* ---- PCD v0.7 files and LZF (host side; pcp_pcd_load_many assembles records on the device) ----------------------------------------------------
 * Replace io::loadPCDFile / savePCDFile* (pcd_helper.h:1343-1497), PCDReader::readHeader / read
 * (pcd_helper.cpp:71-333, 845-1181), PCDWriter::generateHeader / writeASCII / writeBinary /
 * writeBinaryCompressed (pcd_helper.h:323-371, 489-1090) and lzfCompress / lzfDecompress (lzf.cpp:86-415).
 * Errors: negative status, text through pcp_last_error(NULL). */
#define PCP_PCD_ASCII 0
#define PCP_PCD_BINARY 1
#define PCP_PCD_BINARY_COMPRESSED 2
#define PCP_PCD_POINT_XYZRGBA 0 /* records shaped like PointXYZRGBA: x, y, z doubles @0, w @24, rgba @32, stamp_id @36 */
#define PCP_PCD_POINT_XYZ 1     /* records that start with x, y, z doubles (PointC); only those fields are stored */
#define PCP_PCD_MAX_FIELDS 16
typedef struct pcp_pcd_header {
    uint64_t points;
    uint64_t data_offset; /* first byte of the payload */
    uint32_t width, height;
    int32_t data_type;    /* PCP_PCD_ASCII / BINARY / BINARY_COMPRESSED */
    int32_t nfields;
    char names[PCP_PCD_MAX_FIELDS][32];
    int32_t size[PCP_PCD_MAX_FIELDS];
    int32_t count[PCP_PCD_MAX_FIELDS];
    char type[PCP_PCD_MAX_FIELDS]; /* 'F', 'U', 'I' */
    double viewpoint[7];           /* tx ty tz qw qx qy qz */
} pcp_pcd_header;
PCP_API int pcp_pcd_read_header(const char* path, pcp_pcd_header* header);
/* Reads the file's points into `cap` records of `stride` bytes.  Fields are matched to the record's members on
 * name + type + size + count as point_traits.h:271-279 does; other fields are ignored and unmatched members keep
 * the default constructor's values (0, w = 1).  *n = points in the file (also on PCP_ERR_RANGE when cap is too
 * small).  *is_dense = 0 when a binary payload holds a non-finite value (ascii files are not checked, as in the
 * reference). */
PCP_API int pcp_pcd_load(const char* path, int point_kind, void* pts, size_t cap, size_t stride, size_t* n, int* is_dense);
/* The pipelined loader: nfiles PCD files into ONE cloud at pts (host or device memory, capacity cap
 * records of `stride` bytes, a multiple of 8); file f's points become records [offsets[f], offsets[f+1]) (offsets:
 * nfiles + 1 values).  `threads` host threads (<= 0: one per core) read the payloads and LZF-decode the
 * binary_compressed ones, one file each; binary and binary_compressed payloads are uploaded as stored (packed records
 * or field planes: 32 B/point for the reference's clouds instead of 48) and the records are assembled on the device
 * with the same field matching as pcp_pcd_load; ascii payloads are parsed on the host.  *is_dense = AND over the files.
 * With pts in device memory the cloud is ready for pcp_remove_duplicate / pcp_index_build without another copy. */
PCP_API int pcp_pcd_load_many(pcp_ctx* ctx, const char* const* paths, size_t nfiles, int point_kind, void* pts, size_t cap,
                              size_t stride, int threads, size_t* offsets, int* is_dense);
/* Writes n records.  width = height = 0 means an unorganised cloud (n x 1).  mode: PCP_PCD_*; precision: significant
 * digits of the ascii writer (the reference's default is 8).  The payload holds the registered fields only, packed;
 * binary_compressed is limited to 4 GiB of them (32-bit size words). */
PCP_API int pcp_pcd_save(const char* path, int point_kind, const void* pts, size_t n, size_t stride, uint32_t width,
                         uint32_t height, int mode, int precision);
/* liblzf streams as the reference's lzf.cpp writes them (byte-identical); 0 = failure (output too small / corrupt). */
PCP_API unsigned int pcp_lzf_compress(const void* in, unsigned int in_len, void* out, unsigned int out_len);
PCP_API unsigned int pcp_lzf_decompress(const void* in, unsigned int in_len, void* out, unsigned int out_len);

#ifdef __cplusplus
}
#endif
#endif
