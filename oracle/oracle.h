/*
 * oracle.h -- CPU restatement of the PointCloudProcess hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * load this library.  Nothing under pointcloudprocess_b200/ or include/ links or imports it.
 *
 * Parity status (see DESIGN.md section "oracle"):
 *   - voxel grid / remove_duplicate: restates /root/reference voxel_grid.h:811-1056,
 *     concatenate.h:94-162, point_cloud_helper.h:22-90,192-230, point_cloud_helper.cpp:42-73.
 *     Pinned by the reference's derivable known-answer test main_test.cpp:126-154.
 *   - kNN / radius: kd_tree.h:772-903 semantics; the arithmetic lives in FLANN (un-vendored,
 *     version unpinned: "PCL 3rdParty\FLANN", cloudblenddouble.vcxproj:135).  Restated from
 *     FLANN's published L2_Simple functor (diff = a-b; result += diff*diff, x,y,z order).
 *     Pinned by main_test.cpp:156-188 known answer + scipy.spatial.cKDTree cross-check.
 *   - plane fit: calculate_feature.cpp:119-206; eigen-solver is OpenCV 2.4.8 cvEigenVV
 *     (opencv_core248.lib, un-vendored) restated from its published Jacobi (lapack.cpp
 *     JacobiImpl_); pinned against cv2.eigen outputs committed under tests/golden/.
 *
 * All point arguments are (base pointer, count, byte stride); the first 24 bytes of a record
 * are x,y,z doubles (PointXYZRGBA point_type.h:9-82 stride 48, PointC stride 64, packed 24).
 */
#ifndef PCP_ORACLE_H
#define PCP_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* LAS_POINT_PROPERTY, data_struct.h:161-173 (48 bytes) */
typedef struct orc_las_point_property {
    float normal_x, normal_y, normal_z;
    double Distance;
    double curvature;
    int32_t PointID;
    int32_t SegmentID;
    float dis_from_point_plane;
} orc_las_point_property;

/* PlanSegment without the std::vector member, data_struct.h:188-198 */
typedef struct orc_plan_segment {
    float normal_x, normal_y, normal_z;
    float min_value;
    float curvature;
    float Distance;
    double eig[3]; /* lambda1 >= lambda2 >= lambda3 as cvEigenVV returns them (build-defined export) */
} orc_plan_segment;

int orc_num_threads(void);

/* OpenCV JacobiImpl_<double> restated (n x n symmetric, row-major A destroyed; W descending; V rows). */
void orc_jacobi(double* A, int n, double* W, double* V);

/* calculate_feature.cpp:119-206 on h packed xyz triples. */
void orc_plane_fit(const double* xyz, int h, orc_plan_segment* out);

/* kd_tree.h:928-956: indices of finite points in order; returns count. map may be NULL. */
size_t orc_valid_points(const void* pts, size_t n, size_t stride, int32_t* map);

/* Brute-force exact kNN (kd_tree.h:814-845 semantics). idx/d2 are nq*k, padded with -1 / +inf when
 * fewer than k valid points exist. Order key (d2, original index). found[nq]. */
void orc_knn_brute(const void* pts, size_t n, size_t stride, const void* q, size_t nq, size_t qstride,
                   int k, int32_t* idx, double* d2, int32_t* found);

/* kd-tree (leaf <= 15, kd_tree.h:795-797) handle */
typedef struct orc_kdtree orc_kdtree;
orc_kdtree* orc_kdtree_build(const void* pts, size_t n, size_t stride);
void orc_kdtree_free(orc_kdtree*);
size_t orc_kdtree_size(const orc_kdtree*);
void orc_kdtree_knn(const orc_kdtree*, const void* q, size_t nq, size_t qstride, int k,
                    int32_t* idx, double* d2, int32_t* found, int threads);
/* radius: strict d2 < r*r, sorted by (d2, idx); CSR. Pass idx==NULL to only count (offsets[nq+1]). */
void orc_kdtree_radius(const orc_kdtree*, const void* q, size_t nq, size_t qstride, double r,
                       uint32_t max_nn, int64_t* offsets, int32_t* idx, double* d2, int threads);
/* number of queries whose rank-k boundary is an exact distance tie (parity harness reports them) */
int64_t orc_count_boundary_ties(const orc_kdtree*, const void* q, size_t nq, size_t qstride, int k, int threads);

/* a6: per-point kNN(k) -> plane fit. props[n]; eig3 (n*3) / lps3 (n*3) / knn_idx (n*k) optional. */
void orc_features_knn(const void* pts, size_t n, size_t stride, int k, orc_las_point_property* props,
                      double* eig3, float* lps3, int32_t* knn_idx, int threads);
void orc_features_radius(const void* pts, size_t n, size_t stride, double r, uint32_t max_nn,
                         orc_las_point_property* props, double* eig3, float* lps3, int32_t* counts, int threads);
/* a7: calculate_plan_parameter_rpca (calculate_feature.cpp:208-368) with a counter-based generator in place of
 * rand(): 3-point hypotheses, h = 2N/3 subset fits on float-truncated coordinates, MAD inlier test, final fit. */
int orc_rpca_iterations(float Pr, float epi);
int orc_rpca_point(const double* nb_xyz, int N, int iterations, uint64_t seed, uint64_t j, orc_plan_segment* out);
void orc_features_rpca(const void* pts, size_t n, size_t stride, int k, float Pr, float epi, uint64_t seed,
                       orc_las_point_property* props, int threads);

/* TreeExtration::compute_distance_from_point_to_plane, extraction_tree.cpp:47-64 (float-truncated coordinates). */
float orc_point_plane_distance(const double* xyz, float a, float b, float c, float d);
/* TreeExtration::region_growning, extraction_tree.cpp:170-272 (k = 50, curv_t = 0.005, min_points = 5 there).
 * props: n records; SegmentID is rewritten (-1 = UNSEGMENTATION, else the label of a kept region).
 * Kept regions in the order they were found: seg_offsets (capacity n/ (min_points+1) + 2), seg_points (capacity n; the
 * members in the order the reference pushes them into PlanSegment::PointID), seg_plane (4 floats per region: the
 * first seed's normal and Distance).  Returns the number of kept regions. */
int64_t orc_region_growing(const void* pts, size_t n, size_t stride, orc_las_point_property* props, int k,
                           double distanceT, double cosfaT, double curv_t, int min_points,
                           int64_t* seg_offsets, int32_t* seg_points, float* seg_plane, int threads);

/* VoxelGrid::applyFilter, voxel_grid.h:811-1056 (no field filter). Returns M, or -1 on error.
 * out_pts has capacity n records of `stride` bytes (x,y,z double @0, rgba u32 @32, stamp_id u32 @36).
 * point_cell (n*3, or NULL): (i,j,k) of each input point, -1 for skipped points.
 * cell_count (capacity n, or NULL): points per output cell.  overflow32: set to 1 if the reference's
 * 32-bit idx would have overflowed (output then follows the 64-bit key order). */
int64_t orc_voxel_grid(const void* pts, size_t n, size_t stride, const double leaf[3], int downsample_all,
                       int is_dense, void* out_pts, int32_t min_b[3], int32_t div_b[3],
                       int32_t* point_cell, int32_t* cell_count, int* overflow32);

/* point_cloud_helper.cpp:42-63. In place; returns M. center3 (optional) receives the centroid used. */
int64_t orc_remove_duplicate(void* pts, size_t n, size_t stride, float leaf, double* center3);

/* point_cloud_helper.h:192-230 serial-order centroid. */
void orc_centroid(const void* pts, size_t n, size_t stride, int is_dense, double c[3], uint32_t* count);
void orc_minmax(const void* pts, size_t n, size_t stride, int is_dense, double mn[3], double mx[3]);

/* LZF (lzf.cpp:86-415 algorithm, Marc Lehmann HLOG 13) restated. */
unsigned int orc_lzf_compress(const void* in, unsigned int in_len, void* out, unsigned int out_len);
unsigned int orc_lzf_decompress(const void* in, unsigned int in_len, void* out, unsigned int out_len);

#ifdef __cplusplus
}
#endif
#endif
