// shard.cu -- the pieces a cloud needs to be cut into spatial shards: the running serial-order
// sum over chunks (so every shard is centred on the centroid of the WHOLE cloud, point_cloud_helper.cpp:44-52),
// a histogram of voxel columns along one axis (to balance the cut), a stable selection of the points of a
// column / coordinate range (shard extraction, halo extraction), and the exact merge of per-shard neighbour
// candidates for queries whose k-th neighbour reaches beyond the halo.
#include "plane_fit.cuh"
#include "reduce.cuh"
#include "sort.cuh"

namespace {

// voxel column of a coordinate: the expression of voxel_grid.h:936 applied to the centred coordinate
// p + (-c) of transformPointCloud (point_cloud_helper.h:116-124)
__device__ __forceinline__ int column_of(double v, double centre, double inv_leaf, int min_b)
{
    return (int)__dsub_rn(__dmul_rn(__dadd_rn(v, -centre), inv_leaf), (double)min_b);
}

struct SelParams {
    int mode;      // 0: lo <= coordinate < hi;  1: ilo <= voxel column < ihi
    int axis;
    double lo, hi;
    double centre, inv_leaf;
    int min_b;
    int ilo, ihi;
};

__device__ __forceinline__ bool sel_pred(const char* rec, const SelParams& P)
{
    const double* p = reinterpret_cast<const double*>(rec);
    const double x = p[0], y = p[1], z = p[2];
    if (!(isfinite(x) && isfinite(y) && isfinite(z))) return false;
    const double v = P.axis == 0 ? x : (P.axis == 1 ? y : z);
    if (P.mode == 0) return v >= P.lo && v < P.hi;
    const int c = column_of(v, P.centre, P.inv_leaf, P.min_b);
    return c >= P.ilo && c < P.ihi;
}

__global__ void __launch_bounds__(256) sel_flag_kernel(const char* __restrict__ pts, size_t n, size_t stride, SelParams P,
                                                       uint32_t* __restrict__ flag)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) flag[i] = sel_pred(pts + i * stride, P) ? 1u : 0u;
}

// 16-byte granules: records are 16-aligned multiples of 16 bytes (PointXYZRGBA 48, PointC 64) or 24 (8-byte)
__global__ void __launch_bounds__(256) sel_scatter_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                          const uint32_t* __restrict__ flag,
                                                          const uint32_t* __restrict__ scan, char* __restrict__ out,
                                                          uint32_t* __restrict__ m_out)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (flag[i]) {
        const char* s = pts + i * stride;
        char* d = out + (size_t)scan[i] * stride;
        if ((stride & 15) == 0) {
            for (size_t b = 0; b < stride; b += 16) *reinterpret_cast<uint4*>(d + b) = *reinterpret_cast<const uint4*>(s + b);
        } else {
            for (size_t b = 0; b < stride; b += 8) *reinterpret_cast<uint2*>(d + b) = *reinterpret_cast<const uint2*>(s + b);
        }
    }
    if (i == n - 1) *m_out = scan[i] + flag[i];
}

// bins in shared memory (block-private, flushed once) when they fit: one global atomic per point on a few hundred bins
// serialises on the L2 atomic units (2.5 ms for 25 M points)
#define PCP_HIST_SMEM_BINS 4096
__global__ void __launch_bounds__(256) column_hist_kernel(const char* __restrict__ pts, size_t n, size_t stride, int axis,
                                                          double centre, double inv_leaf, int min_b, int col0, int ncols,
                                                          int cols_per_bin, unsigned long long* __restrict__ hist)
{
    __shared__ unsigned int sh[PCP_HIST_SMEM_BINS];
    const bool local = ncols <= PCP_HIST_SMEM_BINS;
    if (local) {
        for (int b = threadIdx.x; b < ncols; b += blockDim.x) sh[b] = 0u;
        __syncthreads();
    }
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const double* p = reinterpret_cast<const double*>(pts + i * stride);
        const double x = p[0], y = p[1], z = p[2];
        if (!(isfinite(x) && isfinite(y) && isfinite(z))) continue;
        const int c = column_of(axis == 0 ? x : (axis == 1 ? y : z), centre, inv_leaf, min_b);
        long long b = ((long long)c - col0) / cols_per_bin;
        if (c < col0) b = 0;
        if (b >= ncols) b = ncols - 1;
        if (local) atomicAdd(&sh[b], 1u);
        else atomicAdd(&hist[b], 1ull);
    }
    if (local) {
        __syncthreads();
        for (int b = threadIdx.x; b < ncols; b += blockDim.x)
            if (sh[b]) atomicAdd(&hist[b], (unsigned long long)sh[b]);
    }
}

} // namespace

extern "C" int pcp_serial_sum(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double sum3[3], uint64_t* count)
{
    if (!ctx || !sum3 || stride < 24 || (n && !pts)) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_serial_sum: bad argument");
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    PCP_TRY(in.open(ctx, pts, n * stride));
    DevBuf<double> s3;
    PCP_TRY(s3.alloc(ctx, 3));
    PCP_TRY(pcp_writeto(ctx, s3.p, sum3, 24));
    PCP_TRY(pcp_serial_sum_device(ctx, in.dptr, n, stride, s3.p, 1));
    pcp_bbox bb;
    PCP_TRY(pcp_bbox_host(ctx, in.dptr, n, stride, &bb));
    PCP_TRY(pcp_readback(ctx, sum3, s3.p, 24));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (count) *count += bb.count;
    return PCP_OK;
}

extern "C" int pcp_column_histogram(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, float leaf,
                                    double centre, int32_t min_b, int32_t col0, int32_t cols_per_bin, int32_t nbins,
                                    uint64_t* hist_inout)
{
    if (!ctx || !hist_inout || stride < 24 || axis < 0 || axis > 2 || nbins < 1 || cols_per_bin < 1 || (n && !pts))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_column_histogram: bad argument");
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    OutStage h;
    PCP_TRY(in.open(ctx, pts, n * stride));
    PCP_TRY(h.open(ctx, hist_inout, (size_t)nbins * 8, true));
    StageTimer t(ctx, "shard.histogram", 8.0 * n);
    unsigned hb = pcp_blocks_for(n, 256 * 16); // 16 points per thread: a block's private bins are flushed once
    if (hb > (unsigned)ctx->sm_count * 8u) hb = (unsigned)ctx->sm_count * 8u;
    column_hist_kernel<<<hb, 256, 0, ctx->stream>>>(static_cast<const char*>(in.dptr), n, stride, axis,
                                                                        centre, 1.0 / (double)leaf, min_b, col0, nbins,
                                                                        cols_per_bin, (unsigned long long*)h.dptr);
    PCP_LAUNCH_CHECK(ctx);
    PCP_TRY(h.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

static int select_impl(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, const SelParams& P, void* out, size_t* m)
{
    *m = 0;
    if (n == 0) return PCP_OK;
    if (n > 0xfffffff0ull) return pcp_fail(ctx, PCP_ERR_RANGE, "select: too many points");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    OutStage o;
    PCP_TRY(in.open(ctx, pts, n * stride));
    PCP_TRY(o.open(ctx, out, n * stride));
    DevBuf<uint32_t> flag, scan, md;
    PCP_TRY(flag.alloc(ctx, n));
    PCP_TRY(scan.alloc(ctx, n));
    PCP_TRY(md.alloc(ctx, 1));
    sel_flag_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(in.dptr), n, stride, P, flag.p);
    PCP_LAUNCH_CHECK(ctx);
    PCP_TRY(pcp_exclusive_scan_u32(ctx, flag.p, scan.p, n));
    sel_scatter_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(in.dptr), n, stride, flag.p,
                                                                        scan.p, static_cast<char*>(o.dptr), md.p);
    PCP_LAUNCH_CHECK(ctx);
    uint32_t mh = 0;
    PCP_TRY(pcp_readback(ctx, &mh, md.p, 4));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    *m = mh;
    PCP_TRY(o.commit((size_t)mh * stride));
    return PCP_OK;
}

extern "C" int pcp_select_range(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, double lo, double hi,
                                void* out_pts, size_t* m)
{
    if (!ctx || !m || stride < 24 || (stride & 7) || axis < 0 || axis > 2 || (n && (!pts || !out_pts)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_select_range: bad argument");
    SelParams P;
    memset(&P, 0, sizeof(P));
    P.mode = 0, P.axis = axis, P.lo = lo, P.hi = hi;
    return select_impl(ctx, pts, n, stride, P, out_pts, m);
}

extern "C" int pcp_select_columns(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, float leaf,
                                  double centre, int32_t min_b, int32_t col_lo, int32_t col_hi, void* out_pts, size_t* m)
{
    if (!ctx || !m || stride < 24 || (stride & 7) || axis < 0 || axis > 2 || !(leaf > 0) || (n && (!pts || !out_pts)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_select_columns: bad argument");
    SelParams P;
    memset(&P, 0, sizeof(P));
    P.mode = 1, P.axis = axis, P.centre = centre, P.inv_leaf = 1.0 / (double)leaf, P.min_b = min_b, P.ilo = col_lo, P.ihi = col_hi;
    return select_impl(ctx, pts, n, stride, P, out_pts, m);
}

// ---- queries whose k-th neighbour may lie beyond the halo ----------------------------------------------------------
namespace {

__global__ void __launch_bounds__(256) bnd_flag_kernel(const char* __restrict__ pts, size_t m, size_t stride, int axis,
                                                       const float* __restrict__ kth, int has_lo, double lo, int has_hi,
                                                       double hi, uint32_t* __restrict__ flag)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= m) return;
    const double v = reinterpret_cast<const double*>(pts + i * stride)[axis];
    const double kd = (double)kth[i];
    bool f = false;
    if (has_lo) {
        const double reach = __dsub_rn(v, lo);
        f |= kd >= __dmul_rn(reach, reach);
    }
    if (has_hi) {
        const double reach = __dsub_rn(hi, v);
        f |= kd >= __dmul_rn(reach, reach);
    }
    flag[i] = f ? 1u : 0u;
}

__global__ void __launch_bounds__(256) bnd_scatter_kernel(const char* __restrict__ pts, size_t m, size_t stride,
                                                          const uint32_t* __restrict__ flag,
                                                          const uint32_t* __restrict__ scan, size_t cap,
                                                          int32_t* __restrict__ out_idx, double* __restrict__ out_xyz,
                                                          uint32_t* __restrict__ count)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= m) return;
    if (flag[i] && scan[i] < cap) {
        const double* p = reinterpret_cast<const double*>(pts + i * stride);
        out_idx[scan[i]] = (int32_t)i;
        out_xyz[3 * (size_t)scan[i] + 0] = p[0];
        out_xyz[3 * (size_t)scan[i] + 1] = p[1];
        out_xyz[3 * (size_t)scan[i] + 2] = p[2];
    }
    if (i == m - 1) *count = scan[i] + flag[i];
}

} // namespace

extern "C" int pcp_select_boundary(pcp_ctx* ctx, const void* pts, size_t m, size_t stride, int axis, const float* kth_d2,
                                   int has_lo, double lo, int has_hi, double hi, size_t cap, int32_t* out_idx,
                                   double* out_xyz, size_t* count)
{
    if (!ctx || !count || stride < 24 || (stride & 7) || axis < 0 || axis > 2 || (m && (!pts || !kth_d2)) ||
        (cap && (!out_idx || !out_xyz)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_select_boundary: bad argument");
    *count = 0;
    if (m == 0 || (!has_lo && !has_hi)) return PCP_OK;
    if (m > 0x7ffffff0ull) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_select_boundary: too many points");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in, kin;
    OutStage oi, ox;
    PCP_TRY(in.open(ctx, pts, m * stride));
    PCP_TRY(kin.open(ctx, kth_d2, m * 4));
    PCP_TRY(oi.open(ctx, out_idx, cap * 4));
    PCP_TRY(ox.open(ctx, out_xyz, cap * 24));
    DevBuf<uint32_t> flag, scan, cd;
    PCP_TRY(flag.alloc(ctx, m));
    PCP_TRY(scan.alloc(ctx, m));
    PCP_TRY(cd.alloc(ctx, 1));
    bnd_flag_kernel<<<pcp_blocks_for(m, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(in.dptr), m, stride, axis,
                                                                     static_cast<const float*>(kin.dptr), has_lo, lo, has_hi,
                                                                     hi, flag.p);
    PCP_LAUNCH_CHECK(ctx);
    PCP_TRY(pcp_exclusive_scan_u32(ctx, flag.p, scan.p, m));
    bnd_scatter_kernel<<<pcp_blocks_for(m, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(in.dptr), m, stride, flag.p,
                                                                        scan.p, cap, static_cast<int32_t*>(oi.dptr),
                                                                        static_cast<double*>(ox.dptr), cd.p);
    PCP_LAUNCH_CHECK(ctx);
    uint32_t c = 0;
    PCP_TRY(pcp_readback(ctx, &c, cd.p, 4));
    *count = c;
    if (c > cap) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_select_boundary: %u queries selected, room for %zu", c, cap);
    PCP_TRY(oi.commit((size_t)c * 4));
    PCP_TRY(ox.commit((size_t)c * 24));
    return PCP_OK;
}

// ---- exact merge of per-shard neighbour candidates -------------------------------------------------------------
namespace {

struct MergeCursor {
    const double* xyz; // [nlists][nq][k][3]
    const double* d2;  // [nlists][nq][k]
    size_t nq, q;
    int k, nlists;
    int head[4];
    double ld, lx, ly, lz; // last emitted (duplicates of it are skipped: a halo point seen by two shards)
    __device__ void reset()
    {
        for (int l = 0; l < 4; l++) head[l] = 0;
        ld = -1.0, lx = ly = lz = 0.0;
    }
    // next neighbour in ascending d2 (lists are sorted); false when exhausted
    __device__ bool next(double& x, double& y, double& z, double& d)
    {
        while (true) {
            int best = -1;
            double bd = 0.0;
            for (int l = 0; l < nlists; l++) {
                if (head[l] >= k) continue;
                const double v = d2[((size_t)l * nq + q) * k + head[l]];
                if (!(v < 1.7e308)) { // +inf padding: this list is exhausted
                    head[l] = k;
                    continue;
                }
                if (best < 0 || v < bd) best = l, bd = v;
            }
            if (best < 0) return false;
            const double* p = xyz + (((size_t)best * nq + q) * k + head[best]) * 3;
            head[best]++;
            if (bd == ld && p[0] == lx && p[1] == ly && p[2] == lz) continue;
            x = p[0], y = p[1], z = p[2], d = bd;
            ld = bd, lx = x, ly = y, lz = z;
            return true;
        }
    }
};

__global__ void __launch_bounds__(128) merge_features_kernel(const double* __restrict__ xyz, const double* __restrict__ d2,
                                                             size_t nq, int k, int nlists,
                                                             const int32_t* __restrict__ point_id, void* props,
                                                             double* eig3, float* lps3)
{
    const size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (q >= nq) return;
    MergeCursor C;
    C.xyz = xyz, C.d2 = d2, C.nq = nq, C.q = q, C.k = k, C.nlists = nlists;
    double x, y, z, d;
    // calculate_plan_parameter_h_points (calculate_feature.cpp:131-164) on the k nearest of the merged lists
    C.reset();
    int N = 0;
    double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
    while (N < k && C.next(x, y, z, d)) {
        XAve += x, YAve += y, ZAve += z;
        N++;
    }
    const size_t jq = (size_t)point_id[q];
    pcp_plane pl;
    const bool fitted = N > 3;
    if (fitted) {
        XAve /= (double)N, YAve /= (double)N, ZAve /= (double)N;
        double s00 = 0, s01 = 0, s02 = 0, s11 = 0, s12 = 0, s22 = 0;
        C.reset();
        for (int e = 0; e < N; e++) {
            C.next(x, y, z, d);
            const double X = x - XAve, Y = y - YAve, Z = z - ZAve;
            s00 += X * X, s01 += X * Y, s02 += X * Z, s11 += Y * Y, s12 += Y * Z, s22 += Z * Z;
        }
        pcp_plane_from_scatter(s00, s01, s02, s11, s12, s22, XAve, YAve, ZAve, pl);
    }
    pcp_store_props(props, jq, fitted, pl);
    pcp_store_extra(eig3, lps3, jq, fitted, pl);
}

} // namespace

extern "C" int pcp_features_merge(pcp_ctx* ctx, size_t nq, int k, int nlists, const double* xyz_lists,
                                  const double* d2_lists, const int32_t* point_id, void* props, double* eig3, float* lps3)
{
    if (!ctx || k < 1 || nlists < 1 || nlists > 4 || (nq && (!xyz_lists || !d2_lists || !point_id || !props)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_features_merge: bad argument");
    if (nq == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (pcp_classify(ctx, props) != PCP_PTR_DEVICE || (eig3 && pcp_classify(ctx, eig3) != PCP_PTR_DEVICE) ||
        (lps3 && pcp_classify(ctx, lps3) != PCP_PTR_DEVICE))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_features_merge: outputs are scattered by PointID and must be device arrays");
    InStage ix, id2, ipid;
    PCP_TRY(ix.open(ctx, xyz_lists, (size_t)nlists * nq * k * 24));
    PCP_TRY(id2.open(ctx, d2_lists, (size_t)nlists * nq * k * 8));
    PCP_TRY(ipid.open(ctx, point_id, nq * 4));
    StageTimer t(ctx, "shard.merge_features");
    merge_features_kernel<<<pcp_blocks_for(nq, 128), 128, 0, ctx->stream>>>(
        (const double*)ix.dptr, (const double*)id2.dptr, nq, k, nlists, (const int32_t*)ipid.dptr, props, eig3, lps3);
    PCP_LAUNCH_CHECK(ctx);
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

// ---- redistribution: stable partition of a cloud by destination shard ---------------------------------------------
namespace {

#define PCP_MAX_PARTS 64

struct PartParams {
    int axis, nparts;
    double centre, inv_leaf;
    int min_b;
    int bounds[PCP_MAX_PARTS + 1];
};

__global__ void __launch_bounds__(256) part_key_kernel(const char* __restrict__ pts, size_t n, size_t stride, PartParams P,
                                                       uint32_t* __restrict__ keys)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* p = reinterpret_cast<const double*>(pts + i * stride);
    const double x = p[0], y = p[1], z = p[2];
    uint32_t key = (uint32_t)P.nparts; // non-finite points belong to no voxel: they go last and are not counted
    if (isfinite(x) && isfinite(y) && isfinite(z)) {
        const int c = column_of(P.axis == 0 ? x : (P.axis == 1 ? y : z), P.centre, P.inv_leaf, P.min_b);
        int part = 0;
        while (part + 1 < P.nparts && c >= P.bounds[part + 1]) part++;
        key = (uint32_t)part;
    }
    keys[i] = key;
}

__global__ void __launch_bounds__(256) part_gather_kernel(const char* __restrict__ pts, size_t stride,
                                                          const uint32_t* __restrict__ order, size_t n, char* __restrict__ out)
{
    // 16-byte granules, consecutive threads move consecutive granules of a record
    const size_t g = stride >> 4;
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= n * g) return;
    const size_t i = t / g, b = (t % g) << 4;
    *reinterpret_cast<uint4*>(out + i * stride + b) = *reinterpret_cast<const uint4*>(pts + (size_t)order[i] * stride + b);
}

__global__ void __launch_bounds__(256) part_count_kernel(const uint32_t* __restrict__ sorted_keys, size_t n, int nparts,
                                                         unsigned long long* __restrict__ first /*[nparts + 2]*/)
{
    // first[p] = first position holding a key >= p
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i > n) return;
    const uint32_t cur = i < n ? sorted_keys[i] : (uint32_t)nparts + 1u;
    const uint32_t prev = i > 0 ? sorted_keys[i - 1] : 0xffffffffu; // "-1"
    for (uint32_t p = prev + 1u; p <= cur && p <= (uint32_t)nparts + 1u; p++) first[p] = i;
}

} // namespace

extern "C" int pcp_partition_columns(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, int axis, float leaf, double centre,
                                     int32_t min_b, const int32_t* bounds, int nparts, void* out_pts, uint64_t* counts)
{
    if (!ctx || !bounds || !counts || nparts < 1 || nparts > PCP_MAX_PARTS || stride < 32 || (stride & 15) || axis < 0 || axis > 2 ||
        !(leaf > 0) || (n && (!pts || !out_pts)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_partition_columns: bad argument (stride must be a multiple of 16, at most %d parts)",
                        PCP_MAX_PARTS);
    for (int p = 0; p < nparts; p++) counts[p] = 0;
    if (n == 0) return PCP_OK;
    if (n >= ((size_t)1 << 30)) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_partition_columns: 2^30 or more points");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    OutStage o;
    PCP_TRY(in.open(ctx, pts, n * stride));
    PCP_TRY(o.open(ctx, out_pts, n * stride));
    PartParams P;
    P.axis = axis, P.nparts = nparts, P.centre = centre, P.inv_leaf = 1.0 / (double)leaf, P.min_b = min_b;
    for (int p = 0; p <= nparts; p++) P.bounds[p] = bounds[p];
    DevBuf<uint32_t> k0, k1, v0, v1;
    DevBuf<unsigned long long> first;
    PCP_TRY(k0.alloc(ctx, n));
    PCP_TRY(k1.alloc(ctx, n));
    PCP_TRY(v0.alloc(ctx, n));
    PCP_TRY(v1.alloc(ctx, n));
    PCP_TRY(first.alloc(ctx, (size_t)nparts + 2));
    StageTimer t(ctx, "shard.partition", 2.0 * stride * n);
    part_key_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(in.dptr), n, stride, P, k0.p);
    PCP_LAUNCH_CHECK(ctx);
    uint32_t *keys = k0.p, *keys_alt = k1.p, *vals = v0.p, *vals_alt = v1.p;
    int bits = 1;
    while ((1 << bits) < nparts + 1) bits++;
    PCP_TRY(pcp_sort_pairs_u32(ctx, &keys, &keys_alt, &vals, &vals_alt, n, bits, true)); // stable: input order kept inside a part
    part_count_kernel<<<pcp_blocks_for(n + 1, 256), 256, 0, ctx->stream>>>(keys, n, nparts, first.p);
    PCP_LAUNCH_CHECK(ctx);
    part_gather_kernel<<<pcp_blocks_for(n * (stride >> 4), 256), 256, 0, ctx->stream>>>(static_cast<const char*>(in.dptr), stride, vals,
                                                                                       n, static_cast<char*>(o.dptr));
    PCP_LAUNCH_CHECK(ctx);
    unsigned long long hf[PCP_MAX_PARTS + 2];
    PCP_TRY(pcp_readback(ctx, hf, first.p, ((size_t)nparts + 2) * 8));
    for (int p = 0; p < nparts; p++) counts[p] = hf[p + 1] - hf[p];
    PCP_TRY(o.commit((size_t)hf[nparts] * stride));
    return PCP_OK;
}

// ... and the other way: a few host bytes into device memory without a place in the copy engine's queue
extern "C" int pcp_poke(pcp_ctx* ctx, void* dev_dst, size_t bytes, const void* host_src)
{
    if (!ctx || !dev_dst || !host_src || bytes > PCP_READBACK_MAX) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_poke: bad argument (at most %d bytes)", PCP_READBACK_MAX);
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return pcp_writeto(ctx, dev_dst, host_src, bytes);
}

// A few bytes of device memory read on the host, ordered behind the work submitted to the context so far, without
// queueing behind a bulk download on a copy lane (a kernel writes them into mapped host memory).
extern "C" int pcp_peek(pcp_ctx* ctx, const void* dev_src, size_t bytes, void* host_dst)
{
    if (!ctx || !dev_src || !host_dst || bytes > PCP_READBACK_MAX) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_peek: bad argument (at most %d bytes)", PCP_READBACK_MAX);
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return pcp_readback(ctx, host_dst, dev_src, bytes);
}
