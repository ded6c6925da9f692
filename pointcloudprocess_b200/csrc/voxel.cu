// voxel.cu -- VoxelGrid<PointXYZRGBA>::applyFilter (voxel_grid.h:811-1056) as sort + segmented reduce,
// and PointCloudHelper::remove_duplicate (point_cloud_helper.cpp:42-63) around it.
//
// Numerics kept from the reference: cell index by truncation of double(p*inv_leaf) - min_b
// (voxel_grid.h:936-938), min_b/max_b by truncation of the bounding box (:835-840) with the DBL_MIN max
// seed (point_cloud_helper.h:65), every field rounded through float before the double sum
// (concatenate.h:153), division by the count (:1034), static_cast write-back (concatenate.h:118) and the
// rgb re-pack (:1045-1050).  Cell keys are 64-bit (the reference's int idx overflows on large grids).
// Inside a cell points are summed in ascending point index (stable LSD sort of (key, index)).
#include "reduce.cuh"
#include "sort.cuh"
#include <cfloat>

namespace {

struct VoxParams {
    double inv[3];
    int min_b[3];
    int div_b[3];
    // remove_duplicate (point_cloud_helper.cpp:42-63) filters the cloud translated by -centroid and translates the
    // result back: with use_shift every coordinate is read as fl(p + shift) and every output written as
    // fl(centroid + unshift) -- the two transformPointCloud passes without a copy of the cloud
    int use_shift;
    double shift[3], unshift[3];
};

template <typename KeyT>
__global__ void __launch_bounds__(256) voxel_key_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                        VoxParams P, KeyT* __restrict__ keys,
                                                        uint32_t* __restrict__ vals, int32_t* __restrict__ point_cell)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* p = reinterpret_cast<const double*>(pts + i * stride);
    double x = p[0], y = p[1], z = p[2];
    KeyT key = ~(KeyT)0; // dropped points sort to the end
    int c0 = -1, c1 = -1, c2 = -1;
    if (isfinite(x) && isfinite(y) && isfinite(z)) {
        if (P.use_shift) x = __dadd_rn(x, P.shift[0]), y = __dadd_rn(y, P.shift[1]), z = __dadd_rn(z, P.shift[2]);
        c0 = (int)__dsub_rn(__dmul_rn(x, P.inv[0]), (double)P.min_b[0]);
        c1 = (int)__dsub_rn(__dmul_rn(y, P.inv[1]), (double)P.min_b[1]);
        c2 = (int)__dsub_rn(__dmul_rn(z, P.inv[2]), (double)P.min_b[2]);
        unsigned long long k = (unsigned long long)c0 + (unsigned long long)c1 * (unsigned long long)P.div_b[0] +
                               (unsigned long long)c2 * (unsigned long long)P.div_b[0] * (unsigned long long)P.div_b[1];
        key = (KeyT)k;
    }
    keys[i] = key;
    if (vals) vals[i] = (uint32_t)i;
    if (point_cell) {
        point_cell[3 * i + 0] = c0;
        point_cell[3 * i + 1] = c1;
        point_cell[3 * i + 2] = c2;
    }
}

template <typename KeyT>
__global__ void __launch_bounds__(256) head_flag_kernel(const KeyT* __restrict__ keys, size_t nv,
                                                        uint32_t* __restrict__ flag)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= nv) return;
    flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}

// scan[] is the exclusive prefix sum of flag[]: a head at sorted position i opens cell scan[i].
__global__ void __launch_bounds__(256) cell_start_kernel(const uint32_t* __restrict__ flag,
                                                         const uint32_t* __restrict__ scan, size_t nv,
                                                         uint32_t* __restrict__ cell_start, uint32_t* __restrict__ m_out)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= nv) return;
    if (flag[i]) cell_start[scan[i]] = (uint32_t)i;
    if (i == nv - 1) {
        uint32_t m = scan[i] + flag[i];
        cell_start[m] = (uint32_t)nv;
        *m_out = m;
    }
}

// One thread per occupied cell (N/M is ~1 for the reference's leaf sizes).
__global__ void __launch_bounds__(128) voxel_centroid_kernel(const char* __restrict__ pts, size_t stride,
                                                             const uint32_t* __restrict__ vals,
                                                             const uint32_t* __restrict__ cell_start,
                                                             const uint32_t* __restrict__ m_ptr, int downsample_all,
                                                             char* __restrict__ out, size_t out_stride,
                                                             int32_t* __restrict__ cell_count, const VoxParams P, int vec)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= *m_ptr) return;
    uint32_t b = cell_start[c], e = cell_start[c + 1];
    double s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // vec & 1: packed 32-byte input records on a 32-byte boundary -- a record is one 256-bit load; vec & 2: the same for
    // the output and one store
    for (uint32_t t = b; t < e; t++) {
        const char* rec = pts + (size_t)vals[t] * stride;
        double p[3];
        uint32_t rgba_in = 0, stamp_in = 0;
        if (vec & 1) {
            const pcp_d4 v = pcp_ld256(rec);
            p[0] = v.a, p[1] = v.b, p[2] = v.c;
            const unsigned long long w = (unsigned long long)__double_as_longlong(v.d);
            rgba_in = (uint32_t)w, stamp_in = (uint32_t)(w >> 32);
        } else {
            const double* pp = reinterpret_cast<const double*>(rec);
            p[0] = pp[0], p[1] = pp[1], p[2] = pp[2];
        }
        if (P.use_shift) p[0] = __dadd_rn(p[0], P.shift[0]), p[1] = __dadd_rn(p[1], P.shift[1]), p[2] = __dadd_rn(p[2], P.shift[2]);
        double v[8];
        if (downsample_all) {
            // PointXYZRGBA keeps rgba / stamp_id at bytes 32 / 36; the packed 32-byte record (a PCD binary payload
            // record: x, y, z, rgba, stamp_id) at 24 / 28
            const size_t ao = stride == 32 ? 24 : 32;
            uint32_t rgba = rgba_in, stamp = stamp_in;
            if (!(vec & 1)) {
                rgba = *reinterpret_cast<const uint32_t*>(rec + ao);
                stamp = *reinterpret_cast<const uint32_t*>(rec + ao + 4);
            }
            v[0] = (double)(float)p[0];
            v[1] = (double)(float)p[1];
            v[2] = (double)(float)p[2];
            v[3] = (double)(float)rgba;
            v[4] = (double)(float)stamp;
            v[5] = (double)((rgba >> 16) & 0xffu);
            v[6] = (double)((rgba >> 8) & 0xffu);
            v[7] = (double)(rgba & 0xffu);
        } else {
            v[0] = p[0], v[1] = p[1], v[2] = p[2];
            v[3] = v[4] = v[5] = v[6] = v[7] = 0;
        }
        if (t == b) {
#pragma unroll
            for (int f = 0; f < 8; f++) s[f] = v[f];
        } else {
#pragma unroll
            for (int f = 0; f < 8; f++) s[f] = __dadd_rn(s[f], v[f]);
        }
    }
    double cnt = (double)(e - b);
#pragma unroll
    for (int f = 0; f < 8; f++) s[f] = __ddiv_rn(s[f], cnt);
    char* o = out + (size_t)c * out_stride;
    double* od = reinterpret_cast<double*>(o);
    if (P.use_shift && isfinite(s[0]) && isfinite(s[1]) && isfinite(s[2])) // transformPointCloud skips the others
        s[0] = __dadd_rn(s[0], P.unshift[0]), s[1] = __dadd_rn(s[1], P.unshift[1]), s[2] = __dadd_rn(s[2], P.unshift[2]);
    uint32_t rgb = 0, stamp = 0;
    if (downsample_all) {
        float r = (float)s[5], g = (float)s[6], bl = (float)s[7];
        rgb = (uint32_t)(((int)r << 16) | ((int)g << 8) | (int)bl);
        stamp = (uint32_t)s[4]; // cvt.rzi.u32.f64 == static_cast<uint32_t> for in-range values
    }
    if ((vec & 2) && downsample_all) { // packed record on a 32-byte boundary: one store
        pcp_st256(o, pcp_d4{s[0], s[1], s[2], __longlong_as_double((long long)((unsigned long long)rgb | ((unsigned long long)stamp << 32)))});
        if (cell_count) cell_count[c] = (int32_t)(e - b);
        return;
    }
    od[0] = s[0];
    od[1] = s[1];
    od[2] = s[2];
    if (out_stride == 32 && downsample_all) { // packed record: no w, no padding
        *reinterpret_cast<uint32_t*>(o + 24) = rgb;
        *reinterpret_cast<uint32_t*>(o + 28) = stamp;
    } else {
        if (out_stride >= 32) od[3] = 1.0;
        if (out_stride >= 40) {
            *reinterpret_cast<uint32_t*>(o + 32) = rgb;
            *reinterpret_cast<uint32_t*>(o + 36) = stamp;
        }
        if (out_stride >= 48) *reinterpret_cast<uint64_t*>(o + 40) = 0ull;
    }
    if (cell_count) cell_count[c] = (int32_t)(e - b);
}

int ceil_log2_u128(double cells)
{
    int b = 0;
    double v = 1.0;
    while (v < cells && b < 64) v *= 2.0, b++;
    return b;
}

} // namespace

// Device-resident core.  dpts: device cloud; dout: device output (capacity n records).
// m_host receives M (synchronises).  d_point_cell / d_cell_count may be NULL.
int pcp_voxel_grid_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, const double leaf[3],
                          int downsample_all, void* dout, size_t out_stride, size_t* m_host, int32_t min_b[3],
                          int32_t div_b[3], int32_t* d_point_cell, int32_t* d_cell_count,
                          const int32_t* force_min_b = nullptr, const int32_t* force_div_b = nullptr,
                          const double* centre3 = nullptr, const pcp_bbox* known_bb = nullptr)
{
    *m_host = 0;
    if (n == 0) return PCP_OK;
    if (n > (size_t)0x7fffffff) return pcp_fail(ctx, PCP_ERR_RANGE, "voxel grid: more than 2^31-1 points");
    pcp_bbox bb;
    VoxParams P;
    P.use_shift = 0;
    for (int d = 0; d < 3; d++) P.shift[d] = P.unshift[d] = 0.0;
    if (known_bb) {
        bb = *known_bb;
    } else {
        StageTimer t(ctx, "voxel.bbox", 24.0 * n);
        PCP_TRY(pcp_bbox_host(ctx, dpts, n, stride, &bb));
    }
    if (bb.count == 0) return PCP_OK;
    if (centre3) {
        // the box of the translated cloud: rounding is monotone, so min fl(x + t) = fl(min x + t)
        P.use_shift = 1;
        for (int d = 0; d < 3; d++) {
            P.shift[d] = -centre3[d];
            P.unshift[d] = centre3[d];
            bb.mn[d] = bb.mn[d] + P.shift[d];
            bb.mx[d] = bb.mx[d] + P.shift[d];
        }
    }
    double cells = 1.0;
    for (int d = 0; d < 3; d++) {
        if (!(leaf[d] > 0)) return pcp_fail(ctx, PCP_ERR_INVALID, "voxel grid: leaf size must be positive");
        P.inv[d] = 1.0 / leaf[d];
        double mx = bb.mx[d] > DBL_MIN ? bb.mx[d] : DBL_MIN; // sic, point_cloud_helper.h:65
        double lo = bb.mn[d] * P.inv[d], hi = mx * P.inv[d];
        if (!(fabs(lo) < 2147483000.0) || !(fabs(hi) < 2147483000.0))
            return pcp_fail(ctx, PCP_ERR_RANGE, "voxel grid: leaf too small, cell index exceeds int range");
        P.min_b[d] = (int)lo;
        int max_b = (int)hi;
        P.div_b[d] = max_b - P.min_b[d] + 1;
        if (force_min_b && force_div_b) { // a shard of a larger grid: the whole cloud's box (same cells, same order)
            if (P.min_b[d] < force_min_b[d] || max_b > force_min_b[d] + force_div_b[d] - 1)
                return pcp_fail(ctx, PCP_ERR_RANGE, "voxel grid: shard exceeds the imposed grid box");
            P.min_b[d] = force_min_b[d];
            P.div_b[d] = force_div_b[d];
        }
        cells *= (double)P.div_b[d];
        if (min_b) min_b[d] = P.min_b[d];
        if (div_b) div_b[d] = P.div_b[d];
    }
    int bits = ceil_log2_u128(cells);
    if (bits > 63) return pcp_fail(ctx, PCP_ERR_RANGE, "voxel grid: more than 2^63 cells");
    size_t nv = (size_t)bb.count;
    DevBuf<uint32_t> v0, v1, flag, scan, cell_start, m_dev;
    PCP_TRY(v0.alloc(ctx, n));
    PCP_TRY(v1.alloc(ctx, n));
    uint32_t *vals = v0.p, *vals_alt = v1.p;
    unsigned nb = pcp_blocks_for(n, 256);
    PCP_TRY(flag.alloc(ctx, nv));
    PCP_TRY(scan.alloc(ctx, nv));
    // dropped (non-finite) points carry an all-ones key: sort one extra bit when there are any
    int sort_bits = (nv < n) ? (bits < 63 ? bits + 1 : 64) : bits;
    if (sort_bits <= 32 && !(nv < n && bits == 32)) {
        DevBuf<uint32_t> k0, k1;
        PCP_TRY(k0.alloc(ctx, n));
        PCP_TRY(k1.alloc(ctx, n));
        uint32_t *keys = k0.p, *keys_alt = k1.p;
        {
            StageTimer t(ctx, "voxel.keys", 24.0 * n);
            voxel_key_kernel<uint32_t><<<nb, 256, 0, ctx->stream>>>(static_cast<const char*>(dpts), n, stride, P, keys,
                                                                   nullptr, d_point_cell); // values: the sort's iota
            PCP_LAUNCH_CHECK(ctx);
        }
        {
            StageTimer t(ctx, "voxel.sort");
            PCP_TRY(pcp_sort_pairs_u32(ctx, &keys, &keys_alt, &vals, &vals_alt, n, nv < n ? 32 : sort_bits, true));
        }
        StageTimer t(ctx, "voxel.segments");
        head_flag_kernel<uint32_t><<<pcp_blocks_for(nv, 256), 256, 0, ctx->stream>>>(keys, nv, flag.p);
        PCP_LAUNCH_CHECK(ctx);
    } else {
        DevBuf<unsigned long long> k0, k1;
        PCP_TRY(k0.alloc(ctx, n));
        PCP_TRY(k1.alloc(ctx, n));
        unsigned long long *keys = k0.p, *keys_alt = k1.p;
        {
            StageTimer t(ctx, "voxel.keys", 24.0 * n);
            voxel_key_kernel<unsigned long long><<<nb, 256, 0, ctx->stream>>>(static_cast<const char*>(dpts), n, stride, P, keys,
                                                                   nullptr, d_point_cell);
            PCP_LAUNCH_CHECK(ctx);
        }
        {
            StageTimer t(ctx, "voxel.sort");
            PCP_TRY(pcp_sort_pairs_u64(ctx, &keys, &keys_alt, &vals, &vals_alt, n, nv < n ? 64 : sort_bits, true));
        }
        StageTimer t(ctx, "voxel.segments");
        head_flag_kernel<unsigned long long><<<pcp_blocks_for(nv, 256), 256, 0, ctx->stream>>>(keys, nv, flag.p);
        PCP_LAUNCH_CHECK(ctx);
    }
    {
        StageTimer t(ctx, "voxel.segments");
        PCP_TRY(pcp_exclusive_scan_u32(ctx, flag.p, scan.p, nv));
        PCP_TRY(cell_start.alloc(ctx, nv + 1));
        PCP_TRY(m_dev.alloc(ctx, 1));
        cell_start_kernel<<<pcp_blocks_for(nv, 256), 256, 0, ctx->stream>>>(flag.p, scan.p, nv, cell_start.p, m_dev.p);
        PCP_LAUNCH_CHECK(ctx);
    }
    {
        StageTimer t(ctx, "voxel.centroids", 32.0 * n);
        // grid sized for the worst case M == nv; threads beyond *m_ptr exit
        voxel_centroid_kernel<<<pcp_blocks_for(nv, 128), 128, 0, ctx->stream>>>(
            static_cast<const char*>(dpts), stride, vals, cell_start.p, m_dev.p, downsample_all,
            static_cast<char*>(dout), out_stride, d_cell_count, P,
            ((stride == 32 && ((uintptr_t)dpts & 31) == 0) ? 1 : 0) | ((out_stride == 32 && ((uintptr_t)dout & 31) == 0) ? 2 : 0));
        PCP_LAUNCH_CHECK(ctx);
    }
    uint32_t m = 0;
    PCP_TRY(pcp_readback(ctx, &m, m_dev.p, 4));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    *m_host = m;
    return PCP_OK;
}

extern "C" int pcp_voxel_grid(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, const double leaf[3],
                              int downsample_all, void* out_pts, size_t* m, int32_t min_b[3], int32_t div_b[3],
                              int32_t* point_cell, int32_t* cell_count)
{
    if (!ctx || !m || !leaf || stride < 24 || (n && (!pts || !out_pts)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_voxel_grid: bad argument");
    if (downsample_all && stride < 32)
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_voxel_grid: downsample_all needs rgba/stamp_id (stride 32 packed, or >= 40)");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    OutStage out, pc, cc;
    PCP_TRY(in.open(ctx, pts, n * stride));
    PCP_TRY(out.open(ctx, out_pts, n * stride));
    PCP_TRY(pc.open(ctx, point_cell, n * 12));
    PCP_TRY(cc.open(ctx, cell_count, n * 4));
    size_t mm = 0;
    PCP_TRY(pcp_voxel_grid_device(ctx, in.dptr, n, stride, leaf, downsample_all, out.dptr, stride, &mm, min_b, div_b,
                                  static_cast<int32_t*>(pc.dptr), static_cast<int32_t*>(cc.dptr)));
    *m = mm;
    PCP_TRY(out.commit(mm * stride));
    PCP_TRY(pc.commit());
    PCP_TRY(cc.commit(mm * 4));
    return PCP_OK;
}

extern "C" int pcp_centroid(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double c[3], uint64_t* count)
{
    if (!ctx || !c || stride < 24) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_centroid: bad argument");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    PCP_TRY(in.open(ctx, pts, n * stride));
    DevBuf<double> s3;
    PCP_TRY(s3.alloc(ctx, 3));
    pcp_bbox bb;
    PCP_TRY(pcp_serial_sum_device(ctx, in.dptr, n, stride, s3.p));
    PCP_TRY(pcp_bbox_host(ctx, in.dptr, n, stride, &bb));
    double h[3];
    PCP_TRY(pcp_readback(ctx, h, s3.p, 24));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    double dn = (double)bb.count;
    for (int d = 0; d < 3; d++) c[d] = bb.count ? h[d] / dn : 0.0;
    if (count) *count = bb.count;
    return PCP_OK;
}

// centre == NULL: the cloud's own serial-order centroid (the reference behaviour); otherwise the given centre
// (a shard of a larger cloud is centred on the centroid of the WHOLE cloud so that voxels line up across shards).
// pts_out == NULL: in place (pts holds *n_inout records on entry and the filtered cloud on return); otherwise the
// input is left alone and the filtered cloud goes to pts_out (capacity *n_inout records): no copy of the cloud at all.
static int remove_duplicate_impl(pcp_ctx* ctx, void* pts, size_t* n_inout, size_t stride, float leaf, const double* centre,
                                 const int32_t* min_b = nullptr, const int32_t* div_b = nullptr, void* pts_out = nullptr)
{
    size_t n = *n_inout;
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    OutStage io; // in-place: host clouds are staged in, filtered on the device, staged back
    InStage in;
    OutStage out;
    DevBuf<char> tmp; // in place: the filtered cloud, copied over the input at the end (the gather must not overwrite its source)
    const char* d = nullptr;
    char* dst = nullptr;
    if (pts_out) {
        PCP_TRY(in.open(ctx, pts, n * stride));
        PCP_TRY(out.open(ctx, pts_out, n * stride));
        d = static_cast<const char*>(in.dptr);
        dst = static_cast<char*>(out.dptr);
    } else {
        PCP_TRY(io.open(ctx, pts, n * stride, true));
        PCP_TRY(tmp.alloc(ctx, n * stride));
        d = static_cast<const char*>(io.dptr);
        dst = tmp.p;
    }
    DevBuf<double> s3;
    DevBuf<pcp_bbox> bb;
    PCP_TRY(s3.alloc(ctx, 3));
    PCP_TRY(bb.alloc(ctx, 1));
    // the two transformPointCloud passes (point_cloud_helper.cpp:52-53, 60-61) are folded into the voxel kernels:
    // coordinates are read as fl(p - c) and the cell centroids written as fl(centroid + c)
    double c3[3];
    pcp_bbox hb;
    {
        StageTimer t(ctx, "dedup.centroid", 24.0 * n);
        if (!centre) PCP_TRY(pcp_serial_sum_device(ctx, d, n, stride, s3.p));
        PCP_TRY(pcp_bbox_device(ctx, d, n, stride, bb.p));
        double h3[3] = {0, 0, 0};
        if (!centre) PCP_TRY(pcp_readback(ctx, h3, s3.p, 24));
        PCP_TRY(pcp_readback(ctx, &hb, bb.p, sizeof(hb)));
        const double dn = (double)hb.count;
        for (int a = 0; a < 3; a++) c3[a] = centre ? centre[a] : h3[a] / dn; // compute3DCentroid: sum / count
    }
    double l = (double)leaf; // const float -> setLeafSize(double,double,double), point_cloud_helper.cpp:57
    double leaf3[3] = {l, l, l};
    size_t m = 0;
    PCP_TRY(pcp_voxel_grid_device(ctx, d, n, stride, leaf3, 1, dst, stride, &m, nullptr, nullptr, nullptr, nullptr,
                                  min_b, div_b, c3, &hb));
    *n_inout = m;
    if (pts_out) {
        PCP_TRY(out.commit(m * stride));
    } else {
        if (m) {
            StageTimer t(ctx, "dedup.copy_back", 96.0 * m);
            PCP_CUDA_TRY(ctx, cudaMemcpyAsync(io.dptr, tmp.p, m * stride, cudaMemcpyDeviceToDevice, ctx->stream));
        }
        PCP_TRY(io.commit(m * stride));
    }
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

static bool rd_stride_ok(size_t stride) { return stride == 32 || stride >= 40; }

extern "C" int pcp_remove_duplicate(pcp_ctx* ctx, void* pts_inout, size_t* n_inout, size_t stride, float leaf)
{
    if (!ctx || !n_inout || !rd_stride_ok(stride) || (*n_inout && !pts_inout))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_remove_duplicate: bad argument");
    return remove_duplicate_impl(ctx, pts_inout, n_inout, stride, leaf, nullptr);
}

extern "C" int pcp_remove_duplicate_centred(pcp_ctx* ctx, void* pts_inout, size_t* n_inout, size_t stride, float leaf,
                                            const double centre[3], const int32_t min_b[3], const int32_t div_b[3])
{
    if (!ctx || !n_inout || !centre || !min_b || !div_b || !rd_stride_ok(stride) || (*n_inout && !pts_inout))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_remove_duplicate_centred: bad argument");
    return remove_duplicate_impl(ctx, pts_inout, n_inout, stride, leaf, centre, min_b, div_b);
}

extern "C" int pcp_remove_duplicate_to(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, float leaf,
                                       const double* centre, const int32_t* min_b, const int32_t* div_b, void* pts_out,
                                       size_t* m)
{
    if (!ctx || !m || !rd_stride_ok(stride) || (n && (!pts || !pts_out)) || ((min_b || div_b) && !(centre && min_b && div_b)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_remove_duplicate_to: bad argument");
    *m = n;
    return remove_duplicate_impl(ctx, const_cast<void*>(pts), m, stride, leaf, centre, min_b, div_b, pts_out);
}
