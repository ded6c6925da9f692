// reduce.cu -- bounding box / valid count over a strided cloud.
// getMinMax3D: point_cloud_helper.h:59-90 (min seeded DBL_MAX, max seeded DBL_MIN -- sic).
#include "reduce.cuh"
#include <cfloat>

namespace {

__device__ __forceinline__ double warp_min(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_max(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// One partial record per block: mn[3], mx[3], count.
__global__ void __launch_bounds__(256) bbox_partial_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                           double* __restrict__ partial)
{
    double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX};
    double mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    unsigned long long cnt = 0;
    double sm[3] = {0.0, 0.0, 0.0};
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const double* p = reinterpret_cast<const double*>(pts + i * stride);
        double x = p[0], y = p[1], z = p[2];
        if (isfinite(x) && isfinite(y) && isfinite(z)) {
            mn[0] = fmin(mn[0], x);
            mn[1] = fmin(mn[1], y);
            mn[2] = fmin(mn[2], z);
            mx[0] = fmax(mx[0], x);
            mx[1] = fmax(mx[1], y);
            mx[2] = fmax(mx[2], z);
            sm[0] += x, sm[1] += y, sm[2] += z;
            cnt++;
        }
    }
    __shared__ double s[8][10];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int d = 0; d < 3; d++) {
        mn[d] = warp_min(mn[d]);
        mx[d] = warp_max(mx[d]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
#pragma unroll
        for (int d = 0; d < 3; d++) sm[d] += __shfl_xor_sync(0xffffffffu, sm[d], o);
    }
    if (lane == 0) {
        for (int d = 0; d < 3; d++) s[w][d] = mn[d], s[w][3 + d] = mx[d], s[w][7 + d] = sm[d];
        s[w][6] = (double)cnt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double r[10];
        for (int d = 0; d < 10; d++) r[d] = s[0][d];
        for (int k = 1; k < 8; k++) {
            for (int d = 0; d < 3; d++) r[d] = fmin(r[d], s[k][d]), r[3 + d] = fmax(r[3 + d], s[k][3 + d]), r[7 + d] += s[k][7 + d];
            r[6] += s[k][6];
        }
        for (int d = 0; d < 10; d++) partial[blockIdx.x * 10 + d] = r[d];
    }
}

__global__ void bbox_final_kernel(const double* __restrict__ partial, int nblocks, pcp_bbox* __restrict__ out)
{
    // single warp
    double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    double cnt = 0, sm[3] = {0.0, 0.0, 0.0};
    for (int b = threadIdx.x; b < nblocks; b += 32) {
        for (int d = 0; d < 3; d++) {
            mn[d] = fmin(mn[d], partial[b * 10 + d]);
            mx[d] = fmax(mx[d], partial[b * 10 + 3 + d]);
            sm[d] += partial[b * 10 + 7 + d];
        }
        cnt += partial[b * 10 + 6]; // counts < 2^53: exact in double
    }
    for (int d = 0; d < 3; d++) mn[d] = warp_min(mn[d]), mx[d] = warp_max(mx[d]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
#pragma unroll
        for (int d = 0; d < 3; d++) sm[d] += __shfl_xor_sync(0xffffffffu, sm[d], o);
    }
    if (threadIdx.x == 0) {
        for (int d = 0; d < 3; d++) out->mn[d] = mn[d], out->mx[d] = mx[d], out->sum[d] = sm[d];
        out->count = (unsigned long long)cnt;
    }
}

} // namespace

int pcp_bbox_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, pcp_bbox* d_out)
{
    int nblocks = ctx->sm_count * 8;
    if ((size_t)nblocks * 256 > n) nblocks = (int)((n + 255) / 256);
    if (nblocks < 1) nblocks = 1;
    DevBuf<double> partial;
    PCP_TRY(partial.alloc(ctx, (size_t)nblocks * 10));
    bbox_partial_kernel<<<nblocks, 256, 0, ctx->stream>>>(static_cast<const char*>(dpts), n, stride, partial.p);
    PCP_LAUNCH_CHECK(ctx);
    bbox_final_kernel<<<1, 32, 0, ctx->stream>>>(partial.p, nblocks, d_out);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

int pcp_bbox_host(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, pcp_bbox* h_out)
{
    DevBuf<pcp_bbox> d;
    PCP_TRY(d.alloc(ctx, 1));
    PCP_TRY(pcp_bbox_device(ctx, dpts, n, stride, d.p));
    PCP_TRY(pcp_readback(ctx, h_out, d.p, sizeof(pcp_bbox)));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

extern "C" int pcp_minmax(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double mn[3], double mx[3])
{
    if (!ctx || !mn || !mx || stride < 24) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_minmax: bad argument");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    PCP_TRY(in.open(ctx, pts, n * stride));
    pcp_bbox bb;
    PCP_TRY(pcp_bbox_host(ctx, in.dptr, n, stride, &bb));
    for (int d = 0; d < 3; d++) {
        // min stays DBL_MAX on an empty cloud; max is seeded with DBL_MIN (smallest positive) in the
        // reference, so an all-negative axis reports 2.2e-308 (point_cloud_helper.h:65)
        mn[d] = bb.count ? bb.mn[d] : DBL_MAX;
        mx[d] = bb.count ? fmax(bb.mx[d], DBL_MIN) : DBL_MIN;
    }
    return PCP_OK;
}


// mn[3], mx[3] (true extremes over the finite points, no DBL_MIN seed), tree-order sums[3] (a prediction for
// pcp_serial_sum_prepare, not a result), count -- one pass; out10 may be a host or a device array of 10 doubles
extern "C" int pcp_cloud_stats(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double* out10)
{
    if (!ctx || !out10 || stride < 24 || (n && !pts)) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_cloud_stats: bad argument");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    PCP_TRY(in.open(ctx, pts, n * stride));
    pcp_bbox bb;
    {
        StageTimer t(ctx, "shard.stats", 24.0 * n);
        PCP_TRY(pcp_bbox_host(ctx, in.dptr, n, stride, &bb));
    }
    double h[10];
    for (int d = 0; d < 3; d++) h[d] = bb.mn[d], h[3 + d] = bb.mx[d], h[6 + d] = bb.sum[d];
    h[9] = (double)bb.count;
    if (pcp_classify(ctx, out10) == PCP_PTR_DEVICE) {
        PCP_TRY(pcp_writeto(ctx, out10, h, sizeof(h)));
    } else {
        memcpy(out10, h, sizeof(h));
    }
    return PCP_OK;
}
