// plane_fit.cu -- CalculateFeature::calculate_plan_parameter_h_points (calculate_feature.cpp:119-206) over
// many point segments at once (one thread per segment; points of a segment are consecutive records).
#include "common.cuh"
#include "plane_fit.cuh"

namespace {
__global__ void __launch_bounds__(128) plane_fit_kernel(const char* __restrict__ pts, size_t stride,
                                                        const long long* __restrict__ off, size_t nseg,
                                                        pcp_plan_segment* __restrict__ out)
{
    size_t s = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (s >= nseg) return;
    const long long b = off[s], e = off[s + 1];
    const long long h = e - b;
    pcp_plan_segment o;
    memset(&o, 0, sizeof(o));
    if (h > 0) {
        double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
        for (long long i = b; i < e; i++) {
            const double* p = reinterpret_cast<const double*>(pts + (size_t)i * stride);
            XAve += p[0];
            YAve += p[1];
            ZAve += p[2];
        }
        XAve /= (double)h;
        YAve /= (double)h;
        ZAve /= (double)h;
        double s00 = 0, s01 = 0, s02 = 0, s11 = 0, s12 = 0, s22 = 0;
        for (long long i = b; i < e; i++) {
            const double* p = reinterpret_cast<const double*>(pts + (size_t)i * stride);
            const double X = p[0] - XAve, Y = p[1] - YAve, Z = p[2] - ZAve;
            s00 += X * X;
            s01 += X * Y;
            s02 += X * Z;
            s11 += Y * Y;
            s12 += Y * Z;
            s22 += Z * Z;
        }
        pcp_plane pl;
        pcp_plane_from_scatter(s00, s01, s02, s11, s12, s22, XAve, YAve, ZAve, pl);
        o.normal_x = pl.nx;
        o.normal_y = pl.ny;
        o.normal_z = pl.nz;
        o.min_value = pl.min_value;
        o.curvature = pl.curvature;
        o.Distance = pl.Distance;
        o.eig[0] = pl.l1;
        o.eig[1] = pl.l2;
        o.eig[2] = pl.l3;
    }
    out[s] = o;
}
} // namespace

extern "C" int pcp_plane_fit(pcp_ctx* ctx, const void* pts, size_t stride, const int64_t* seg_offsets, size_t nseg,
                             pcp_plan_segment* out)
{
    if (!ctx || stride < 24 || (nseg && (!pts || !seg_offsets || !out)))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_plane_fit: bad argument");
    if (nseg == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage off;
    PCP_TRY(off.open(ctx, seg_offsets, (nseg + 1) * 8));
    // the number of points is the last offset
    long long total = 0;
    PCP_CUDA_TRY(ctx, cudaMemcpyAsync(&total, (const long long*)off.dptr + nseg, 8, cudaMemcpyDefault, ctx->stream));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (total < 0) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_plane_fit: negative offset");
    InStage in;
    OutStage o;
    PCP_TRY(in.open(ctx, pts, (size_t)total * stride));
    PCP_TRY(o.open(ctx, out, nseg * sizeof(pcp_plan_segment)));
    {
        StageTimer t(ctx, "plane_fit.kernel", 24.0 * total);
        plane_fit_kernel<<<pcp_blocks_for(nseg, 128), 128, 0, ctx->stream>>>(
            static_cast<const char*>(in.dptr), stride, (const long long*)off.dptr, nseg, (pcp_plan_segment*)o.dptr);
        PCP_LAUNCH_CHECK(ctx);
    }
    PCP_TRY(o.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}
