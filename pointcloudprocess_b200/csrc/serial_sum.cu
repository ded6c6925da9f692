// serial_sum.cu -- left-to-right double accumulation of x, y, z over the finite points of a cloud,
// bit-identical to the reference's serial loop in compute3DCentroid (point_cloud_helper.h:203-226).
//
// Floating-point addition is not associative, so a tree reduction gives a centroid that differs from the
// reference's in the last bits -- enough to move ~1e-5 of the points across a voxel boundary after
// remove_duplicate's centring (point_cloud_helper.cpp:44-52).  This kernel therefore keeps the serial
// ORDER of additions.  v1: one warp per axis; lanes fetch 32 consecutive points (coalesced), then the
// running sum is advanced through the 32 values by warp-shuffle broadcast (every lane keeps the same sum).
#include "reduce.cuh"

namespace {

__global__ void __launch_bounds__(96) serial_sum_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                        double* __restrict__ out3)
{
    const int axis = threadIdx.x >> 5; // warp 0,1,2 -> x,y,z
    const int lane = threadIdx.x & 31;
    double s = 0.0;
    // software pipeline: fetch batch b+1 while batch b is folded in
    size_t i = lane;
    double x = 0, y = 0, z = 0;
    bool have = i < n;
    if (have) {
        const double* p = reinterpret_cast<const double*>(pts + i * stride);
        x = p[0], y = p[1], z = p[2];
    }
    for (size_t base = 0; base < n; base += 32) {
        double cx = x, cy = y, cz = z;
        bool chave = have;
        size_t ni = base + 32 + lane;
        have = ni < n;
        if (have) {
            const double* p = reinterpret_cast<const double*>(pts + ni * stride);
            x = p[0], y = p[1], z = p[2];
        }
        bool ok = chave && isfinite(cx) && isfinite(cy) && isfinite(cz);
        double v = axis == 0 ? cx : (axis == 1 ? cy : cz);
        unsigned okmask = __ballot_sync(0xffffffffu, ok);
        if (okmask == 0xffffffffu) {
#pragma unroll
            for (int j = 0; j < 32; j++) s = __dadd_rn(s, __shfl_sync(0xffffffffu, v, j));
        } else {
            while (okmask) {
                int j = __ffs(okmask) - 1;
                okmask &= okmask - 1;
                s = __dadd_rn(s, __shfl_sync(0xffffffffu, v, j));
            }
        }
    }
    if (lane == 0) out3[axis] = s;
}

} // namespace

int pcp_serial_sum_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, double* d_sum3)
{
    serial_sum_kernel<<<1, 96, 0, ctx->stream>>>(static_cast<const char*>(dpts), n, stride, d_sum3);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}
