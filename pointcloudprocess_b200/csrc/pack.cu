// pack.cu -- the packed forms of the two records that cross PCIe: the 32-byte cloud record (the binary PCD payload
// record: x, y, z doubles, rgba, stamp_id) and the 24-byte LAS_POINT_PROPERTY (include/pcp_cuda.h).  Lossless both
// ways; the 48-byte forms are the reference's structs (point_type.h:9-82, data_struct.h:161-173).
#include "common.cuh"
#include <thread>

// ---- packed records for the trip over PCIe -------------------------------------------------------------------------
// LAS_POINT_PROPERTY carries 24 bytes of information in 48: Distance and curvature are PlanSegment's FLOAT members
// widened to double (calculate_feature.cpp:344-345), SegmentID and dis_from_point_plane are not set by this path.
namespace {
__global__ void __launch_bounds__(256) props_pack_kernel(const uint4* __restrict__ props, size_t n,
                                                         pcp_point_property_packed* __restrict__ out)
{
    const size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint4 a = props[3 * j], b = props[3 * j + 1], c = props[3 * j + 2];
    pcp_point_property_packed o;
    o.normal_x = __uint_as_float(a.x), o.normal_y = __uint_as_float(a.y), o.normal_z = __uint_as_float(a.z);
    o.Distance = (float)__longlong_as_double((long long)(((unsigned long long)b.y << 32) | b.x)); // exact: a widened float
    o.curvature = (float)__longlong_as_double((long long)(((unsigned long long)b.w << 32) | b.z));
    o.PointID = (int32_t)c.x;
    out[j] = o;
}
} // namespace

extern "C" int pcp_props_pack(pcp_ctx* ctx, const void* props, size_t n, pcp_point_property_packed* packed)
{
    if (!ctx || (n && (!props || !packed))) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_props_pack: bad argument");
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    OutStage o;
    PCP_TRY(in.open(ctx, props, n * 48));
    PCP_TRY(o.open(ctx, packed, n * sizeof(pcp_point_property_packed)));
    {
        StageTimer t(ctx, "props.pack", 72.0 * n);
        props_pack_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const uint4*>(in.dptr), n,
                                                                         static_cast<pcp_point_property_packed*>(o.dptr));
        PCP_LAUNCH_CHECK(ctx);
    }
    PCP_TRY(o.commit());
    return PCP_OK; // device outputs: asynchronous on the context's stream
}

// host side: packed records -> LAS_POINT_PROPERTY (threads <= 0: one per core)
extern "C" int pcp_props_unpack(const pcp_point_property_packed* packed, size_t n, void* props, int threads)
{
    if (n && (!packed || !props)) return PCP_ERR_INVALID;
    unsigned nt = threads > 0 ? (unsigned)threads : std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (n < 65536) nt = 1;
    auto work = [&](size_t lo, size_t hi) {
        pcp_point_property* o = static_cast<pcp_point_property*>(props);
        for (size_t j = lo; j < hi; j++) {
            const pcp_point_property_packed p = packed[j];
            pcp_point_property q;
            memset(&q, 0, sizeof(q));
            q.normal_x = p.normal_x, q.normal_y = p.normal_y, q.normal_z = p.normal_z;
            q.Distance = (double)p.Distance, q.curvature = (double)p.curvature;
            q.PointID = p.PointID;
            o[j] = q;
        }
    };
    if (nt == 1) {
        work(0, n);
        return PCP_OK;
    }
    std::vector<std::thread> th;
    const size_t per = (n + nt - 1) / nt;
    for (unsigned t = 0; t < nt; t++) {
        const size_t lo = t * per, hi = lo + per < n ? lo + per : n;
        if (lo < hi) th.emplace_back(work, lo, hi);
    }
    for (auto& x : th) x.join();
    return PCP_OK;
}

// ---- PointXYZRGBA (48 bytes) <-> packed cloud record (32 bytes) ---------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) cloud_pack_kernel(const char* __restrict__ in, size_t n, char* __restrict__ out)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* s = reinterpret_cast<const uint4*>(in + i * 48);
    const uint4 a = s[0], b = s[1], c = s[2]; // x y | z w | rgba stamp pad
    uint4* d = reinterpret_cast<uint4*>(out + i * 32);
    d[0] = a;
    d[1] = make_uint4(b.x, b.y, c.x, c.y);
}
__global__ void __launch_bounds__(256) cloud_unpack_kernel(const char* __restrict__ in, size_t n, char* __restrict__ out)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* s = reinterpret_cast<const uint4*>(in + i * 32);
    const uint4 a = s[0], b = s[1];
    uint4* d = reinterpret_cast<uint4*>(out + i * 48);
    const unsigned long long one = (unsigned long long)__double_as_longlong(1.0); // w = 1 (point_type.h:23)
    d[0] = a;
    d[1] = make_uint4(b.x, b.y, (unsigned)one, (unsigned)(one >> 32));
    d[2] = make_uint4(b.z, b.w, 0u, 0u);
}
} // namespace

static int cloud_repack(pcp_ctx* ctx, const void* in, size_t n, void* out, bool pack)
{
    if (!ctx || (n && (!in || !out))) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_cloud_pack/unpack: bad argument");
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage i;
    OutStage o;
    PCP_TRY(i.open(ctx, in, n * (pack ? 48 : 32)));
    PCP_TRY(o.open(ctx, out, n * (pack ? 32 : 48)));
    StageTimer t(ctx, pack ? "cloud.pack" : "cloud.unpack", 80.0 * n);
    if (pack)
        cloud_pack_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(i.dptr), n, static_cast<char*>(o.dptr));
    else
        cloud_unpack_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(i.dptr), n, static_cast<char*>(o.dptr));
    PCP_LAUNCH_CHECK(ctx);
    PCP_TRY(o.commit());
    return PCP_OK;
}
extern "C" int pcp_cloud_pack(pcp_ctx* ctx, const void* pts48, size_t n, void* pts32) { return cloud_repack(ctx, pts48, n, pts32, true); }
extern "C" int pcp_cloud_unpack(pcp_ctx* ctx, const void* pts32, size_t n, void* pts48) { return cloud_repack(ctx, pts32, n, pts48, false); }
