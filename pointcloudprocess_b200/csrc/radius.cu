// radius.cu -- KdTreeFLANN::radiusSearch (kd_tree.h:863-903) and radius-neighbourhood eigen-features.
// Neighbours are the points with d2 < r*r (strict, FLANN's RadiusResultSet), sorted by (d2, index),
// capped at max_nn (0 or > valid points = unlimited, kd_tree.h:873-883).
// The ball of radius r is covered by the 3x3x3 block of level-L cells, L = first level with cell edge >= r.
#include "knn_warp.cuh"
#include "sort.cuh"
#include <cub/cub.cuh>
#include <cstdlib>

namespace {

struct RadParams {
    const pcp_rec* rec;
    const unsigned long long* ckey;
    const uint32_t* cstart;
    uint32_t ncell;
    pcp_grid g;
    int level;
    double r2;
    // queries: external points, or (qpts == NULL) the index's own sorted points
    const char* qpts;
    size_t qstride;
    size_t nq;
};

__device__ __forceinline__ void rad_query(const RadParams& P, size_t qi, double& qx, double& qy, double& qz)
{
    if (P.qpts) {
        const double* q = reinterpret_cast<const double*>(P.qpts + qi * P.qstride);
        qx = q[0], qy = q[1], qz = q[2];
    } else {
        const pcp_rec q = pcp_load_rec(P.rec, (uint32_t)qi);
        qx = q.x, qy = q.y, qz = q.z;
    }
}

// lanes 0..26 receive the point range of one neighbour cell each
__device__ __forceinline__ void rad_ranges(const RadParams& P, int lane, double qx, double qy, double qz, uint32_t& lo,
                                           uint32_t& hi)
{
    lo = hi = 0;
    const int cx = min(max(pcp_cell_coord(qx, P.g.org[0], P.g.inv_h, 0), 0), P.g.dims[0] - 1) >> P.level;
    const int cy = min(max(pcp_cell_coord(qy, P.g.org[1], P.g.inv_h, 0), 0), P.g.dims[1] - 1) >> P.level;
    const int cz = min(max(pcp_cell_coord(qz, P.g.org[2], P.g.inv_h, 0), 0), P.g.dims[2] - 1) >> P.level;
    if (lane < 27) {
        int dx, dy, dz;
        pcp_nb_offset(lane, dx, dy, dz);
        const int nx = cx + dx, ny = cy + dy, nz = cz + dz;
        const int mx = (P.g.dims[0] - 1) >> P.level, my = (P.g.dims[1] - 1) >> P.level, mz = (P.g.dims[2] - 1) >> P.level;
        if (nx >= 0 && ny >= 0 && nz >= 0 && nx <= mx && ny <= my && nz <= mz)
            pcp_level_range(P.ckey, P.cstart, P.ncell, P.level, nx, ny, nz, lo, hi);
    }
}

// mode 0: count;  mode 1: fill (d2, idx) at off[qi];  warp per query, grid-stride
template <int MODE>
__global__ void __launch_bounds__(128) radius_scan_kernel(RadParams P, unsigned long long* __restrict__ cnt,
                                                          const unsigned long long* __restrict__ off,
                                                          double* __restrict__ out_d2, int32_t* __restrict__ out_idx)
{
    const int lane = threadIdx.x & 31;
    const size_t warp0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t qi = warp0; qi < P.nq; qi += nwarp) {
        double qx, qy, qz;
        rad_query(P, qi, qx, qy, qz);
        unsigned long long total = 0;
        if (isfinite(qx) && isfinite(qy) && isfinite(qz)) {
            uint32_t lo, hi;
            rad_ranges(P, lane, qx, qy, qz, lo, hi);
            unsigned long long wbase = MODE ? off[qi] : 0ull;
            for (int nb = 0; nb < 27; nb++) {
                const uint32_t s0 = __shfl_sync(0xffffffffu, lo, nb);
                const uint32_t s1 = __shfl_sync(0xffffffffu, hi, nb);
                for (uint32_t t0 = s0; t0 < s1; t0 += 32) {
                    const uint32_t t = t0 + lane;
                    bool in = false;
                    double d2 = 0;
                    int ci = 0;
                    if (t < s1) {
                        const pcp_rec r = pcp_load_rec(P.rec, t);
                        d2 = pcp_dist2(qx, qy, qz, r.x, r.y, r.z);
                        ci = (int)r.idx;
                        in = d2 < P.r2;
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, in);
                    if (MODE && in) {
                        const unsigned long long w = wbase + total + __popc(m & ((1u << lane) - 1));
                        out_d2[w] = d2;
                        out_idx[w] = ci;
                    }
                    total += __popc(m);
                }
            }
        }
        if (!MODE && lane == 0) cnt[qi] = total;
    }
}

// Per-query moments about the query point (warp per query): n, sum d, sum d d^T  (d = p - q).
__global__ void __launch_bounds__(128) radius_moments_kernel(RadParams P, double* __restrict__ mom /*[nq][10]*/)
{
    const int lane = threadIdx.x & 31;
    const size_t warp0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t qi = warp0; qi < P.nq; qi += nwarp) {
        double qx, qy, qz;
        rad_query(P, qi, qx, qy, qz);
        double a[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t lo, hi;
        rad_ranges(P, lane, qx, qy, qz, lo, hi);
        for (int nb = 0; nb < 27; nb++) {
            const uint32_t s0 = __shfl_sync(0xffffffffu, lo, nb);
            const uint32_t s1 = __shfl_sync(0xffffffffu, hi, nb);
            for (uint32_t t = s0 + lane; t < s1; t += 32) {
                const pcp_rec r = pcp_load_rec(P.rec, t);
                const double d2 = pcp_dist2(qx, qy, qz, r.x, r.y, r.z);
                if (d2 < P.r2) {
                    const double dx = r.x - qx, dy = r.y - qy, dz = r.z - qz;
                    a[0] += 1.0;
                    a[1] += dx;
                    a[2] += dy;
                    a[3] += dz;
                    a[4] += dx * dx;
                    a[5] += dx * dy;
                    a[6] += dx * dz;
                    a[7] += dy * dy;
                    a[8] += dy * dz;
                    a[9] += dz * dz;
                }
            }
        }
#pragma unroll
        for (int f = 0; f < 10; f++)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a[f] += __shfl_xor_sync(0xffffffffu, a[f], o);
        if (lane < 10) {
            double v = a[0];
#pragma unroll
            for (int f = 1; f < 10; f++)
                if (lane == f) v = a[f];
            mom[qi * 10 + lane] = v;
        }
    }
}

// The same moments for the index's own points, cell-cooperative: a warp takes one base cell, lane j holds query j of
// the cell (chunks of 32).  The balls of all its queries lie inside the block of (2R+1)^3 level-L cells around the
// cell (R * edge >= r); the block's points are streamed once per chunk through shared memory as FP64 coordinates and
// every lane runs over them (broadcast reads), accumulating its query's moments under the predicate d2 < r2.
// Against the warp-per-query kernel: the candidates are loaded once per cell instead of once per query, and no
// lane idles on the 7/8 of a 3x3x3 block that lies outside the ball of a single query -- it works on its own query.
constexpr int RC_WARPS = 4;
constexpr int RC_MAXCELLS = 128; // >= (2R+1)^3, R <= 2
constexpr int RC_CHUNK = 192;    // candidates staged at a time

struct RadCellParams {
    RadParams R;
    const uint4* cinfo; // per base cell: x, y, z, first point
    int blockR;         // R
    uint32_t* work_counter;
};

__global__ void __launch_bounds__(32 * RC_WARPS) radius_moments_cell_kernel(RadCellParams P, double* __restrict__ mom)
{
    __shared__ uint32_t s_pre[RC_WARPS][RC_MAXCELLS]; // exclusive prefix of the block cells' point counts
    __shared__ uint32_t s_lo[RC_WARPS][RC_MAXCELLS];  // first point of each block cell
    __shared__ double s_xyz[RC_WARPS][3][RC_CHUNK];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* pre = s_pre[warp];
    uint32_t* los = s_lo[warp];
    double* sx = s_xyz[warp][0];
    double* sy = s_xyz[warp][1];
    double* sz = s_xyz[warp][2];
    const RadParams& Q = P.R;
    const int L = Q.level, R = P.blockR, W = 2 * R + 1, NB = W * W * W;
    const int mx = (Q.g.dims[0] - 1) >> L, my = (Q.g.dims[1] - 1) >> L, mz = (Q.g.dims[2] - 1) >> L;
    const double r2 = Q.r2;
    while (true) {
        uint32_t cell = 0;
        if (lane == 0) cell = atomicAdd(P.work_counter, 1u);
        cell = __shfl_sync(0xffffffffu, cell, 0);
        if (cell >= Q.ncell) return;
        const uint4 ce = __ldg(P.cinfo + cell);
        const uint32_t first = ce.w, nqc = __ldg(Q.cstart + cell + 1) - first;
        const int cx = (int)ce.x >> L, cy = (int)ce.y >> L, cz = (int)ce.z >> L;
        // ---- the block's point ranges and their prefix ----
        uint32_t run = 0;
        __syncwarp();
        for (int b0 = 0; b0 < RC_MAXCELLS; b0 += 32) {
            const int b = b0 + lane;
            uint32_t lo = 0, hi = 0;
            if (b < NB) {
                const int dx = b % W - R, dy = (b / W) % W - R, dz = b / (W * W) - R;
                const int nx = cx + dx, ny = cy + dy, nz = cz + dz;
                if (nx >= 0 && ny >= 0 && nz >= 0 && nx <= mx && ny <= my && nz <= mz)
                    pcp_level_range(Q.ckey, Q.cstart, Q.ncell, L, nx, ny, nz, lo, hi);
            }
            uint32_t inc = hi - lo;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += v;
            }
            pre[b] = run + inc - (hi - lo);
            los[b] = lo;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        const uint32_t T = run;
        __syncwarp();
        for (uint32_t q0 = 0; q0 < nqc; q0 += 32) {
            const bool have = q0 + lane < nqc;
            const pcp_rec q = pcp_load_rec(Q.rec, first + (have ? q0 + lane : 0));
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0, a5 = 0, a6 = 0, a7 = 0, a8 = 0, a9 = 0;
            for (uint32_t c0 = 0; c0 < T; c0 += RC_CHUNK) {
                const uint32_t n = min((uint32_t)RC_CHUNK, T - c0);
                __syncwarp();
                for (uint32_t i = lane; i < n; i += 32) {
                    const uint32_t f = c0 + i;
                    int t = 0;
#pragma unroll
                    for (int step = RC_MAXCELLS / 2; step; step >>= 1)
                        if (pre[t + step] <= f) t += step;
                    const pcp_rec r = pcp_load_rec(Q.rec, los[t] + (f - pre[t]));
                    sx[i] = r.x, sy[i] = r.y, sz[i] = r.z;
                }
                __syncwarp();
#pragma unroll 4
                for (uint32_t i = 0; i < n; i++) {
                    const double dx = sx[i] - q.x, dy = sy[i] - q.y, dz = sz[i] - q.z;
                    const double xx = dx * dx, yy = dy * dy, zz = dz * dz;
                    const double d2 = (xx + yy) + zz; // == pcp_dist2: (q - p)^2 == (p - q)^2 exactly
                    if (d2 < r2) {
                        a0 += 1.0, a1 += dx, a2 += dy, a3 += dz;
                        a4 += xx, a5 += dx * dy, a6 += dx * dz, a7 += yy, a8 += dy * dz, a9 += zz;
                    }
                }
            }
            if (have) {
                double* m = mom + (size_t)(first + q0 + lane) * 10;
                m[0] = a0, m[1] = a1, m[2] = a2, m[3] = a3, m[4] = a4;
                m[5] = a5, m[6] = a6, m[7] = a7, m[8] = a8, m[9] = a9;
            }
        }
    }
}

// thread per query: moments -> mean / scatter -> plane fit -> outputs at the ORIGINAL index
__global__ void __launch_bounds__(128) radius_features_kernel(const pcp_rec* __restrict__ rec,
                                                              const double* __restrict__ mom, size_t nq, void* props,
                                                              double* eig3, float* lps3, int32_t* counts)
{
    const size_t qi = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    const pcp_rec q = pcp_load_rec(rec, (uint32_t)qi);
    const double* a = mom + qi * 10;
    const double n = a[0];
    const size_t jq = (size_t)q.idx;
    if (counts) counts[jq] = (int32_t)n;
    pcp_plane pl;
    const bool fitted = n > 3.0;
    if (fitted) {
        const double mx = a[1] / n, my = a[2] / n, mz = a[3] / n;
        // sum (d-m)(d-m)^T = sum d d^T - n m m^T
        const double s00 = a[4] - n * mx * mx, s01 = a[5] - n * mx * my, s02 = a[6] - n * mx * mz;
        const double s11 = a[7] - n * my * my, s12 = a[8] - n * my * mz, s22 = a[9] - n * mz * mz;
        pcp_plane_from_scatter(s00, s01, s02, s11, s12, s22, q.x + mx, q.y + my, q.z + mz, pl);
    }
    pcp_store_props(props, jq, fitted, pl);
    pcp_store_extra(eig3, lps3, jq, fitted, pl);
}

__global__ void __launch_bounds__(256) cap_offsets_kernel(const unsigned long long* __restrict__ cnt, size_t nq,
                                                          unsigned long long cap, unsigned long long* __restrict__ capped)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < nq) capped[i] = cnt[i] < cap ? cnt[i] : cap;
}

__global__ void __launch_bounds__(256) rad_props_fallback_kernel(void* props, double* eig3, float* lps3, size_t n)
{
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n) return;
    pcp_plane pl;
    pcp_store_props(props, j, false, pl);
    pcp_store_extra(eig3, lps3, j, false, pl);
}

__global__ void __launch_bounds__(256) counts_fallback_kernel(int32_t* counts, size_t n)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) counts[i] = 0;
}

__global__ void __launch_bounds__(128) copy_capped_kernel(const unsigned long long* __restrict__ full_off,
                                                          const long long* __restrict__ user_off, size_t nq,
                                                          const double* __restrict__ d2, const int32_t* __restrict__ idx,
                                                          double* __restrict__ out_d2, int32_t* __restrict__ out_idx)
{
    const int lane = threadIdx.x & 31;
    const size_t warp0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t qi = warp0; qi < nq; qi += nwarp) {
        const unsigned long long s = full_off[qi];
        const long long o = user_off[qi], e = user_off[qi + 1];
        for (long long t = o + lane; t < e; t += 32) {
            out_idx[t] = idx[s + (t - o)];
            if (out_d2) out_d2[t] = d2[s + (t - o)];
        }
    }
}

int setup(const pcp_index* ix, double r, RadParams& P)
{
    memset(&P, 0, sizeof(P));
    P.rec = ix->rec;
    P.ckey = ix->ckey;
    P.cstart = ix->cstart;
    P.ncell = ix->ncell;
    P.g = ix->grid;
    P.r2 = r * r; // kd_tree.h:888
    int level = 0;
    while (level < ix->grid.top_level && ldexp(ix->grid.h, level) < r * (1.0 + 1e-9) + 1e-12) level++;
    P.level = level;
    return PCP_OK;
}

// counts (device, nq) for external or own queries
int radius_counts(pcp_ctx* ctx, const RadParams& P, unsigned long long* d_cnt)
{
    int blocks = ctx->sm_count * 8;
    radius_scan_kernel<0><<<blocks, 128, 0, ctx->stream>>>(P, d_cnt, nullptr, nullptr, nullptr);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

} // namespace

extern "C" int pcp_radius_count(const pcp_index* ix, const void* q, size_t nq, size_t qstride, double r,
                                uint32_t max_nn, int64_t* offsets)
{
    if (!ix || !offsets || qstride < 24 || (nq && !q) || !(r >= 0))
        return pcp_fail(ix ? ix->ctx : nullptr, PCP_ERR_INVALID, "pcp_radius_count: bad argument");
    pcp_ctx* ctx = ix->ctx;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    OutStage out;
    PCP_TRY(in.open(ctx, q, nq * qstride));
    PCP_TRY(out.open(ctx, offsets, (nq + 1) * 8));
    DevBuf<unsigned long long> cnt, capped;
    PCP_TRY(cnt.alloc(ctx, nq + 1));
    PCP_TRY(capped.alloc(ctx, nq + 1));
    PCP_CUDA_TRY(ctx, cudaMemsetAsync(cnt.p, 0, (nq + 1) * 8, ctx->stream));
    if (nq && ix->n) {
        RadParams P;
        setup(ix, r, P);
        P.qpts = static_cast<const char*>(in.dptr);
        P.qstride = qstride;
        P.nq = nq;
        StageTimer t(ctx, "radius.count");
        PCP_TRY(radius_counts(ctx, P, cnt.p));
    }
    unsigned long long cap = (max_nn == 0 || max_nn > ix->n) ? (unsigned long long)ix->n : max_nn;
    PCP_CUDA_TRY(ctx, cudaMemsetAsync(capped.p, 0, (nq + 1) * 8, ctx->stream));
    if (nq) {
        cap_offsets_kernel<<<pcp_blocks_for(nq, 256), 256, 0, ctx->stream>>>(cnt.p, nq, cap, capped.p);
        PCP_LAUNCH_CHECK(ctx);
    }
    PCP_TRY(pcp_exclusive_scan_u64(ctx, capped.p, (unsigned long long*)out.dptr, nq + 1));
    PCP_TRY(out.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

extern "C" int pcp_radius_fill(const pcp_index* ix, const void* q, size_t nq, size_t qstride, double r, uint32_t max_nn,
                               const int64_t* offsets, int32_t* idx, double* d2)
{
    if (!ix || !offsets || qstride < 24 || (nq && !q) || !(r >= 0))
        return pcp_fail(ix ? ix->ctx : nullptr, PCP_ERR_INVALID, "pcp_radius_fill: bad argument");
    pcp_ctx* ctx = ix->ctx;
    if (nq == 0 || ix->n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in, uoff;
    PCP_TRY(in.open(ctx, q, nq * qstride));
    PCP_TRY(uoff.open(ctx, offsets, (nq + 1) * 8));
    RadParams P;
    setup(ix, r, P);
    P.qpts = static_cast<const char*>(in.dptr);
    P.qstride = qstride;
    P.nq = nq;
    // full (uncapped) lists
    DevBuf<unsigned long long> cnt, foff;
    PCP_TRY(cnt.alloc(ctx, nq + 1));
    PCP_TRY(foff.alloc(ctx, nq + 1));
    PCP_CUDA_TRY(ctx, cudaMemsetAsync(cnt.p, 0, (nq + 1) * 8, ctx->stream));
    PCP_TRY(radius_counts(ctx, P, cnt.p));
    PCP_TRY(pcp_exclusive_scan_u64(ctx, cnt.p, foff.p, nq + 1));
    unsigned long long total = 0;
    PCP_TRY(pcp_readback(ctx, &total, foff.p + nq, 8));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (total == 0) return PCP_OK;
    if (total > 0x7fffffffull) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_radius_fill: more than 2^31-1 neighbours in one call");
    DevBuf<double> fd2, fd2b;
    DevBuf<int32_t> fidx, fidxb;
    PCP_TRY(fd2.alloc(ctx, total));
    PCP_TRY(fd2b.alloc(ctx, total));
    PCP_TRY(fidx.alloc(ctx, total));
    PCP_TRY(fidxb.alloc(ctx, total));
    {
        StageTimer t(ctx, "radius.fill");
        radius_scan_kernel<1><<<ctx->sm_count * 8, 128, 0, ctx->stream>>>(P, nullptr, foff.p, fd2.p, fidx.p);
        PCP_LAUNCH_CHECK(ctx);
    }
    {
        // order every list by (d2, index): stable segmented sort by index, then by d2
        StageTimer t(ctx, "radius.sort");
        size_t tb = 0, tb2 = 0;
        cub::DoubleBuffer<int32_t> ki(fidx.p, fidxb.p);
        cub::DoubleBuffer<double> vd(fd2.p, fd2b.p);
        PCP_CUDA_TRY(ctx, cub::DeviceSegmentedSort::StableSortPairs(nullptr, tb, ki, vd, (int)total, (int)nq, foff.p,
                                                                   foff.p + 1, ctx->stream));
        cub::DoubleBuffer<double> kd(fd2.p, fd2b.p);
        cub::DoubleBuffer<int32_t> vi(fidx.p, fidxb.p);
        PCP_CUDA_TRY(ctx, cub::DeviceSegmentedSort::StableSortPairs(nullptr, tb2, kd, vi, (int)total, (int)nq, foff.p,
                                                                   foff.p + 1, ctx->stream));
        DevBuf<char> tmp;
        PCP_TRY(tmp.alloc(ctx, tb > tb2 ? tb : tb2));
        PCP_CUDA_TRY(ctx, cub::DeviceSegmentedSort::StableSortPairs(tmp.p, tb, ki, vd, (int)total, (int)nq, foff.p,
                                                                   foff.p + 1, ctx->stream));
        cub::DoubleBuffer<double> kd2(vd.Current(), vd.Alternate());
        cub::DoubleBuffer<int32_t> vi2(ki.Current(), ki.Alternate());
        PCP_CUDA_TRY(ctx, cub::DeviceSegmentedSort::StableSortPairs(tmp.p, tb2, kd2, vi2, (int)total, (int)nq, foff.p,
                                                                   foff.p + 1, ctx->stream));
        ctx->launches += 6;
        OutStage oi, od;
        // user lists are addressed through the user's (capped) offsets
        long long utotal = 0;
        PCP_TRY(pcp_readback(ctx, &utotal, (const long long*)uoff.dptr + nq, 8));
        PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
        PCP_TRY(oi.open(ctx, idx, (size_t)utotal * 4));
        PCP_TRY(od.open(ctx, d2, (size_t)utotal * 8));
        copy_capped_kernel<<<ctx->sm_count * 8, 128, 0, ctx->stream>>>(foff.p, (const long long*)uoff.dptr, nq,
                                                                       kd2.Current(), vi2.Current(), (double*)od.dptr,
                                                                       (int32_t*)oi.dptr);
        PCP_LAUNCH_CHECK(ctx);
        PCP_TRY(oi.commit());
        PCP_TRY(od.commit());
    }
    (void)max_nn; // the cap is already encoded in the caller's offsets
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

extern "C" int pcp_features_radius(const pcp_index* ix, double r, uint32_t max_nn, void* props, double* eig3,
                                   float* lps3, int32_t* counts)
{
    if (!ix || !props || !(r >= 0))
        return pcp_fail(ix ? ix->ctx : nullptr, PCP_ERR_INVALID, "pcp_features_radius: bad argument");
    pcp_ctx* ctx = ix->ctx;
    if (max_nn != 0 && max_nn <= ix->n)
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_features_radius: a max_nn cap is not supported by the fused path; "
                                              "use pcp_radius_count/fill + pcp_plane_fit");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const size_t n0 = ix->n_orig, nv = ix->n;
    if (n0 == 0) return PCP_OK;
    OutStage o_props, o_eig, o_lps, o_cnt;
    PCP_TRY(o_props.open(ctx, props, n0 * 48));
    PCP_TRY(o_eig.open(ctx, eig3, n0 * 24));
    PCP_TRY(o_lps.open(ctx, lps3, n0 * 12));
    PCP_TRY(o_cnt.open(ctx, counts, n0 * 4));
    if (nv < n0) { // dropped points get the N<=3 fallback
        rad_props_fallback_kernel<<<pcp_blocks_for(n0, 256), 256, 0, ctx->stream>>>(o_props.dptr, (double*)o_eig.dptr,
                                                                                  (float*)o_lps.dptr, n0);
        PCP_LAUNCH_CHECK(ctx);
        if (o_cnt.dptr) {
            counts_fallback_kernel<<<pcp_blocks_for(n0, 256), 256, 0, ctx->stream>>>((int32_t*)o_cnt.dptr, n0);
            PCP_LAUNCH_CHECK(ctx);
        }
    }
    if (nv) {
        RadParams P;
        setup(ix, r, P);
        P.nq = nv;
        DevBuf<double> mom;
        PCP_TRY(mom.alloc(ctx, nv * 10));
        // the level whose block of (2R+1)^3 cells, R <= 2, covers every ball of a base cell's queries
        int L = 0;
        while (L < ix->grid.top_level && !(r / ldexp(ix->grid.h, L) < 1.999)) L++;
        const int blockR = (int)floor(r / ldexp(ix->grid.h, L)) + 1;
        if (blockR <= 2 && !getenv("PCP_RADIUS_WARP")) {
            StageTimer t(ctx, "radius.moments_kernel", 72.0 * nv);
            DevBuf<uint32_t> counter;
            PCP_TRY(counter.alloc(ctx, 1));
            PCP_CUDA_TRY(ctx, cudaMemsetAsync(counter.p, 0, 4, ctx->stream));
            RadCellParams C;
            C.R = P;
            C.R.level = L;
            C.cinfo = ix->cinfo;
            C.blockR = blockR;
            C.work_counter = counter.p;
            int per_sm = 0;
            PCP_CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radius_moments_cell_kernel, 32 * RC_WARPS, 0));
            if (per_sm < 1) per_sm = 1;
            radius_moments_cell_kernel<<<ctx->sm_count * per_sm, 32 * RC_WARPS, 0, ctx->stream>>>(C, mom.p);
            PCP_LAUNCH_CHECK(ctx);
        } else { // the whole grid is smaller than the ball: warp per query over the 3x3x3 block of the covering level
            StageTimer t(ctx, "radius.moments_kernel", 72.0 * nv);
            radius_moments_kernel<<<ctx->sm_count * 16, 128, 0, ctx->stream>>>(P, mom.p);
            PCP_LAUNCH_CHECK(ctx);
        }
        {
            StageTimer t(ctx, "radius.features_kernel");
            radius_features_kernel<<<pcp_blocks_for(nv, 128), 128, 0, ctx->stream>>>(
                ix->rec, mom.p, nv, o_props.dptr, (double*)o_eig.dptr, (float*)o_lps.dptr, (int32_t*)o_cnt.dptr);
            PCP_LAUNCH_CHECK(ctx);
        }
    }
    PCP_TRY(o_props.commit());
    PCP_TRY(o_eig.commit());
    PCP_TRY(o_lps.commit());
    PCP_TRY(o_cnt.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}
