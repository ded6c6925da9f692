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

// The same moments for the index's own points, cell-cooperative: a warp takes one base cell.
//   stage    the balls of all the cell's queries lie inside the block of (2R+1)^3 level-L cells around it
//            (R * edge >= r); the block's points are read once, those farther than r from the cell's box are dropped
//            (no query of the cell can reach them: 3.5x fewer at r = 0.5, h = 0.37), the rest are compacted into shared
//            memory as FP32 and FP64 coordinates relative to the cell origin o;
//   search   8 lanes per query, four queries of the cell in flight per warp; a lane runs over every 8th staged
//            candidate.  The predicate d2 < r2 is decided in FP32 wherever the FP32 distance is farther from r2 than
//            its error bound (`margin`, from the block extent); the few pairs inside the band are decided by the exact
//            FP64 expression of the reference (FLANN L2_Simple on the original coordinates): the neighbour SETS are exact;
//   moments  accumulated about o in FP64 (u = p - o) and moved to the query at the end: sum (p-q) = m1 - n t,
//            sum (p-q)(p-q)^T = m2 - m1 t^T - t m1^T + n t t^T, t = q - o; three xor-shuffle levels reduce a group.
// Against the warp-per-query kernel: the candidates are loaded once per cell instead of once per query, from a
// tighter region, and the FP64 work is confined to the accumulation.
constexpr int RC_WARPS = 4;
constexpr int RC_MAXCELLS = 128; // >= (2R+1)^3, R <= 2
constexpr int RC_CHUNK = 256;    // staged candidates at a time
constexpr int RC_G = 8;          // lanes per query

struct RadCellParams {
    RadParams R;
    const uint4* cinfo; // per base cell: x, y, z, first point
    const pcp_hslot* htab; // base cell -> (first point, count)
    uint32_t hmask;
    const pcp_hslot* ltab; // (level, cell) -> run of base cells
    uint32_t lmask;
    int blockR;         // R
    float r2f, margin;  // FP32 decision band: |d2f - r2f| <= margin goes to the exact test
    float rstage2;      // squared reach of the staging filter: (r + slack)^2
    uint32_t* work_counter;
};

struct RadCellSmem {
    float4 f[RC_CHUNK];     // u as floats, w = sorted position
    double ux[RC_CHUNK], uy[RC_CHUNK], uz[RC_CHUNK];
    uint32_t pre[RC_MAXCELLS]; // exclusive prefix of the block cells' point counts
    uint32_t lo[RC_MAXCELLS];  // first point of each block cell
};

__global__ void __launch_bounds__(32 * RC_WARPS) radius_moments_cell_kernel(RadCellParams P, double* __restrict__ mom)
{
    extern __shared__ __align__(16) unsigned char rc_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    RadCellSmem& S = reinterpret_cast<RadCellSmem*>(rc_smem)[warp];
    const RadParams& Q = P.R;
    const int L = Q.level, R = P.blockR, W = 2 * R + 1, NB = W * W * W;
    const int mx = (Q.g.dims[0] - 1) >> L, my = (Q.g.dims[1] - 1) >> L, mz = (Q.g.dims[2] - 1) >> L;
    const double r2 = Q.r2;
    const float r2f = P.r2f, margin = P.margin, hf = (float)Q.g.h;
    const int gl = lane & (RC_G - 1), grp = lane / RC_G;
    while (true) {
        uint32_t cell = 0;
        if (lane == 0) cell = atomicAdd(P.work_counter, 1u);
        cell = __shfl_sync(0xffffffffu, cell, 0);
        if (cell >= Q.ncell) return;
        const uint4 ce = __ldg(P.cinfo + cell);
        const uint32_t first = ce.w, nqc = __ldg(Q.cstart + cell + 1) - first;
        const int cx = (int)ce.x >> L, cy = (int)ce.y >> L, cz = (int)ce.z >> L;
        const double ox = Q.g.org[0] + (double)ce.x * Q.g.h, oy = Q.g.org[1] + (double)ce.y * Q.g.h,
                     oz = Q.g.org[2] + (double)ce.z * Q.g.h;
        // ---- the block's point ranges and their prefix ----
        uint32_t run = 0;
        __syncwarp();
        for (int b0 = 0; b0 < RC_MAXCELLS; b0 += 32) {
            const int b = b0 + lane;
            uint32_t lo = 0, hi = 0;
            if (b < NB) {
                const int dx = b % W - R, dy = (b / W) % W - R, dz = b / (W * W) - R;
                const int nx = cx + dx, ny = cy + dy, nz = cz + dz;
                if (nx >= 0 && ny >= 0 && nz >= 0 && nx <= mx && ny <= my && nz <= mz) {
                    // hash lookups (one or two probes) instead of two binary searches over the cell keys
                    if (L == 0) {
                        uint32_t cs = 0, cc = 0;
                        if (pcp_hash_find(P.htab, P.hmask, pcp_morton(nx, ny, nz), cs, cc)) lo = cs, hi = cs + cc;
                    } else if (L < PCP_HLEVELS) {
                        uint32_t a = 0, cn = 0;
                        if (pcp_hash_find(P.ltab, P.lmask, PCP_LKEY(L, pcp_morton(nx, ny, nz)), a, cn))
                            lo = __ldg(Q.cstart + a), hi = __ldg(Q.cstart + a + cn);
                    } else {
                        pcp_level_range(Q.ckey, Q.cstart, Q.ncell, L, nx, ny, nz, lo, hi);
                    }
                }
            }
            uint32_t inc = hi - lo;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += v;
            }
            S.pre[b] = run + inc - (hi - lo);
            S.lo[b] = lo;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        const uint32_t T = run;
        __syncwarp();
        uint32_t next = 0; // next block point to stage
        bool first_chunk = true;
        while (next < T) {
            // ---- stage: up to RC_CHUNK points within reach of the cell's box ----
            uint32_t n = 0;
            __syncwarp();
            while (next < T && n + 32 <= RC_CHUNK) {
                const uint32_t f = next + lane;
                bool keep = false;
                double ux = 0, uy = 0, uz = 0;
                uint32_t pos = 0;
                if (f < T) {
                    int t = 0;
#pragma unroll
                    for (int step = RC_MAXCELLS / 2; step; step >>= 1)
                        if (S.pre[t + step] <= f) t += step;
                    pos = S.lo[t] + (f - S.pre[t]);
                    const pcp_rec r = pcp_load_rec(Q.rec, pos);
                    ux = r.x - ox, uy = r.y - oy, uz = r.z - oz;
                    // distance to the box [0,h]^3 (the cell's queries lie in it up to rounding: covered by the slack)
                    const float fx = (float)ux, fy = (float)uy, fz = (float)uz;
                    const float ex = fmaxf(fmaxf(-fx, fx - hf), 0.f), ey = fmaxf(fmaxf(-fy, fy - hf), 0.f),
                                ez = fmaxf(fmaxf(-fz, fz - hf), 0.f);
                    keep = __fmaf_rn(ez, ez, __fmaf_rn(ey, ey, ex * ex)) <= P.rstage2;
                }
                const unsigned mask = __ballot_sync(0xffffffffu, keep);
                if (keep) {
                    const uint32_t at = n + __popc(mask & ((1u << lane) - 1u));
                    S.f[at] = make_float4((float)ux, (float)uy, (float)uz, __uint_as_float(pos));
                    S.ux[at] = ux, S.uy[at] = uy, S.uz[at] = uz;
                }
                n += __popc(mask);
                next += 32;
            }
            __syncwarp();
            // ---- the cell's queries, four at a time ----
            for (uint32_t q0 = 0; q0 < nqc; q0 += 32 / RC_G) {
                const uint32_t qj = q0 + grp;
                const bool have = qj < nqc;
                const pcp_rec q = pcp_load_rec(Q.rec, first + (have ? qj : 0));
                const double tx = q.x - ox, ty = q.y - oy, tz = q.z - oz;
                const float qfx = (float)tx, qfy = (float)ty, qfz = (float)tz;
                double a1 = 0, a2 = 0, a3 = 0, a4 = 0, a5 = 0, a6 = 0, a7 = 0, a8 = 0, a9 = 0;
                unsigned cnt = 0;
#pragma unroll 2
                for (uint32_t i = gl; i < n; i += RC_G) {
                    const float4 cf = S.f[i];
                    const float dx = cf.x - qfx, dy = cf.y - qfy, dz = cf.z - qfz;
                    const float d2f = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
                    bool in = d2f < r2f - margin;
                    if (!in && d2f <= r2f + margin) { // inside the FP32 error band: the reference's own expression decides
                        const pcp_rec r = pcp_load_rec(Q.rec, __float_as_uint(cf.w));
                        in = pcp_dist2(q.x, q.y, q.z, r.x, r.y, r.z) < r2;
                    }
                    if (in && have) {
                        const double ux = S.ux[i], uy = S.uy[i], uz = S.uz[i];
                        cnt++;
                        a1 += ux, a2 += uy, a3 += uz;
                        a4 += ux * ux, a5 += ux * uy, a6 += ux * uz, a7 += uy * uy, a8 += uy * uz, a9 += uz * uz;
                    }
                }
#pragma unroll
                for (int o = 1; o < RC_G; o <<= 1) {
                    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
                    a1 += __shfl_xor_sync(0xffffffffu, a1, o), a2 += __shfl_xor_sync(0xffffffffu, a2, o);
                    a3 += __shfl_xor_sync(0xffffffffu, a3, o), a4 += __shfl_xor_sync(0xffffffffu, a4, o);
                    a5 += __shfl_xor_sync(0xffffffffu, a5, o), a6 += __shfl_xor_sync(0xffffffffu, a6, o);
                    a7 += __shfl_xor_sync(0xffffffffu, a7, o), a8 += __shfl_xor_sync(0xffffffffu, a8, o);
                    a9 += __shfl_xor_sync(0xffffffffu, a9, o);
                }
                if (have) {
                    // about the query: d = p - q = u - t; lane k of the group owns value k (lanes 0, 1 also 8, 9)
                    const double nn = (double)cnt;
                    double v[10];
                    v[0] = nn;
                    v[1] = a1 - nn * tx, v[2] = a2 - nn * ty, v[3] = a3 - nn * tz;
                    v[4] = a4 - 2.0 * a1 * tx + nn * tx * tx;
                    v[5] = a5 - a1 * ty - a2 * tx + nn * tx * ty;
                    v[6] = a6 - a1 * tz - a3 * tx + nn * tx * tz;
                    v[7] = a7 - 2.0 * a2 * ty + nn * ty * ty;
                    v[8] = a8 - a2 * tz - a3 * ty + nn * ty * tz;
                    v[9] = a9 - 2.0 * a3 * tz + nn * tz * tz;
                    double* m = mom + (size_t)(first + qj) * 10;
                    double mine = v[0], extra = gl == 0 ? v[8] : v[9];
#pragma unroll
                    for (int k = 1; k < RC_G; k++)
                        if (gl == k) mine = v[k];
                    if (first_chunk) {
                        m[gl] = mine;
                        if (gl < 2) m[8 + gl] = extra;
                    } else {
                        m[gl] += mine;
                        if (gl < 2) m[8 + gl] += extra;
                    }
                }
            }
            first_chunk = false;
        }
        if (T == 0 && lane < 10) // cannot happen for an occupied cell (it holds its own points); keep mom defined
            for (uint32_t qj = 0; qj < nqc; qj++) mom[(size_t)(first + qj) * 10 + lane] = 0.0;
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
            C.htab = ix->htab, C.hmask = ix->hmask, C.ltab = ix->ltab, C.lmask = ix->lmask;
            C.blockR = blockR;
            C.work_counter = counter.p;
            {
                // FP32 error of d2 near the boundary: coordinates relative to the cell origin reach (R + 2) level-L edges;
                // each float coordinate carries one rounding of that magnitude (2^-24 relative), the difference another,
                // plus the rounding of the double subtraction p - o (epsd); the three products and two sums of d2 add
                // 2^-22 relative.  Doubled for safety.
                const double ext = (double)(blockR + 2) * ldexp(ix->grid.h, L);
                const double ec = 3.0 * 5.97e-8 * ext + 2.0 * ix->epsd;
                const double band = 2.0 * (2.0 * sqrt(3.0) * r * ec + 3.0 * ec * ec + 2.4e-7 * r * r);
                C.r2f = (float)(r * r);
                C.margin = (float)(band + 1.2e-7 * r * r) * 1.001f; // + the rounding of r2 to float
                // staging filter: a query may sit a rounding outside its cell's box and the box distance is FP32
                const double reach = r * (1.0 + 1e-5) + 4.0 * ec + 1e-6 * ext;
                C.rstage2 = (float)(reach * reach) * 1.0001f;
            }
            const size_t smem = (size_t)RC_WARPS * sizeof(RadCellSmem);
            PCP_CUDA_TRY(ctx, cudaFuncSetAttribute(radius_moments_cell_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            int per_sm = 0;
            PCP_CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, radius_moments_cell_kernel, 32 * RC_WARPS, smem));
            if (per_sm < 1) per_sm = 1;
            radius_moments_cell_kernel<<<ctx->sm_count * per_sm, 32 * RC_WARPS, smem, ctx->stream>>>(C, mom.p);
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
