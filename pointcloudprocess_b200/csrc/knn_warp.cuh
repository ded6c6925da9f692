// knn_warp.cuh -- warp-per-query exact kNN with level escalation: KdTreeFLANN::nearestKSearch
// (kd_tree.h:814-845) for arbitrary query points, and the fall-back for queries the cell kernel could not
// prove exact.  The 32 lanes stream the points of the 3x3x3 block of level-l cells (coalesced 32-byte
// records), keep the best candidates as a warp-distributed sorted list (entry e lives in lane e%32, slot
// e/32) and insert with ballot + shuffle; ordering is (d2, original index).  If the k-th distance is not
// provably inside the block the query moves one level up (cell edge doubles) and rescans.
#pragma once
#include "knn_cell.cuh"

struct pcp_wknn_params {
    const pcp_rec* rec;
    const unsigned long long* ckey;
    const uint32_t* cstart;
    uint32_t ncell;
    // hash tables of the index (may be NULL): cell ranges by one or two probes instead of two binary searches over the
    // cell keys (46 dependent loads per level at 6 M cells -- the latency that bounded this kernel)
    const pcp_hslot* htab;
    uint32_t hmask;
    const pcp_hslot* ltab;
    uint32_t lmask;
    pcp_grid g;
    int k, kk;
    // queries: either sorted positions (qlist, with count in *qcount_dev or qcount) or external points
    const uint32_t* qlist;
    const uint32_t* qcount_dev;
    uint32_t qcount;
    const char* qpts;
    size_t qstride;
    const uint32_t* qsel; // optional, with qpts: query qi is external point qsel[qi], and that is its output row
    // outputs, row r = query ordinal in the list: kk sorted positions (and squared distances)
    uint32_t* out_pos; // [nq][k]
    double* out_d2;    // [nq][k] or NULL
    uint32_t* work_counter;
    unsigned long long* dbg; // optional (PCP_KNN_DEBUG): [level] queries settled there, [16] points scanned
};

template <int KPL> struct pcp_wlist {
    double d[KPL];
    int idx[KPL];
    uint32_t pos[KPL];
};

__device__ __forceinline__ bool pcp_less(double d, int i, double e, int j) { return d < e || (d == e && i < j); }

template <int KPL>
__device__ __forceinline__ void pcp_wlist_insert(pcp_wlist<KPL>& L, int lane, double xd, int xi, uint32_t xp)
{
    int p = 0;
#pragma unroll
    for (int s = 0; s < KPL; s++) p += __popc(__ballot_sync(0xffffffffu, pcp_less(L.d[s], L.idx[s], xd, xi)));
#pragma unroll
    for (int s = KPL - 1; s >= 0; s--) {
        double pd = __shfl_up_sync(0xffffffffu, L.d[s], 1);
        int pi = __shfl_up_sync(0xffffffffu, L.idx[s], 1);
        uint32_t pp = __shfl_up_sync(0xffffffffu, L.pos[s], 1);
        if (s > 0) {
            double cd = __shfl_sync(0xffffffffu, L.d[s - 1], 31);
            int ci = __shfl_sync(0xffffffffu, L.idx[s - 1], 31);
            uint32_t cp = __shfl_sync(0xffffffffu, L.pos[s - 1], 31);
            if (lane == 0) pd = cd, pi = ci, pp = cp;
        }
        const int e = s * 32 + lane;
        if (e > p) {
            L.d[s] = pd;
            L.idx[s] = pi;
            L.pos[s] = pp;
        } else if (e == p) {
            L.d[s] = xd;
            L.idx[s] = xi;
            L.pos[s] = xp;
        }
    }
}

template <int KPL> __device__ __forceinline__ void pcp_wlist_kth(const pcp_wlist<KPL>& L, int kk, double& kd, int& ki)
{
    const int e = kk - 1;
    double d = L.d[0];
    int i = L.idx[0];
#pragma unroll
    for (int s = 1; s < KPL; s++)
        if ((e >> 5) == s) d = L.d[s], i = L.idx[s];
    kd = __shfl_sync(0xffffffffu, d, e & 31);
    ki = __shfl_sync(0xffffffffu, i, e & 31);
}

template <int KPL> static __global__ void __launch_bounds__(128) knn_warp_kernel(pcp_wknn_params P)
{
    const int lane = threadIdx.x & 31;
    const uint32_t nq = P.qcount_dev ? *P.qcount_dev : P.qcount;
    while (true) {
        uint32_t qi = 0;
        if (lane == 0) qi = atomicAdd(P.work_counter, 1u);
        qi = __shfl_sync(0xffffffffu, qi, 0);
        if (qi >= nq) return;
        double qx, qy, qz;
        int level = 0;
        const uint32_t row = (!P.qlist && P.qsel) ? __ldg(P.qsel + qi) : qi;
        if (P.qlist) {
            uint32_t code = __ldg(P.qlist + qi);
            level = (code & 0x80000000u) ? 1 : 0;
            const pcp_rec q = pcp_load_rec(P.rec, code & 0x7fffffffu);
            qx = q.x, qy = q.y, qz = q.z;
        } else {
            const double* qp = reinterpret_cast<const double*>(P.qpts + (size_t)row * P.qstride);
            qx = qp[0], qy = qp[1], qz = qp[2];
        }
        pcp_wlist<KPL> L;
        int found = 0;
        const bool finite_q = isfinite(qx) && isfinite(qy) && isfinite(qz);
        if (finite_q && P.kk > 0) {
            // base-cell coordinates of the query, clamped into the grid
            const int bx = min(max(pcp_cell_coord(qx, P.g.org[0], P.g.inv_h, 0), 0), P.g.dims[0] - 1);
            const int by = min(max(pcp_cell_coord(qy, P.g.org[1], P.g.inv_h, 0), 0), P.g.dims[1] - 1);
            const int bz = min(max(pcp_cell_coord(qz, P.g.org[2], P.g.inv_h, 0), 0), P.g.dims[2] - 1);
            if (level > P.g.top_level) level = P.g.top_level;
            // the points [s0, s1) against the list: four chunks of 32 points in flight per lane (the scan is bound by the
            // latency of its loads; insertions are rare once the list is warm, and the result does not depend on the order)
            auto scan_range = [&](pcp_wlist<KPL>& Ls, double& kd, int& ki, uint32_t s0, uint32_t s1) {
                constexpr int WU = 4;
                for (uint32_t c0 = s0; c0 < s1; c0 += 32 * WU) {
                    pcp_rec r[WU];
#pragma unroll
                    for (int u = 0; u < WU; u++) {
                        const uint32_t t = c0 + 32 * u + lane;
                        if (t < s1) r[u] = pcp_load_rec(P.rec, t);
                    }
#pragma unroll
                    for (int u = 0; u < WU; u++) {
                        const uint32_t t0 = c0 + 32 * u;
                        if (t0 >= s1) break;
                        const uint32_t t = t0 + lane;
                        double d2 = DBL_MAX;
                        int ci = 0x7fffffff;
                        if (t < s1) {
                            d2 = pcp_dist2(qx, qy, qz, r[u].x, r[u].y, r[u].z);
                            ci = (int)r[u].idx;
                        }
                        unsigned qual = __ballot_sync(0xffffffffu, t < s1 && pcp_less(d2, ci, kd, ki));
                        while (qual) {
                            const int src = __ffs(qual) - 1;
                            qual &= qual - 1;
                            const double xd = __shfl_sync(0xffffffffu, d2, src);
                            const int xi = __shfl_sync(0xffffffffu, ci, src);
                            if (!pcp_less(xd, xi, kd, ki)) continue; // threshold moved since the ballot
                            pcp_wlist_insert<KPL>(Ls, lane, xd, xi, t0 + src);
                            pcp_wlist_kth<KPL>(Ls, P.kk, kd, ki);
                        }
                    }
                }
            };
            // (measured and dropped: with the far kernel's bound, streaming only the base cells that the bound sphere
            // touches -- the overflow queries' bounds are corner bounds, their boxes thousands of cells: no faster)
            const bool settled = false;
            for (; !settled && level <= P.g.top_level; level++) {
#pragma unroll
                for (int s = 0; s < KPL; s++) L.d[s] = DBL_MAX, L.idx[s] = 0x7fffffff, L.pos[s] = 0xffffffffu;
                const int cx = bx >> level, cy = by >> level, cz = bz >> level;
                uint32_t lo = 0, hi = 0;
                if (lane < 27) {
                    int dx, dy, dz;
                    pcp_nb_offset(lane, dx, dy, dz);
                    const int nx = cx + dx, ny = cy + dy, nz = cz + dz;
                    const int mx = (P.g.dims[0] - 1) >> level, my = (P.g.dims[1] - 1) >> level,
                              mz = (P.g.dims[2] - 1) >> level;
                    if (nx >= 0 && ny >= 0 && nz >= 0 && nx <= mx && ny <= my && nz <= mz) {
                        if (P.htab && level == 0) {
                            uint32_t cs = 0, cc = 0;
                            if (pcp_hash_find(P.htab, P.hmask, pcp_morton(nx, ny, nz), cs, cc)) lo = cs, hi = cs + cc;
                        } else if (P.ltab && level < PCP_HLEVELS) {
                            uint32_t a = 0, cn = 0;
                            if (pcp_hash_find(P.ltab, P.lmask, PCP_LKEY(level, pcp_morton(nx, ny, nz)), a, cn))
                                lo = __ldg(P.cstart + a), hi = __ldg(P.cstart + a + cn);
                        } else {
                            pcp_level_range(P.ckey, P.cstart, P.ncell, level, nx, ny, nz, lo, hi);
                        }
                    }
                }
                double kd = DBL_MAX;
                int ki = 0x7fffffff;
                for (int nb = 0; nb < 27; nb++) {
                    const uint32_t s0 = __shfl_sync(0xffffffffu, lo, nb);
                    const uint32_t s1 = __shfl_sync(0xffffffffu, hi, nb);
                    scan_range(L, kd, ki, s0, s1);
                }
                const double g2 = pcp_guarantee2(P.g, level, cx, cy, cz, qx, qy, qz);
                if (P.dbg) {
                    const unsigned scanned = __reduce_add_sync(0xffffffffu, hi - lo);
                    if (lane == 0) atomicAdd(P.dbg + 16, (unsigned long long)scanned);
                }
                if (level == P.g.top_level || (kd < DBL_MAX && kd < g2)) break;
            }
            found = P.kk;
            if (P.dbg && lane == 0) atomicAdd(P.dbg + min(level, 15), 1ull);
        }
        // write the list: entry e -> out[qi*k + e]
#pragma unroll
        for (int s = 0; s < KPL; s++) {
            const int e = s * 32 + lane;
            if (e < P.k) {
                const bool have = e < found && L.d[s] < DBL_MAX;
                P.out_pos[(size_t)row * P.k + e] = have ? L.pos[s] : 0xffffffffu;
                if (P.out_d2) P.out_d2[(size_t)row * P.k + e] = have ? L.d[s] : __longlong_as_double(0x7ff0000000000000ll);
            }
        }
    }
}

// ---- the plane fit of the deferred rows in two kernels ----------------------------------------------------------
// The fit is two different programs: gathering twenty records and summing them in order (loads, little state), and
// the Jacobi iteration (sixty live doubles, no loads).  In one kernel the first waits on memory at the occupancy the
// second's registers allow, and the FP64 pipe idles half the time (54 % busy).  Split, the moments pass runs at
// full occupancy and hands nine doubles per row to a pass that only computes.  mom: [10][rows] (column-major, so both
// passes are coalesced): s00 s01 s02 s11 s12 s22 XAve YAve ZAve, and N << 32 | original index (N = -1: no row).
static __global__ void __launch_bounds__(256) moments_from_rows_kernel(const pcp_rec* __restrict__ rec,
                                                                       const uint32_t* __restrict__ nbr_pos, int k, int kk,
                                                                       uint32_t row0, uint32_t nrows, double* __restrict__ mom,
                                                                       int32_t* knn_idx, float* kth_d2)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nrows) return;
    const uint32_t qi = row0 + t;
    const uint32_t* row = nbr_pos + (size_t)qi * k;
    long long tag = -1ll << 32;
    if (row[0] != 0xffffffffu) {
        const size_t jq = (size_t)rec[qi].idx;
        int N = 0;
        for (int e = 0; e < kk; e++)
            if (row[e] != 0xffffffffu) N++;
        if (knn_idx)
            for (int e = 0; e < k; e++) knn_idx[jq * k + e] = e < N ? (int32_t)rec[row[e]].idx : -1;
        if (kth_d2 && N > 0) {
            const pcp_rec q = pcp_load_rec(rec, qi), r = pcp_load_rec(rec, row[N - 1]);
            kth_d2[jq] = __double2float_ru(pcp_dist2(q.x, q.y, q.z, r.x, r.y, r.z));
        }
        tag = ((long long)N << 32) | (long long)(uint32_t)jq;
        if (N > 3) {
            double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
            for (int e = 0; e < N; e++) {
                const pcp_rec r = pcp_load_rec(rec, row[e]);
                XAve += r.x;
                YAve += r.y;
                ZAve += r.z;
            }
            XAve /= (double)N;
            YAve /= (double)N;
            ZAve /= (double)N;
            double s00 = 0, s01 = 0, s02 = 0, s11 = 0, s12 = 0, s22 = 0;
            for (int e = 0; e < N; e++) {
                const pcp_rec r = pcp_load_rec(rec, row[e]);
                const double X = r.x - XAve, Y = r.y - YAve, Z = r.z - ZAve;
                s00 += X * X;
                s01 += X * Y;
                s02 += X * Z;
                s11 += Y * Y;
                s12 += Y * Z;
                s22 += Z * Z;
            }
            mom[0 * (size_t)nrows + t] = s00, mom[1 * (size_t)nrows + t] = s01, mom[2 * (size_t)nrows + t] = s02;
            mom[3 * (size_t)nrows + t] = s11, mom[4 * (size_t)nrows + t] = s12, mom[5 * (size_t)nrows + t] = s22;
            mom[6 * (size_t)nrows + t] = XAve, mom[7 * (size_t)nrows + t] = YAve, mom[8 * (size_t)nrows + t] = ZAve;
        }
    }
    mom[9 * (size_t)nrows + t] = __longlong_as_double(tag);
}

static __global__ void __launch_bounds__(128) plane_from_moments_kernel(const double* __restrict__ mom, uint32_t nrows, void* props,
                                                                        double* eig3, float* lps3)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nrows) return;
    const long long tag = __double_as_longlong(mom[9 * (size_t)nrows + t]);
    const int N = (int)(tag >> 32);
    if (N < 0) return;
    const size_t jq = (size_t)(uint32_t)tag;
    pcp_plane pl;
    const bool fitted = N > 3;
    if (fitted)
        pcp_plane_from_scatter<true>(mom[0 * (size_t)nrows + t], mom[1 * (size_t)nrows + t], mom[2 * (size_t)nrows + t],
                                     mom[3 * (size_t)nrows + t], mom[4 * (size_t)nrows + t], mom[5 * (size_t)nrows + t],
                                     mom[6 * (size_t)nrows + t], mom[7 * (size_t)nrows + t], mom[8 * (size_t)nrows + t], pl);
    pcp_store_props(props, jq, fitted, pl);
    pcp_store_extra(eig3, lps3, jq, fitted, pl);
}

// ---- the same search, one BLOCK per query ----------------------------------------------------------------------
// For a handful of queries (what the far kernel could not settle: a point metres above a dense surface, whose block
// holds 10^5 points) the warp kernel's time is the time of its longest query, and a single warp keeps only 4 KB of
// loads in flight.  Here the W warps of a block share one query: every warp scans its W-th of each of the 27 cell
// ranges into a list of its own; the lists meet in shared memory and warp 0 merges them (the k nearest of a union are
// among the k nearest of its parts), decides whether the level certifies, and tells the others.  Same order
// (d2, original index), same levels, same answer as knn_warp_kernel.
struct pcp_bent {
    double d;
    int idx;
    uint32_t pos;
};

template <int KPL, int W> static __global__ void __launch_bounds__(32 * W) knn_block_kernel(pcp_wknn_params P)
{
    extern __shared__ __align__(16) unsigned char bsm_raw[];
    pcp_bent* const part = reinterpret_cast<pcp_bent*>(bsm_raw); // [W - 1][kk]
    __shared__ uint32_t s_qi;
    __shared__ int s_done;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t nq = P.qcount_dev ? *P.qcount_dev : P.qcount;
    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) s_qi = atomicAdd(P.work_counter, 1u);
        __syncthreads();
        const uint32_t qi = s_qi;
        if (qi >= nq) return;
        double qx, qy, qz;
        int level = 0;
        const uint32_t row = (!P.qlist && P.qsel) ? __ldg(P.qsel + qi) : qi;
        if (P.qlist) {
            const uint32_t code = __ldg(P.qlist + qi);
            level = (code & 0x80000000u) ? 1 : 0;
            const pcp_rec q = pcp_load_rec(P.rec, code & 0x7fffffffu);
            qx = q.x, qy = q.y, qz = q.z;
        } else {
            const double* qp = reinterpret_cast<const double*>(P.qpts + (size_t)row * P.qstride);
            qx = qp[0], qy = qp[1], qz = qp[2];
        }
        pcp_wlist<KPL> L;
        int found = 0;
        const bool finite_q = isfinite(qx) && isfinite(qy) && isfinite(qz);
        if (finite_q && P.kk > 0) {
            const int bx = min(max(pcp_cell_coord(qx, P.g.org[0], P.g.inv_h, 0), 0), P.g.dims[0] - 1);
            const int by = min(max(pcp_cell_coord(qy, P.g.org[1], P.g.inv_h, 0), 0), P.g.dims[1] - 1);
            const int bz = min(max(pcp_cell_coord(qz, P.g.org[2], P.g.inv_h, 0), 0), P.g.dims[2] - 1);
            if (level > P.g.top_level) level = P.g.top_level;
            double kd = DBL_MAX;
            int ki = 0x7fffffff;
            // candidate (xd, xi, xp) of lane src against the list
            auto offer = [&](unsigned qual, double d2, int ci, uint32_t pos) {
                while (qual) {
                    const int src = __ffs(qual) - 1;
                    qual &= qual - 1;
                    const double xd = __shfl_sync(0xffffffffu, d2, src);
                    const int xi = __shfl_sync(0xffffffffu, ci, src);
                    const uint32_t xp = __shfl_sync(0xffffffffu, pos, src);
                    if (!pcp_less(xd, xi, kd, ki)) continue; // threshold moved since the ballot
                    pcp_wlist_insert<KPL>(L, lane, xd, xi, xp);
                    pcp_wlist_kth<KPL>(L, P.kk, kd, ki);
                }
            };
            for (;; level++) {
#pragma unroll
                for (int s = 0; s < KPL; s++) L.d[s] = DBL_MAX, L.idx[s] = 0x7fffffff, L.pos[s] = 0xffffffffu;
                kd = DBL_MAX, ki = 0x7fffffff;
                const int cx = bx >> level, cy = by >> level, cz = bz >> level;
                uint32_t lo = 0, hi = 0;
                if (lane < 27) {
                    int dx, dy, dz;
                    pcp_nb_offset(lane, dx, dy, dz);
                    const int nx = cx + dx, ny = cy + dy, nz = cz + dz;
                    const int mx = (P.g.dims[0] - 1) >> level, my = (P.g.dims[1] - 1) >> level,
                              mz = (P.g.dims[2] - 1) >> level;
                    if (nx >= 0 && ny >= 0 && nz >= 0 && nx <= mx && ny <= my && nz <= mz) {
                        if (P.htab && level == 0) {
                            uint32_t cs = 0, cc = 0;
                            if (pcp_hash_find(P.htab, P.hmask, pcp_morton(nx, ny, nz), cs, cc)) lo = cs, hi = cs + cc;
                        } else if (P.ltab && level < PCP_HLEVELS) {
                            uint32_t a = 0, cn = 0;
                            if (pcp_hash_find(P.ltab, P.lmask, PCP_LKEY(level, pcp_morton(nx, ny, nz)), a, cn))
                                lo = __ldg(P.cstart + a), hi = __ldg(P.cstart + a + cn);
                        } else {
                            pcp_level_range(P.ckey, P.cstart, P.ncell, level, nx, ny, nz, lo, hi);
                        }
                    }
                }
                for (int nb = 0; nb < 27; nb++) {
                    const uint32_t s0 = __shfl_sync(0xffffffffu, lo, nb);
                    const uint32_t s1 = __shfl_sync(0xffffffffu, hi, nb);
                    // this warp's part of the range, in chunks of 4 x 32 points in flight
                    const unsigned long long len = s1 - s0;
                    const uint32_t a = s0 + (uint32_t)(len * (unsigned)warp / W), b = s0 + (uint32_t)(len * (unsigned)(warp + 1) / W);
                    constexpr int WU = 4;
                    for (uint32_t c0 = a; c0 < b; c0 += 32 * WU) {
                        pcp_rec r[WU];
#pragma unroll
                        for (int u = 0; u < WU; u++) {
                            const uint32_t t = c0 + 32 * u + lane;
                            if (t < b) r[u] = pcp_load_rec(P.rec, t);
                        }
#pragma unroll
                        for (int u = 0; u < WU; u++) {
                            const uint32_t t0 = c0 + 32 * u;
                            if (t0 >= b) break;
                            const uint32_t t = t0 + lane;
                            double d2 = DBL_MAX;
                            int ci = 0x7fffffff;
                            if (t < b) {
                                d2 = pcp_dist2(qx, qy, qz, r[u].x, r[u].y, r[u].z);
                                ci = (int)r[u].idx;
                            }
                            offer(__ballot_sync(0xffffffffu, t < b && pcp_less(d2, ci, kd, ki)), d2, ci, t);
                        }
                    }
                }
                // the lists of warps 1 .. W-1 go to warp 0
                if (warp > 0) {
#pragma unroll
                    for (int s = 0; s < KPL; s++) {
                        const int e = s * 32 + lane;
                        if (e < P.kk) part[(size_t)(warp - 1) * P.kk + e] = pcp_bent{L.d[s], L.idx[s], L.pos[s]};
                    }
                }
                __syncthreads();
                if (warp == 0) {
                    for (int w = 0; w < W - 1; w++)
                        for (int e0 = 0; e0 < P.kk; e0 += 32) {
                            const int e = e0 + lane;
                            pcp_bent x{DBL_MAX, 0x7fffffff, 0xffffffffu};
                            if (e < P.kk) x = part[(size_t)w * P.kk + e];
                            offer(__ballot_sync(0xffffffffu, x.d < DBL_MAX && pcp_less(x.d, x.idx, kd, ki)), x.d, x.idx, x.pos);
                        }
                    const double g2 = pcp_guarantee2(P.g, level, cx, cy, cz, qx, qy, qz);
                    if (lane == 0) s_done = (level >= P.g.top_level || (kd < DBL_MAX && kd < g2)) ? 1 : 0;
                }
                __syncthreads();
                if (s_done) break;
            }
            found = P.kk;
        }
        if (warp == 0) {
#pragma unroll
            for (int s = 0; s < KPL; s++) {
                const int e = s * 32 + lane;
                if (e < P.k) {
                    const bool have = e < found && L.d[s] < DBL_MAX;
                    P.out_pos[(size_t)row * P.k + e] = have ? L.pos[s] : 0xffffffffu;
                    if (P.out_d2) P.out_d2[(size_t)row * P.k + e] = have ? L.d[s] : __longlong_as_double(0x7ff0000000000000ll);
                }
            }
        }
    }
}

// thread per fall-back query: neighbour positions -> plane fit -> LAS_POINT_PROPERTY
static __global__ void __launch_bounds__(128) features_from_list_kernel(const pcp_rec* __restrict__ rec,
                                                                 const uint32_t* __restrict__ qlist,
                                                                 const uint32_t* __restrict__ qcount_dev,
                                                                 const uint32_t* __restrict__ nbr_pos, int k, int kk,
                                                                 void* props, double* eig3, float* lps3,
                                                                 int32_t* knn_idx, float* kth_d2, uint32_t all_rows = 0)
{
    const uint32_t qi = blockIdx.x * blockDim.x + threadIdx.x;
    // qlist == NULL: one row per sorted position (the deferred features of the dense / far kernels); a row that was
    // never written starts with 0xffffffff (its query is settled elsewhere, or the point is not a query)
    if (qi >= (qlist ? *qcount_dev : all_rows)) return;
    const uint32_t qpos = qlist ? (qlist[qi] & 0x7fffffffu) : qi;
    const uint32_t* row = nbr_pos + (size_t)qi * k;
    if (!qlist && row[0] == 0xffffffffu) return;
    const size_t jq = (size_t)rec[qpos].idx;
    int N = 0;
    for (int e = 0; e < kk; e++)
        if (row[e] != 0xffffffffu) N++;
    if (knn_idx)
        for (int e = 0; e < k; e++) knn_idx[jq * k + e] = e < N ? (int32_t)rec[row[e]].idx : -1;
    if (kth_d2 && N > 0) {
        const pcp_rec q = pcp_load_rec(rec, qpos), r = pcp_load_rec(rec, row[N - 1]);
        kth_d2[jq] = __double2float_ru(pcp_dist2(q.x, q.y, q.z, r.x, r.y, r.z));
    }
    pcp_plane pl;
    const bool fitted = N > 3;
    if (fitted) {
        double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
        for (int e = 0; e < N; e++) {
            const pcp_rec r = pcp_load_rec(rec, row[e]);
            XAve += r.x;
            YAve += r.y;
            ZAve += r.z;
        }
        XAve /= (double)N;
        YAve /= (double)N;
        ZAve /= (double)N;
        double s00 = 0, s01 = 0, s02 = 0, s11 = 0, s12 = 0, s22 = 0;
        for (int e = 0; e < N; e++) {
            const pcp_rec r = pcp_load_rec(rec, row[e]);
            const double X = r.x - XAve, Y = r.y - YAve, Z = r.z - ZAve;
            s00 += X * X;
            s01 += X * Y;
            s02 += X * Z;
            s11 += Y * Y;
            s12 += Y * Z;
            s22 += Z * Z;
        }
        pcp_plane_from_scatter<true>(s00, s01, s02, s11, s12, s22, XAve, YAve, ZAve, pl);
    }
    pcp_store_props(props, jq, fitted, pl);
    pcp_store_extra(eig3, lps3, jq, fitted, pl);
}
