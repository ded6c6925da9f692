// knn_far.cuh -- exact kNN + features for the queries the dense kernel (knn_fast.cuh) could not settle:
// points in sparse cells, queries whose k-th neighbour reaches beyond the 3x3x3 block of base cells, and
// queries whose FP32 ordering was ambiguous.  One warp per query; an implicit-octree search over the
// Morton-sorted cells with sphere pruning at base-cell granularity:
//   level   the smallest level whose 3x3x3 block of level cells certifies the known upper bound U of the
//           k-th squared distance (U comes from the dense kernel), or -- without a bound -- the smallest
//           level whose block holds k points;
//   bound   without U: the occupied base cells of the block are binned by their FARTHEST corner distance,
//           weighted by their point count; the bucket at which the count reaches k bounds the k-th distance;
//   prune   base cells whose NEAREST corner distance exceeds the bound are dropped; the survivors'
//           point ranges form the candidate set (typically a few dozen cells);
//   search  pcp_fk_search (knn_fast.cuh): FP32 histogram filter on coordinates relative to the query, exact
//           FP64 distances and (d2, index) ranking of the short list, certification against the block faces.
// If the exact k-th distance still reaches beyond the block, it becomes the new U and the level goes up.
#pragma once
#include "knn_fast.cuh"

#define PCP_FAR_NSEL 128   // selected base cells per attempt (256, with a 128-entry cell cache to stay at 4 blocks per SM: 1.4 k instead of 7 k queries overflow to the streaming kernel, but the kernel itself 11.6 -> 12.6 ms)
#define PCP_FAR_CCACHE 256 // cells whose (nearest-corner distance, first point, count) are kept between passes

// The cell cache and the staged candidates are never alive together (the cache carries the corner distances from the
// bound pass to the selection; staging starts after it), and they lie side by side -- [ccache | the base struct's
// candidate array] -- so the far kernel stages PCP_FAR_CAND = CCACHE + PCP_FK_CAND candidates at a time out of the
// shared memory of 256: a selection of 256..512 points (the mean is 283) is staged once, not once per pass.
#define PCP_FAR_CAND (PCP_FAR_CCACHE + PCP_FK_CAND)
template <int CAP> struct pcp_far_smem : pcp_fk_smem<CAP, 32> {
    uint32_t* selpre;   // [NSEL]  exclusive prefix of the selected cells' point counts (0xffffffff beyond)
    uint32_t* selstart; // [NSEL]  first point of the selected cell
    float4* ccache;     // [CCACHE] nearest-corner d2, first point, count of the block's cells (pass A -> B);
                        //          later the first CCACHE of the PCP_FAR_CAND staged candidates
    static __host__ __device__ size_t bytes(int k)
    {
        return pcp_fk_smem<CAP, 32>::bytes(k) + PCP_FAR_NSEL * 8 + PCP_FAR_CCACHE * 16;
    }
    __device__ pcp_far_smem(unsigned char* base, int k) : pcp_fk_smem<CAP, 32>(base + PCP_FAR_CCACHE * 16, k)
    {
        ccache = reinterpret_cast<float4*>(base);
        selpre = reinterpret_cast<uint32_t*>(base + PCP_FAR_CCACHE * 16 + pcp_fk_smem<CAP, 32>::bytes(k));
        selstart = selpre + PCP_FAR_NSEL;
    }
};

// One staging round of the far kernel: candidates c0 .. c0+CAND of the selected cells' points, as floats relative to
// the query.  A function of its own taking scalars (not a member: `this` would pin the source struct in local
// memory): it has three call sites, and the far kernel's code has to stay near the instruction cache's size.
static __device__ __noinline__ void pcp_far_stage(const pcp_rec* __restrict__ rec, float4* cand, const uint32_t* selpre,
                                                  const uint32_t* selstart, uint32_t T, double qx, double qy, double qz,
                                                  uint32_t c0, int lane)
{
    __syncwarp();
#pragma unroll 1
    for (int part = 0; part < PCP_FAR_CAND / 64 && c0 + part * 64 < T; part++) {
        pcp_rec r[2];
        uint32_t ps[2];
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const uint32_t fl = c0 + part * 64 + lane + 32 * i;
            ps[i] = 0;
            if (fl < T) {
                int c = 0;
#pragma unroll
                for (int step = PCP_FAR_NSEL / 2; step; step >>= 1)
                    if (selpre[c + step] <= fl) c += step;
                ps[i] = selstart[c] + (fl - selpre[c]);
                r[i] = pcp_load_rec(rec, ps[i]);
            }
        }
#pragma unroll
        for (int i = 0; i < 2; i++)
            if (c0 + part * 64 + lane + 32 * i < T)
                cand[part * 64 + lane + 32 * i] = make_float4((float)(r[i].x - qx), (float)(r[i].y - qy),
                                                               (float)(r[i].z - qz), __uint_as_float(ps[i]));
    }
    __syncwarp();
}

// candidates = the points of the selected cells, staged CAND at a time as floats relative to the query
struct pcp_far_src {
    const pcp_rec* rec;
    float4* cand;
    const uint32_t* selpre;
    const uint32_t* selstart;
    uint32_t T;
    double qx, qy, qz;
    bool resident; // T <= CAND and already staged

    __device__ __forceinline__ void stage(uint32_t c0, int lane) const
    {
        pcp_far_stage(rec, cand, selpre, selstart, T, qx, qy, qz, c0, lane);
    }
    template <class F> __device__ __forceinline__ void for_each(int lane, F f) const
    {
        for (uint32_t c0 = 0; c0 < T; c0 += PCP_FAR_CAND) {
            if (!resident) stage(c0, lane);
            const int n = (int)min((uint32_t)PCP_FAR_CAND, T - c0);
            for (int f0 = 0; f0 < n; f0 += 32) {
                const int i = f0 + lane;
                const bool valid = i < n;
                const float4 c = cand[valid ? i : 0];
                f(valid, c.x, c.y, c.z, __float_as_uint(c.w));
            }
        }
    }
};

// base-cell index range [ca, cb) of the level-`level` cell (cx,cy,cz) (binary search in the sorted cell keys)
__device__ __forceinline__ void pcp_level_cells(const unsigned long long* __restrict__ ckey, uint32_t ncell, int level,
                                                int cx, int cy, int cz, uint32_t& ca, uint32_t& cb)
{
    const unsigned long long pre = pcp_morton(cx, cy, cz);
    const int sh = 3 * level;
    const unsigned long long klo = sh >= 64 ? 0ull : (pre << sh);
    const unsigned long long khi = sh >= 64 ? PCP_EMPTY_KEY : ((pre + 1) << sh);
    uint32_t a = 0, b = ncell;
    while (a < b) {
        const uint32_t m = (a + b) >> 1;
        if (__ldg(ckey + m) < klo) a = m + 1; else b = m;
    }
    ca = a;
    b = ncell;
    while (a < b) {
        const uint32_t m = (a + b) >> 1;
        if (__ldg(ckey + m) < khi) a = m + 1; else b = m;
    }
    cb = a;
}

// Query position in cell units: integer cell + fraction in [0,1) (float), so that a cell's offset from the
// query is an exact integer difference minus the fraction.
struct pcp_qcell {
    int ix, iy, iz;
    float fx, fy, fz;
    float h;
};
// squared distance from q to the nearest / farthest point of base cell (cx,cy,cz), with outward margins
__device__ __forceinline__ void pcp_cell_dist2(const pcp_qcell& Q, int cx, int cy, int cz, float& mind2, float& maxd2)
{
    const float lo0 = ((float)(cx - Q.ix) - Q.fx) * Q.h, lo1 = ((float)(cy - Q.iy) - Q.fy) * Q.h,
                lo2 = ((float)(cz - Q.iz) - Q.fz) * Q.h;
    const float hi0 = lo0 + Q.h, hi1 = lo1 + Q.h, hi2 = lo2 + Q.h;
    const float n0 = fmaxf(fmaxf(lo0, -hi0), 0.f), n1 = fmaxf(fmaxf(lo1, -hi1), 0.f), n2 = fmaxf(fmaxf(lo2, -hi2), 0.f);
    const float x0 = fmaxf(fabsf(lo0), fabsf(hi0)), x1 = fmaxf(fabsf(lo1), fabsf(hi1)), x2 = fmaxf(fabsf(lo2), fabsf(hi2));
    const float slack = 4e-6f * Q.h; // float rounding of lo/hi and of the cell assignment
    const float m0 = fmaxf(n0 - slack, 0.f), m1 = fmaxf(n1 - slack, 0.f), m2 = fmaxf(n2 - slack, 0.f);
    mind2 = __fmaf_rn(m2, m2, __fmaf_rn(m1, m1, m0 * m0)) * 0.99999f;
    maxd2 = __fmaf_rn(x2 + slack, x2 + slack, __fmaf_rn(x1 + slack, x1 + slack, (x0 + slack) * (x0 + slack))) * 1.00001f;
}

// first histogram bucket (64 buckets, 2 per lane) at which the cumulative count reaches kk, or 64
__device__ __forceinline__ int pcp_hist_find32(const uint32_t* hist, int lane, uint32_t kk)
{
    const uint32_t h0 = hist[2 * lane], h1 = hist[2 * lane + 1];
    uint32_t incl = h0 + h1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const unsigned bal = __ballot_sync(0xffffffffu, incl >= kk);
    if (!bal) return PCP_FK_NB;
    const int L = __ffs(bal) - 1;
    const uint32_t c0L = __shfl_sync(0xffffffffu, incl - h1, L);
    return c0L >= kk ? 2 * L : 2 * L + 1;
}

struct pcp_far_args {
    const uint32_t* list;    // sorted positions of the queries
    const float* listU;      // upper bound of the k-th squared distance; +inf: none; -1: none, dense block
    const uint32_t* count;   // number of entries (device)
    uint32_t* work_counter;
    uint32_t* fail_list;     // queries for the streaming FP64 kernel (list overflow)
    uint32_t* fail_count;
    // external queries (pcp_knn): query qi is the point at ext_q + qi * ext_stride (x, y, z doubles), no list, no bound;
    // its neighbour row is row qi of nbr_out
    const char* ext_q = nullptr;
    size_t ext_stride = 0;
};

template <int CAP, bool DEFER = false>
static __global__ void __launch_bounds__(32 * PCP_FK_WARPS, 4) knn_far_kernel(pcp_fknn_params P, pcp_far_args A)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const pcp_far_smem<CAP> S(smem_raw + pcp_far_smem<CAP>::bytes(P.k) * warp, P.k);
    S.sqpos[lane] = 0xffffffffu;
    __syncwarp();
    int fill = 0;
    const uint32_t nq = *A.count;
    const float e1 = (float)(2.0 * P.epsd) * 1.0001f;
    const float INF = __int_as_float(0x7f800000);
    while (true) {
        uint32_t qi = 0;
        if (lane == 0) qi = atomicAdd(A.work_counter, 1u);
        qi = __shfl_sync(0xffffffffu, qi, 0);
        if (qi >= nq) break;
        const uint32_t qpos = A.ext_q ? qi : __ldg(A.list + qi);
        if (!A.ext_q && qpos == 0xffffffffu) continue; // settled by the third dense pass (knn_wide.cuh)
        float U = A.ext_q ? INF : __ldg(A.listU + qi);
        pcp_rec q;
        if (A.ext_q) {
            const double* qp = reinterpret_cast<const double*>(A.ext_q + (size_t)qi * A.ext_stride);
            q.x = qp[0], q.y = qp[1], q.z = qp[2], q.idx = 0;
            if (!(isfinite(q.x) && isfinite(q.y) && isfinite(q.z))) { // the streaming kernel writes the empty row
                if (lane == 0) A.fail_list[atomicAdd(A.fail_count, 1u)] = qpos;
                continue;
            }
        } else {
            q = pcp_load_rec(P.rec, qpos);
        }
        const int bx = min(max(pcp_cell_coord(q.x, P.g.org[0], P.g.inv_h, 0), 0), P.g.dims[0] - 1);
        const int by = min(max(pcp_cell_coord(q.y, P.g.org[1], P.g.inv_h, 0), 0), P.g.dims[1] - 1);
        const int bz = min(max(pcp_cell_coord(q.z, P.g.org[2], P.g.inv_h, 0), 0), P.g.dims[2] - 1);
        int level = (U == INF) ? 1 : 0;
        if (U < 0.f) { // PCP_FAR_HINT: no bound, start level given
            level = min((int)(-U - 1.0f), P.g.top_level);
            U = INF;
        }
        pcp_qcell Q;
        {
            const double tx = (q.x - P.g.org[0]) * P.g.inv_h, ty = (q.y - P.g.org[1]) * P.g.inv_h,
                         tz = (q.z - P.g.org[2]) * P.g.inv_h;
            Q.ix = bx, Q.iy = by, Q.iz = bz;
            Q.fx = (float)(tx - (double)bx), Q.fy = (float)(ty - (double)by), Q.fz = (float)(tz - (double)bz);
            Q.h = (float)P.g.h;
        }
        if (fill == 32) {
            pcp_fk_flush<pcp_far_smem<CAP>, DEFER>(P, S, lane);
            fill = 0;
        }
        const int slot = fill;
        int st = PCP_FK_BEYOND;
        int partial = 0; // attempts on a truncated cell selection (see below)
        for (; level <= P.g.top_level; level++) {
            const int lx = bx >> level, ly = by >> level, lz = bz >> level;
            const double g2 = pcp_guarantee2(P.g, level, lx, ly, lz, q.x, q.y, q.z);
            if (U != INF && !((double)U < g2) && level < P.g.top_level) continue; // this block cannot certify U
            // ---- the occupied base cells of the 27 level cells ----
            uint32_t ca = 0, cb = 0;
            if (lane < 27) {
                int dx, dy, dz;
                pcp_nb_offset(lane, dx, dy, dz);
                const int nx = lx + dx, ny = ly + dy, nz = lz + dz;
                const int mx = (P.g.dims[0] - 1) >> level, my = (P.g.dims[1] - 1) >> level, mz = (P.g.dims[2] - 1) >> level;
                bool take = nx >= 0 && ny >= 0 && nz >= 0 && nx <= mx && ny <= my && nz <= mz;
                if (take && U != INF) {
                    // a level cell the bound sphere does not reach holds no neighbour: its base cells need not be
                    // enumerated at all (a query near the middle of its cell keeps a few of the 27)
                    const double H = ldexp(P.g.h, level), sl = 1e-6 * H; // sl: rounding of the cell assignment
                    const double ax = P.g.org[0] + (double)nx * H, ay = P.g.org[1] + (double)ny * H,
                                 az = P.g.org[2] + (double)nz * H;
                    const double ex = fmax(fmax(ax - q.x, q.x - (ax + H)) - sl, 0.0),
                                 ey = fmax(fmax(ay - q.y, q.y - (ay + H)) - sl, 0.0),
                                 ez = fmax(fmax(az - q.z, q.z - (az + H)) - sl, 0.0);
                    take = ex * ex + ey * ey + ez * ez <= (double)U * 1.000001;
                }
                if (take) {
                    if (level < PCP_HLEVELS) {
                        uint32_t cn = 0;
                        if (pcp_hash_find(P.ltab, P.lmask, PCP_LKEY(level, pcp_morton(nx, ny, nz)), ca, cn)) cb = ca + cn;
                    } else {
                        pcp_level_cells(P.ckey, P.ncell, level, nx, ny, nz, ca, cb);
                    }
                }
            }
            uint32_t cpre = cb - ca;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, cpre, o);
                if (lane >= o) cpre += v;
            }
            const uint32_t NC = __shfl_sync(0xffffffffu, cpre, 31);
            PCP_FK_DBG(10, 1);
            PCP_FK_DBG(11, NC);
            __syncwarp();
            S.tab[lane] = lane < 27 ? cpre - (cb - ca) : 0xffffffffu;
            S.tab[32 + lane] = ca;
            __syncwarp();
            float Ueff = U;
            if (U == INF) {
                uint32_t tl = (cb > ca) ? __ldg(P.cstart + cb) - __ldg(P.cstart + ca) : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) tl += __shfl_xor_sync(0xffffffffu, tl, o);
                if (tl < (uint32_t)P.kk && level < P.g.top_level) {
                    PCP_FK_DBG(16, 1);
                    continue;
                }
                // ---- bound: cells binned by farthest-corner distance, weighted by point count ----
                const double H = ldexp(P.g.h, level);
                const float topA = (float)(27.0 * H * H) * 1.001f;
                const int key0 = (__float_as_int(topA) >> 21) - (PCP_FK_NB - 2);
                S.hist[lane] = 0;
                S.hist[lane + 32] = 0;
                __syncwarp();
                for (uint32_t c0 = 0; c0 < NC; c0 += 32) {
                    const uint32_t fc = c0 + lane;
                    if (fc < NC) {
                        int t = 0;
#pragma unroll
                        for (int step = 16; step; step >>= 1)
                            if (S.tab[t + step] <= fc) t += step;
                        const uint32_t ci = S.tab[32 + t] + (fc - S.tab[t]);
                        const uint4 ce = __ldg(P.cinfo + ci);
                        const uint32_t cnt = __ldg(P.cstart + ci + 1) - ce.w;
                        float mn, mx;
                        pcp_cell_dist2(Q, (int)ce.x, (int)ce.y, (int)ce.z, mn, mx);
                        atomicAdd(&S.hist[pcp_fk_bucket(mx, key0)], cnt);
                        if (fc < PCP_FAR_CCACHE) S.ccache[fc] = make_float4(mn, __uint_as_float(ce.w), __uint_as_float(cnt), 0.f);
                    }
                }
                __syncwarp();
                const int bA = pcp_hist_find32(S.hist, lane, (uint32_t)P.kk);
                Ueff = bA >= PCP_FK_NB - 1 ? topA : pcp_fk_edge(bA, key0);
                __syncwarp();
            }
            // ---- prune: keep the base cells the sphere of radius sqrt(Ueff) touches ----
            // (inside = true: only the cells that lie entirely inside the sphere)
            auto select_cells = [&](float bound, bool inside, bool cached) -> uint32_t {
                uint32_t nsel = 0;
                __syncwarp();
                for (int i = lane; i < PCP_FAR_NSEL; i += 32) S.selpre[i] = 0xffffffffu;
                __syncwarp();
                for (uint32_t c0 = 0; c0 < NC; c0 += 32) {
                    const uint32_t fc = c0 + lane;
                    bool sel = false;
                    uint32_t cs = 0, cc = 0;
                    if (fc < NC) {
                        if (cached && fc < PCP_FAR_CCACHE) {
                            const float4 e = S.ccache[fc];
                            sel = e.x <= bound;
                            cs = __float_as_uint(e.y), cc = __float_as_uint(e.z);
                        } else {
                            int t = 0;
#pragma unroll
                            for (int step = 16; step; step >>= 1)
                                if (S.tab[t + step] <= fc) t += step;
                            const uint32_t ci = S.tab[32 + t] + (fc - S.tab[t]);
                            const uint4 ce = __ldg(P.cinfo + ci);
                            float mn, mx;
                            pcp_cell_dist2(Q, (int)ce.x, (int)ce.y, (int)ce.z, mn, mx);
                            sel = (inside ? mx : mn) <= bound;
                            if (sel) cs = ce.w, cc = __ldg(P.cstart + ci + 1) - cs;
                        }
                    }
                    const unsigned mask = __ballot_sync(0xffffffffu, sel);
                    const uint32_t at = nsel + __popc(mask & ((1u << lane) - 1u));
                    if (sel && at < PCP_FAR_NSEL) S.selstart[at] = cs, S.selpre[at] = cc;
                    nsel += __popc(mask);
                }
                __syncwarp();
                return nsel;
            };
            // exclusive prefix of the selected cells' counts; returns the number of candidates
            auto prefix_counts = [&](uint32_t nsel) -> uint32_t {
                uint32_t T = 0;
                for (uint32_t c0 = 0; c0 < nsel; c0 += 32) {
                    const uint32_t i = c0 + lane;
                    const uint32_t cnt = i < nsel ? S.selpre[i] : 0u;
                    uint32_t inc = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
                        if (lane >= o) inc += v;
                    }
                    if (i < nsel) S.selpre[i] = T + inc - cnt;
                    T += __shfl_sync(0xffffffffu, inc, 31);
                }
                __syncwarp();
                return T;
            };
            // Too many cells for the corner bound (a sparse query next to a dense surface): the cells that lie
            // entirely inside that sphere hold k points by construction, and their exact k-th distance is a much
            // tighter bound; search those, take it and run this level again.
            uint32_t nsel = 0;
            bool inner = false, truncated = false;
#pragma unroll 1
            for (int pass = 0; pass < 2; pass++) {
                nsel = select_cells(Ueff, pass == 1, pass == 0 && U == INF); // pass A filled the cache
                if (nsel <= PCP_FAR_NSEL || U != INF) break;
                inner = pass == 0;
            }
            if (nsel > PCP_FAR_NSEL) {
                // Still too many (a point metres above a dense surface: the corner bound's sphere cuts a disc of
                // hundreds of cells out of it).  ANY k points bound the k-th distance: search the cells that did fit,
                // take their exact k-th as the bound -- now the true scale of the neighbourhood, a handful of cells
                // -- and run this level again.  Twice at most; then the streaming kernel.
                if (partial >= 2) {
                    PCP_FK_DBG(23, 1);
                    st = PCP_FK_OVERFLOW;
                    break;
                }
                partial++;
                nsel = PCP_FAR_NSEL;
                inner = true;
                truncated = true;
            }
            const uint32_t T = prefix_counts(nsel);
            PCP_FK_DBG(17, inner);
            PCP_FK_DBG(12, 1);
            PCP_FK_DBG(13, T);
            PCP_FK_DBG(14, nsel);
            pcp_far_src src{P.rec, S.ccache, S.selpre, S.selstart, T, q.x, q.y, q.z, T <= PCP_FAR_CAND};
            if (src.resident) src.stage(0, lane);
            double dk;
            // (the points of a cell whose nearest corner lies inside the sphere reach a cell diagonal beyond it: the
            // truncated selection is searched up to there, or its k-th may fall off the histogram)
            const float rtop = sqrtf(Ueff) + 1.8f * (float)P.g.h;
            const float top = truncated ? rtop * rtop * 1.001f : Ueff * 1.001f;
            st = pcp_fk_search<32, CAP>(src, S, P.rec, lane, true, slot, P.kk, 0.f, 0.f, 0.f, q.x, q.y, q.z, top,
                                        inner ? 0.0 : g2, e1, dk);
            if (inner) {
#ifdef PCP_FAR_TRACE
                if (st != PCP_FK_UNCERT && lane == 0) printf("far inner: qpos %u st %d level %d Ueff %g T %u nsel %u truncated %d dk %g\n", qpos, st, level, (double)Ueff, T, nsel, (int)truncated, dk);
#endif
                if (st != PCP_FK_UNCERT) {
                    st = PCP_FK_OVERFLOW;
                    break;
                }
                const float Unew = __double2float_ru(dk) * 1.000001f;
                if (truncated && !(Unew < Ueff)) { // no tighter than the corner bound: nothing gained by another round
                    st = PCP_FK_OVERFLOW;
                    break;
                }
                U = Unew;
                level--;
                continue;
            }
            PCP_FK_DBG(15, st == PCP_FK_UNCERT);
            PCP_FK_DBG(6, !((double)Ueff < g2));
            PCP_FK_DBG(7, !((double)Ueff < g2) && st == PCP_FK_OK);
            if (st == PCP_FK_OK) U = __double2float_ru(dk);
            if (st == PCP_FK_OK || st == PCP_FK_OVERFLOW) break;
            if (st == PCP_FK_UNCERT) U = __double2float_ru(dk) * 1.000001f; // exact k-th of the scanned set: a true bound
        }
        if (st == PCP_FK_OK) {
            S.sqpos[slot] = qpos;
            S.skd[slot] = U; // set below from the exact k-th distance of the successful attempt
            fill++;
        } else {
            PCP_FK_DBG(18, st == PCP_FK_OVERFLOW);
            PCP_FK_DBG(19, st == PCP_FK_BEYOND);
            PCP_FK_DBG(20, st == PCP_FK_UNCERT);
            PCP_FK_DBG(21, U == INF);
            PCP_FK_DBG(22, level);
#ifdef PCP_FAR_TRACE
            if (lane == 0) printf("far fail: qpos %u st %d level %d U %g partial %d q=(%.3f %.3f %.3f)\n", qpos, st, level, (double)U, partial, q.x, q.y, q.z);
#endif
            if (lane == 0) A.fail_list[atomicAdd(A.fail_count, 1u)] = qpos;
        }
    }
    pcp_fk_flush<pcp_far_smem<CAP>, DEFER>(P, S, lane);
}
