// knn_fast.cuh -- the hot kernel: exact kNN(k) + plane-fit features in one pass over the Morton-sorted cloud.
//
// Persistent warps pull "items" (up to 32 consecutive queries of one base cell) from a global counter.
//   stage    the FP32 copies of the points of the 3x3x3 block of cells around the item's cell are staged
//            once per item in shared memory (all queries of the item share them);
//   search   a query is handled by a group of G lanes (G = 8: four queries of the item in flight per warp;
//            G = 32 for sparse cells and for the level escalation):
//     pass 1   the lanes evaluate candidates in FP32 (coordinates relative to the grid origin) and bin the
//              squared distances into a 64-bucket shared-memory histogram; a group scan finds the bucket b*
//              at which the cumulative count reaches k;
//     pass 2   candidates with d2 <= edge(b*+1) + 3*delta are compacted (ballot) into a short list, delta
//              being a rigorous bound on the FP32 error, so the list provably contains the true k nearest;
//     refine   the listed candidates get their exact FP64 squared distance (FLANN L2_Simple order, no FMA)
//              and are ranked by (d2, original index); ranks < k are the answer, already in order;
//     certify  edge(b*+1) + 2*delta must lie inside the distance from the query to the faces of the
//              scanned block, otherwise the query moves up one level (cell edge doubles) and scans point
//              ranges straight from global memory with the whole warp;
//   features default (DEFER): the neighbour positions of a settled query are written as a row of a scratch in
//            global memory and features_from_list_kernel (knn_warp.cuh) fits the planes afterwards -- the search
//            kernels stay small enough for the instruction cache and need no parked lists in shared memory.
//            Fused variant (PCP_FK_FUSED=1, kept for comparison): finished queries are parked in 32 shared-memory
//            slots; when the slots are full every lane takes one query and runs calculate_plan_parameter_h_points
//            (calculate_feature.cpp:119-206) on its k neighbours in ascending-distance order.
// Queries that overflow the list (mass ties / duplicates) go to the streaming FP64 kernel in knn_warp.cuh.
#pragma once
#include "knn_cell.cuh"

#define PCP_FK_NB 64      // histogram buckets of the exact search (last one = beyond range)
#define PCP_FD_NB 32      // ... of the dense kernel: 8 octaves below the certifiable radius, nearer ones share bucket 0
#define PCP_FK_CAND 256   // staged candidates per item
#define PCP_FK_WARPS 4
#define PCP_FK_STAGE_U 2  // candidates a lane loads before it converts and stores them
#ifndef PCP_FK_MINB
#define PCP_FK_MINB 6
#endif

// lists the FP32 search could not order (see pcp_fd_search): entries of PCP_RR_HEAD + CAP words
struct pcp_rr {
    uint32_t* list = nullptr;
    uint32_t* count = nullptr;
    uint32_t cap = 0;
};

struct pcp_fknn_params {
    const pcp_rec* rec;
    const float4* fpt; // (float)(p - grid origin); w unused
    const pcp_hslot* htab;
    uint32_t hmask;
    const pcp_hslot* ltab; // level table (index.cuh)
    uint32_t lmask;
    const uint4* cinfo;    // per base cell: x, y, z, first point
    const unsigned long long* ckey;
    const uint32_t* cstart;
    uint32_t ncell;
    const uint32_t* item_cell;
    const uint32_t* item_first;
    const uint32_t* item_meta; // level << 8 | queries
    uint32_t nitem;
    pcp_grid g;
    int k, kk;
    double eps0;           // bound on |float coordinate - true coordinate relative to the origin|
    double epsd;           // bound on the rounding of a double subtraction p - origin
    long long query_limit; // points with original index >= query_limit are candidates only (halo points)
    void* props;
    double* eig3;
    float* lps3;
    int32_t* knn_idx;
    float* kth_d2;           // optional: upper bound of every query's k-th squared distance (shard boundary test)
    uint32_t* fail_list;     // far list: sorted positions ...
    float* fail_u;           // ... and the known upper bound of the k-th squared distance (+inf: none)
    uint32_t* fail_count;
    uint32_t* work_counter;  // item dispenser
    // second pass over the items whose 3x3x3 block held more points than the first pass stages: the first pass
    // appends them to dense_list (when set) instead of sending their queries to the far list; the second pass
    // (a kernel instance with a larger CAND) reads its items from item_list[0 .. *nitem_ptr)
    const uint32_t* item_list = nullptr;
    const uint32_t* nitem_ptr = nullptr;
    uint32_t* dense_list = nullptr;
    uint32_t* dense_count = nullptr;
    // deferred features (kernel instances with DEFER): the flush writes the neighbour POSITIONS of the settled queries
    // to nbr_out[sorted position * k ...] and a separate thread-per-query kernel fits the planes.  The fused flush
    // (gathers + FP64 scatter + Jacobi, ~2200 instructions = 35 KB) shares no code with the search loops, and warps
    // of one SM sit in both at once: the instruction working set exceeded the 32 KB L1.5 instruction cache and
    // 'no instruction' was the top stall of the fused kernels (6.5 cycles per issue).
    uint32_t* nbr_out = nullptr;
    pcp_rr rr;               // re-rank list (deferred rows only)
    // the third pass (knn_wide.cuh): one record per leaf that put queries on the far list -- x: its first entry there,
    // y: entries | level << 8
    uint2* wide_list = nullptr;
    uint32_t* wide_count = nullptr;
    uint32_t* wide_work = nullptr;
    unsigned long long* dbg; // optional counters (PCP_KNN_DEBUG=1), NULL in production
};

#define PCP_FK_DBG(i, v)                                                                                     \
    do {                                                                                                     \
        if (P.dbg && lane == 0) atomicAdd(P.dbg + (i), (unsigned long long)(v));                             \
    } while (0)

// per-warp shared memory.  CAP = list capacity per query; the staged path runs 32/G queries at once, each
// with its own histogram and list, the full-warp path uses set 0.
template <int CAP, int GS> struct pcp_fk_smem {
    static constexpr int NSET = 32 / GS;
    float4* cand;    // [PCP_FK_CAND]
    double* ld2;     // [NSET][CAP]
    uint32_t* hist;  // [NSET][PCP_FK_NB]
    uint32_t* lpos;  // [NSET][CAP]
    int* lidx;       // [NSET][CAP]
    uint32_t* nbr;   // [k][32]   neighbour positions of the parked queries
    uint32_t* sqpos; // [32]      sorted position of the parked query (0xffffffff = empty slot)
    uint32_t* tab;   // [64]      flat-index table of the 27 cell ranges: prefix[32], start[32]
    float* skd;      // [32]      k-th squared distance (upper bound) of the parked query
    static __host__ __device__ size_t bytes(int k)
    {
        return (size_t)PCP_FK_CAND * 16 + (size_t)NSET * CAP * 16 + (size_t)NSET * PCP_FK_NB * 4 + (size_t)32 * k * 4 + 128 + 256 + 128;
    }
    __device__ pcp_fk_smem(unsigned char* base, int k)
    {
        cand = reinterpret_cast<float4*>(base);
        ld2 = reinterpret_cast<double*>(base + PCP_FK_CAND * 16);
        hist = reinterpret_cast<uint32_t*>(base + PCP_FK_CAND * 16 + NSET * CAP * 8);
        lpos = hist + NSET * PCP_FK_NB;
        lidx = reinterpret_cast<int*>(lpos + NSET * CAP);
        nbr = reinterpret_cast<uint32_t*>(lidx + NSET * CAP);
        sqpos = nbr + 32 * k;
        tab = sqpos + 32;
        skd = reinterpret_cast<float*>(tab + 64);
    }
};

// candidate sources: every lane of the warp must call f(valid, x, y, z, pos) the same number of times
template <int G> struct pcp_fk_staged {
    const float4* cand;
    int T;
    template <class F> __device__ __forceinline__ void for_each(int gl, F f) const
    {
        for (int f0 = 0; f0 < T; f0 += G) {
            const int i = f0 + gl;
            const bool valid = i < T;
            const float4 c = cand[valid ? i : 0];
            f(valid, c.x, c.y, c.z, __float_as_uint(c.w));
        }
    }
};

// Full warp: the points of 27 cell ranges, flattened and staged through shared memory CAND at a time with
// PCP_FK_CAND/32 independent loads in flight per lane (one memory latency per chunk instead of one per round).
// tab[0..31] = exclusive prefix of the range sizes (0xffffffff beyond the 27th), tab[32..63] = range starts.
struct pcp_fk_chunked {
    const float4* fpt;
    float4* cand;
    const uint32_t* tab;
    uint32_t T;
    template <class F> __device__ __forceinline__ void for_each(int lane, F f) const
    {
        constexpr int R = PCP_FK_CAND / 32;
        for (uint32_t c0 = 0; c0 < T; c0 += PCP_FK_CAND) {
            __syncwarp();
            float4 v[R];
            uint32_t ps[R];
#pragma unroll
            for (int i = 0; i < R; i++) {
                const uint32_t fl = c0 + lane + 32 * i;
                int nb = 0;
#pragma unroll
                for (int step = 16; step; step >>= 1)
                    if (tab[nb + step] <= fl) nb += step;
                ps[i] = tab[32 + nb] + (fl - tab[nb]);
                if (fl < T) v[i] = __ldg(fpt + ps[i]);
            }
#pragma unroll
            for (int i = 0; i < R; i++) {
                v[i].w = __uint_as_float(ps[i]);
                if (c0 + lane + 32 * i < T) cand[lane + 32 * i] = v[i];
            }
            __syncwarp();
            const int n = (int)min((uint32_t)PCP_FK_CAND, T - c0);
            for (int f0 = 0; f0 < n; f0 += 32) {
                const int i = f0 + lane;
                const bool valid = i < n;
                const float4 c = cand[valid ? i : 0];
                f(valid, c.x, c.y, c.z, __float_as_uint(c.w));
            }
        }
    }
};

// status of one search attempt
#define PCP_FK_OK 0
#define PCP_FK_BEYOND 1   // the k-th neighbour lies beyond the histogram range
#define PCP_FK_UNCERT 2   // found, but not provably inside the scanned block
#define PCP_FK_OVERFLOW 3 // more candidates than the list holds (ties / duplicates)
#define PCP_FK_IDLE 4     // the group had no query

// Logarithmic distance buckets, 4 per octave of d2, read straight off the float bits: no range to guess.
// key(d2) = bits >> 21; bucket = clamp(key - (key(top) - 62), 0, 63); bucket 63 = beyond `top`.
__device__ __forceinline__ int pcp_fk_bucket(float d2, int key0)
{
    return min(max((__float_as_int(d2) >> 21) - key0, 0), PCP_FK_NB - 1);
}
__device__ __forceinline__ int pcp_fd_bucket(float d2, int key0)
{
    return min(max((__float_as_int(d2) >> 21) - key0, 0), PCP_FD_NB - 1);
}
// exclusive upper edge of bucket b (b <= 62)
__device__ __forceinline__ float pcp_fk_edge(int b, int key0) { return __int_as_float((key0 + b + 1) << 21); }

// One search attempt (warp-collective; every group of G lanes works on its own query, or idles).
// `top`: squared distance beyond which candidates are of no interest (they land in the last bucket).
// On PCP_FK_OK the kk nearest (sorted positions) are in nbr[e*32 + slot], ascending by (d2, original index),
// and the exact k-th squared distance is < g2.  On PCP_FK_UNCERT the list holds the kk nearest of the
// scanned candidates and dk is their exact k-th squared distance: an upper bound of the true one.
template <int G, int CAP, int GS, class Src>
__device__ __forceinline__ int pcp_fk_search(const Src& src, const pcp_fk_smem<CAP, GS>& S, const pcp_rec* __restrict__ rec,
                                             int lane, bool active, int slot, int kk, float qfx, float qfy, float qfz,
                                             double qx, double qy, double qz, float top, double g2, float e1, double& dk)
{
    constexpr int SL = CAP / G;        // list entries per lane
    constexpr int NBL = PCP_FK_NB / G; // histogram buckets per lane
    const int gl = lane & (G - 1), gbase = lane & ~(G - 1), set = (G == 32) ? 0 : lane / G;
    const unsigned gmask = (G == 32) ? 0xffffffffu : ((1u << G) - 1u);
    uint32_t* hist = S.hist + set * PCP_FK_NB;
    uint32_t* lpos = S.lpos + set * CAP;
    int* lidx = S.lidx + set * CAP;
    double* ld2 = S.ld2 + set * CAP;
    const int key0 = (__float_as_int(top) >> 21) - (PCP_FK_NB - 2);
#pragma unroll
    for (int i = 0; i < NBL; i++) hist[gl + i * G] = 0;
    __syncwarp();
    // ---- pass 1: histogram ----
    src.for_each(gl, [&](bool valid, float x, float y, float z, uint32_t) {
        const float dx = qfx - x, dy = qfy - y, dz = qfz - z;
        const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
        if (valid && active) atomicAdd(&hist[pcp_fk_bucket(d2, key0)], 1u);
    });
    __syncwarp();
    // ---- scan: first bucket at which the cumulative count reaches kk ----
    uint32_t hb[NBL];
    if constexpr (NBL == 8) {
        const uint4 a = *reinterpret_cast<const uint4*>(hist + gl * 8);
        const uint4 b = *reinterpret_cast<const uint4*>(hist + gl * 8 + 4);
        hb[0] = a.x, hb[1] = a.y, hb[2] = a.z, hb[3] = a.w, hb[4] = b.x, hb[5] = b.y, hb[6] = b.z, hb[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < NBL; i++) hb[i] = hist[gl * NBL + i];
    }
    uint32_t tot = 0;
#pragma unroll
    for (int i = 0; i < NBL; i++) tot += hb[i];
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < G; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o, G);
        if (gl >= o) incl += v;
    }
    int st = active ? PCP_FK_OK : PCP_FK_IDLE;
    const unsigned bal = (__ballot_sync(0xffffffffu, incl >= (uint32_t)kk) >> gbase) & gmask;
    if (!bal && active) st = PCP_FK_BEYOND;
    const int L = bal ? __ffs(bal) - 1 : 0;
    int myb = PCP_FK_NB;
    {
        uint32_t c = incl - tot;
#pragma unroll
        for (int i = 0; i < NBL; i++) {
            c += hb[i];
            if (myb == PCP_FK_NB && c >= (uint32_t)kk) myb = gl * NBL + i;
        }
    }
    const int bstar = __shfl_sync(0xffffffffu, myb, gbase + L);
    if (st == PCP_FK_OK && bstar >= PCP_FK_NB - 1) st = PCP_FK_BEYOND;
    // every true kk-nearest has FP32 d2 <= edge + 3*delta (delta: FP32 error bound at that distance)
    float edge = pcp_fk_edge(min(bstar, PCP_FK_NB - 2), key0);
    if constexpr (G == 32) {
        // A bucket spans 19% of d2; behind a gap in the distances (a sparse query over a dense surface) it can
        // hold far more than the list does.  Cut that bucket into 64 linear slices and take the slice at which
        // the count reaches kk.
        uint32_t below = 0, upto = 0; // candidates in the buckets before b*, and up to and including it
        {
            uint32_t c = incl - tot;
#pragma unroll
            for (int i = 0; i < NBL; i++) {
                if (gl * NBL + i == bstar) below = c, upto = c + hb[i];
                c += hb[i];
            }
            below = __shfl_sync(0xffffffffu, below, L);
            upto = __shfl_sync(0xffffffffu, upto, L);
        }
        // (Up to three times: a point metres above a dense plane sees hundreds of points at distances that differ
        // in the fourth digit -- one slice of the bucket, 0.3 % of d2, still holds more than the list.)
        float lo = bstar ? pcp_fk_edge(bstar - 1, key0) : 0.0f;
#pragma unroll 1
        // (+ 4: the list takes the candidates up to edge + 3 delta -- a few more than the buckets count)
        for (int round = 0; round < 3 && st == PCP_FK_OK && upto + 4 > (uint32_t)CAP; round++) {
            const float w = (edge - lo) * (1.0f / PCP_FK_NB), iw = 1.0f / w;
            if (!(w > 0.0f)) break;
            __syncwarp();
            hist[lane] = 0, hist[lane + 32] = 0;
            __syncwarp();
            uint32_t nb = 0; // candidates below the interval, counted again (round 0: the buckets before b*)
            src.for_each(gl, [&](bool valid, float x, float y, float z, uint32_t) {
                const float dx = qfx - x, dy = qfy - y, dz = qfz - z;
                const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
                if (valid && d2 < lo) nb++;
                if (valid && d2 >= lo && d2 < edge)
                    atomicAdd(&hist[min(max((int)((d2 - lo) * iw), 0), PCP_FK_NB - 1)], 1u);
            });
            __syncwarp();
            below = __reduce_add_sync(0xffffffffu, nb);
            if (below >= (uint32_t)kk) break; // (cannot happen; the interval as it is holds the k-th in any case)
            const uint32_t s0 = hist[2 * lane], s1 = hist[2 * lane + 1];
            uint32_t si = s0 + s1;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, si, o);
                if (lane >= o) si += v;
            }
            const uint32_t need = (uint32_t)kk - below; // >= 1
            const unsigned sb = __ballot_sync(0xffffffffu, si >= need);
            if (!sb) break;
            const int SLn = __ffs(sb) - 1;
            const uint32_t before = __shfl_sync(0xffffffffu, si - s1, SLn); // cumulative through slice 2*SLn
            const uint32_t through = __shfl_sync(0xffffffffu, si, SLn);     // ... through slice 2*SLn + 1
            const int slice = before >= need ? 2 * SLn : 2 * SLn + 1;
            upto = below + (before >= need ? before : through);
            // slice(d2) <= slice for every d2 below the new edge (and a little beyond: harmless); slice(d2) < slice
            // for every d2 below the new lower end
            edge = fminf(edge, __fmaf_rn((float)(slice + 1), w, lo) * 1.000002f);
            lo = __fmaf_rn((float)slice * 0.999998f, w, lo);
        }
    }
    const float delta = 4.0f * sqrtf(edge) * e1 + 4.0f * e1 * e1 + 1e-6f * edge;
    const float thr = (edge + 3.0f * delta) * 1.000001f;
    bool live = st == PCP_FK_OK;
    // ---- pass 2: compact every candidate that can be among the kk nearest ----
    uint32_t m = 0;
    src.for_each(gl, [&](bool valid, float x, float y, float z, uint32_t pos) {
        const float dx = qfx - x, dy = qfy - y, dz = qfz - z;
        const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
        const bool in = live && valid && d2 <= thr;
        const unsigned mask = (__ballot_sync(0xffffffffu, in) >> gbase) & gmask;
        const uint32_t at = m + __popc(mask & ((1u << gl) - 1u));
        if (in && at < (uint32_t)CAP) lpos[at] = pos;
        m += __popc(mask);
    });
#ifdef PCP_FAR_TRACE
    if constexpr (G == 32) {
        if (live && (m > (uint32_t)CAP || m < (uint32_t)kk) && lane == 0)
            printf("fk overflow: m %u kk %d edge %g thr %g top %g bstar %d\n", m, kk, (double)edge, (double)thr, (double)top, bstar);
    }
#endif
    if (live && (m > (uint32_t)CAP || m < (uint32_t)kk)) st = PCP_FK_OVERFLOW, live = false; // (m < kk: cannot happen)
    __syncwarp();
    // ---- refine: exact squared distances, rank by (d2, original index) ----
    double d2s[SL];
    int ids[SL];
    uint32_t ps[SL];
#pragma unroll
    for (int s = 0; s < SL; s++) {
        const int e = s * G + gl;
        d2s[s] = DBL_MAX, ids[s] = 0x7fffffff, ps[s] = 0;
        if (live && e < (int)m) {
            ps[s] = lpos[e];
            const pcp_rec r = pcp_load_rec(rec, ps[s]);
            d2s[s] = pcp_dist2(qx, qy, qz, r.x, r.y, r.z);
            ids[s] = (int)r.idx;
            ld2[e] = d2s[s];
            lidx[e] = ids[s];
        }
    }
    __syncwarp();
    const uint32_t mmax = __reduce_max_sync(0xffffffffu, live ? m : 0u);
    int rank[SL];
#pragma unroll
    for (int s = 0; s < SL; s++) rank[s] = 0;
    for (uint32_t j = 0; j < mmax; j++) {
        const bool have = j < m;
        const double dj = ld2[have ? j : 0];
        const int ij = lidx[have ? j : 0];
#pragma unroll
        for (int s = 0; s < SL; s++) rank[s] += (have && (dj < d2s[s] || (dj == d2s[s] && ij < ids[s]))) ? 1 : 0;
    }
    double dkl = -1.0;
#pragma unroll
    for (int s = 0; s < SL; s++)
        if (live && s * G + gl < (int)m && rank[s] < kk) {
            S.nbr[rank[s] * 32 + slot] = ps[s];
            if (rank[s] == kk - 1) dkl = d2s[s];
        }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) dkl = fmax(dkl, __shfl_xor_sync(0xffffffffu, dkl, o));
    dk = dkl;
    if (live && !(dkl < g2)) st = PCP_FK_UNCERT;
    __syncwarp();
    return st;
}

// features of the parked queries: lane = slot
template <class SM, bool DEFER = false>
__device__ __forceinline__ void pcp_fk_flush(const pcp_fknn_params& P, const SM& S, int lane)
{
    __syncwarp();
    const uint32_t qpos = S.sqpos[lane];
    if constexpr (DEFER) {
        // rows of k positions, one slot after the other: lanes = entries, so a row is written as one segment
#pragma unroll 1
        for (int sl = 0; sl < 32; sl++) {
            const uint32_t qp = S.sqpos[sl];
            if (qp == 0xffffffffu) continue;
            for (int e = lane; e < P.k; e += 32)
                P.nbr_out[(size_t)qp * P.k + e] = e < P.kk ? S.nbr[e * 32 + sl] : 0xffffffffu;
        }
        __syncwarp(); // (kth_d2: the features kernel writes the exact k-th distance, a tighter bound than S.skd)
        S.sqpos[lane] = 0xffffffffu;
        __syncwarp();
        return;
    }
    if (qpos != 0xffffffffu) {
        const size_t jq = (size_t)P.rec[qpos].idx;
        const int N = P.kk;
        if (P.knn_idx) {
            for (int e = 0; e < P.k; e++)
                P.knn_idx[jq * P.k + e] = e < N ? (int32_t)P.rec[S.nbr[e * 32 + lane]].idx : -1;
        }
        pcp_plane pl;
        const bool fitted = N > 3; // calculate_feature.cpp:238
        if (fitted) {
            double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
            for (int e = 0; e < N; e++) {
                const pcp_rec r = pcp_load_rec(P.rec, S.nbr[e * 32 + lane]);
                XAve += r.x;
                YAve += r.y;
                ZAve += r.z;
            }
            XAve /= (double)N;
            YAve /= (double)N;
            ZAve /= (double)N;
            double s00 = 0, s01 = 0, s02 = 0, s11 = 0, s12 = 0, s22 = 0;
            for (int e = 0; e < N; e++) {
                const pcp_rec r = pcp_load_rec(P.rec, S.nbr[e * 32 + lane]);
                const double X = r.x - XAve, Y = r.y - YAve, Z = r.z - ZAve;
                s00 += X * X;
                s01 += X * Y;
                s02 += X * Z;
                s11 += Y * Y;
                s12 += Y * Z;
                s22 += Z * Z;
            }
            pcp_plane_from_scatter(s00, s01, s02, s11, s12, s22, XAve, YAve, ZAve, pl);
        }
        pcp_store_props(P.props, jq, fitted, pl);
        pcp_store_extra(P.eig3, P.lps3, jq, fitted, pl);
        if (P.kth_d2) P.kth_d2[jq] = S.skd[lane];
    }
    S.sqpos[lane] = 0xffffffffu;
    __syncwarp();
}

// ---- the dense kernel ------------------------------------------------------------------------------------
// FP32 only, coordinates relative to the item's cell origin (error ~1e-7 h, see pcp_fd_e1).  Queries whose FP32
// ordering is not provably the exact one (two candidates closer in d2 than the error bound, or a k-th distance
// too close to the certified radius) are not settled here: they go to the far list with their upper bound.
// CAND: staged candidates per item.  320 candidates and lists for 4 queries of 32 entries come to 6656 bytes per
// warp when the neighbour rows go straight to global memory (DEFER): 8 blocks of 4 warps per SM; with the parked
// lists of the fused variant (k = 20) 9472 bytes and 6 blocks.  (The 4 query groups of a warp read the same 8
// candidate pairs per step: two 128-byte wavefronts, LDS.128 + LDS.64, for 16 candidates.)
template <int CAP, int GS, int CAND, bool DEFER = false> struct pcp_fd_smem {
    static constexpr int NSET = 32 / GS;
    static_assert(CAND % (2 * GS) == 0 && 2 * CAND + 4 < 65536 && (NSET * CAP) % 2 == 0, "pcp_fd_smem: layout");
    __device__ __forceinline__ void put(uint32_t i, float x, float y, float z, uint32_t pos) const
    {
        float* a = reinterpret_cast<float*>(cand) + 4 * (i >> 1) + (i & 1u);
        a[0] = x, a[2] = y, a[2 * CAND] = z, a[2 * CAND + 2] = __uint_as_float(pos);
    }
    // the list keeps a candidate as the word offset of its position inside the second half: 4 * pair + 2 + (0 | 1)
    __device__ __forceinline__ uint32_t pos_at(uint32_t word) const
    {
        return __float_as_uint(reinterpret_cast<const float*>(cand)[2 * CAND + word]);
    }
    float4* cand;    // [CAND]  candidates in PAIRS for the packed FP32 pipe: cand[p] = (x0, x1, y0, y1) and
                     //         cand[CAND/2 + p] = (z0, z1, pos0, pos1) of candidates 2p, 2p+1; x, y, z relative to the
                     //         leaf origin, pos = sorted position
    uint32_t* hist;  // [NSET][PCP_FD_NB]
    float* ld2;      // [NSET][CAP]
    uint16_t* lpos;  // [NSET][CAP]  candidate (see pos_at)
    uint32_t* tab;   // [64]  exclusive prefix of the staged runs' lengths (0xffffffff beyond the 27), first points
    uint32_t* nbr;   // [k][32]  (DEFER: none of the three -- the rows go straight to global memory)
    uint32_t* sqpos; // [32]
    float* skd;      // [32]
    static __host__ __device__ size_t bytes(int k)
    {
        return (size_t)CAND * 16 + (size_t)NSET * PCP_FD_NB * 4 + (size_t)NSET * CAP * 6 + 256 +
               (DEFER ? 0 : (size_t)32 * k * 4 + 256);
    }
    __device__ pcp_fd_smem(unsigned char* base, int k)
    {
        cand = reinterpret_cast<float4*>(base);
        hist = reinterpret_cast<uint32_t*>(cand + CAND);
        ld2 = reinterpret_cast<float*>(hist + NSET * PCP_FD_NB);
        lpos = reinterpret_cast<uint16_t*>(ld2 + NSET * CAP);
        tab = reinterpret_cast<uint32_t*>(lpos + NSET * CAP);
        nbr = tab + 64;
        sqpos = nbr + 32 * k;
        skd = reinterpret_cast<float*>(sqpos + 32);
    }
    // sorted position of the f-th point of the staged runs
    __device__ __forceinline__ uint32_t position(uint32_t f) const
    {
        int t = 0;
#pragma unroll
        for (int step = 16; step; step >>= 1)
            if (tab[t + step] <= f) t += step;
        return tab[32 + t] + (f - tab[t]);
    }
};

// far-list entries without a bound carry the level the far kernel should start at, as -(1 + level)
#define PCP_FAR_HINT(level) (-1.0f - (float)(level))
#define PCP_FD_OK 0
#define PCP_FD_BEYOND 1 // fewer than kk candidates inside `top`, or list overflow: no bound known
#define PCP_FD_AMBIG 2  // kk nearest found among the staged candidates but order / certification not provable in FP32
#define PCP_FD_IDLE 4
#define PCP_FD_RERANK 5 // as AMBIG for the order alone: handed to knn_rerank_kernel with its list
#define PCP_RR_HEAD 8   // words before the positions of a re-rank entry: qpos, m, edge, amb2, g2, delta

// acc + (x < 0): the sign bit as the increment (PTX, so that the compiler does not turn it back into a compare, a
// select and two moves)
__device__ __forceinline__ uint32_t pcp_add_sign(uint32_t acc, int x)
{
    uint32_t r;
    asm("{\n\t.reg .u32 t;\n\tshr.u32 t, %1, 31;\n\tadd.u32 %0, %2, t;\n\t}" : "=r"(r) : "r"(x), "r"(acc));
    return r;
}

// Bitonic sort, ascending, of G x SL (key, payload) pairs held SL per lane by the G lanes of a group: element
// i = gl SL + s.  Compare-exchange distance j < SL: inside the lane; else with lane gl ^ (j / SL), same slot.
template <int G, int SL>
__device__ __forceinline__ void pcp_fd_bitonic(int (&key)[SL], uint32_t (&pay)[SL], int gl)
{
    constexpr int N = G * SL;
#pragma unroll
    for (int k = 2; k <= N; k <<= 1) {
        // direction of the run of k elements that holds element i: ascending iff (i & k) == 0
        const bool up_lane = k >= N || ((gl * SL) & k) == 0; // (k >= SL: the same for every slot of the lane)
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j < SL) {
#pragma unroll
                for (int a = 0; a < SL; a++) {
                    const int b = a ^ j;
                    if (b > a) {
                        const bool up = k < SL ? ((a & k) == 0) : up_lane;
                        const bool sw = (key[a] > key[b]) == up;
                        const int ka = key[a], kb = key[b];
                        const uint32_t pa = pay[a], pb = pay[b];
                        key[a] = sw ? kb : ka, key[b] = sw ? ka : kb;
                        pay[a] = sw ? pb : pa, pay[b] = sw ? pa : pb;
                    }
                }
            } else {
                const int lj = j / SL;
                const bool keepmin = ((gl & lj) == 0) == up_lane;
#pragma unroll
                for (int a = 0; a < SL; a++) {
                    const int ok = __shfl_xor_sync(0xffffffffu, key[a], lj);
                    const uint32_t op = __shfl_xor_sync(0xffffffffu, pay[a], lj);
                    const int nk = keepmin ? min(key[a], ok) : max(key[a], ok); // (selects: no branch around the shuffles)
                    pay[a] = nk != key[a] ? op : pay[a];
                    key[a] = nk;
                }
            }
        }
    }
}

// Every group of G lanes searches the staged block for its own query.  dku: upper bound of the k-th squared
// distance (valid for OK and AMBIG).
// The kk nearest go to nbr_base[nbr_off + rank * nbr_stride] (a column of the parked lists, or the query's row in
// global memory: a 32-bit offset from the uniform base, not a pointer of its own to keep in registers).
template <int G, int CAP, int CAND, bool DEFER>
__device__ __forceinline__ int pcp_fd_search(int T, const pcp_fd_smem<CAP, G, CAND, DEFER>& S, int lane, bool active,
                                             uint32_t* nbr_base, uint32_t nbr_off, int nbr_stride, int kk, float qfx,
                                             float qfy, float qfz,
                                             float top, float g2, float e1, float& dku, const pcp_rr& rr, uint32_t qpos)
{
    constexpr int SL = CAP / G;
    constexpr int NBL = PCP_FD_NB / G;
    static_assert(NBL == 8 || NBL == 4, "pcp_fd_search: 8 or 16 lanes per query");
    const int gl = lane & (G - 1), gbase = lane & ~(G - 1), set = (G == 32) ? 0 : lane / G;
    const unsigned gmask = (G == 32) ? 0xffffffffu : ((1u << G) - 1u);
    uint32_t* hist = S.hist + set * PCP_FD_NB;
    float* ld2 = S.ld2 + set * CAP;
    uint16_t* lpos = S.lpos + set * CAP;
    const int key0 = (__float_as_int(top) >> 21) - (PCP_FD_NB - 2);
#pragma unroll
    for (int i = 0; i < NBL; i++) hist[gl + i * G] = 0;
    __syncwarp();
    // ---- pass 1: histogram of FP32 squared distances ----
    // The staged block is padded to a multiple of G with points at "infinity" (beyond `top` for every query), so
    // the loops carry no bounds test; the increment is one predicated instruction, not a divergent region (some
    // lane of the warp takes it at almost every step, so a branch around it saves nothing and costs three).
    // Two candidates per lane and step, their distances on the packed FP32 pipe (half the FP instructions of an
    // issue-bound loop).  (c - q)^2 here, (q - c)^2 in the reference order: the same value bit for bit.
    const int Tpad = (T + 2 * G - 1) & ~(2 * G - 1);
    const float4* cpa = S.cand + gl;
    const float4* cpb = S.cand + CAND / 2 + gl;
    const float2 nqx = make_float2(-qfx, -qfx), nqy = make_float2(-qfy, -qfy), nqz = make_float2(-qfz, -qfz);
    {
        const uint32_t hist_s = (uint32_t)__cvta_generic_to_shared(hist);
        const uint32_t dummy_s = (uint32_t)__cvta_generic_to_shared(S.tab + lane); // free once the block is staged
        const int key1 = active ? key0 : -0x40000000; // an idle group counts nothing: every bucket number "beyond"
        for (int p0 = 0; p0 < Tpad / 2; p0 += G) {
            const float4 a = cpa[p0], bb = cpb[p0];
            const float2 dx = __fadd2_rn(make_float2(a.x, a.y), nqx), dy = __fadd2_rn(make_float2(a.z, a.w), nqy),
                         dz = __fadd2_rn(make_float2(bb.x, bb.y), nqz);
            const float2 d2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
            // (the last bucket = beyond `top` is never counted: nothing is decided there, and it is the one address
            // most lanes would collide on; lanes with nothing to count hit a word of their own)
            const int b0 = (__float_as_int(d2.x) >> 21) - key1, b1 = (__float_as_int(d2.y) >> 21) - key1;
            const bool t0 = b0 < PCP_FD_NB - 1, t1 = b1 < PCP_FD_NB - 1;
            asm volatile("red.shared.add.u32 [%0], 1;" : : "r"(t0 ? hist_s + 4u * (uint32_t)max(b0, 0) : dummy_s) : "memory");
            asm volatile("red.shared.add.u32 [%0], 1;" : : "r"(t1 ? hist_s + 4u * (uint32_t)max(b1, 0) : dummy_s) : "memory");
        }
    }
    __syncwarp();
    // ---- scan: bucket b* of the kk-th candidate; hist[b] becomes the write cursor (exclusive prefix) ----
    uint32_t hb[NBL];
    {
        const uint4 a = *reinterpret_cast<const uint4*>(hist + gl * NBL);
        hb[0] = a.x, hb[1] = a.y, hb[2] = a.z, hb[3] = a.w;
        if constexpr (NBL == 8) {
            const uint4 b = *reinterpret_cast<const uint4*>(hist + gl * 8 + 4);
            hb[4] = b.x, hb[5] = b.y, hb[6] = b.z, hb[7] = b.w;
        }
    }
    uint32_t tot = 0;
#pragma unroll
    for (int i = 0; i < NBL; i++) tot += hb[i];
    uint32_t incl = tot;
#pragma unroll
    for (int o = 1; o < G; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o, G);
        if (gl >= o) incl += v;
    }
    int st = active ? PCP_FD_OK : PCP_FD_IDLE;
    const unsigned bal = (__ballot_sync(0xffffffffu, incl >= (uint32_t)kk) >> gbase) & gmask;
    if (!bal && active) st = PCP_FD_BEYOND;
    const int L = bal ? __ffs(bal) - 1 : 0;
    int myb = PCP_FD_NB;
    uint32_t mym = 0;
    {
        uint32_t c = incl - tot;
        uint32_t ex[NBL];
#pragma unroll
        for (int i = 0; i < NBL; i++) {
            ex[i] = c;
            c += hb[i];
            if (myb == PCP_FD_NB && c >= (uint32_t)kk) myb = gl * NBL + i, mym = c;
        }
        uint4 a;
        a.x = ex[0], a.y = ex[1], a.z = ex[2], a.w = ex[3];
        *reinterpret_cast<uint4*>(hist + gl * NBL) = a;
        if constexpr (NBL == 8) {
            uint4 b;
            b.x = ex[4], b.y = ex[5], b.z = ex[6], b.w = ex[7];
            *reinterpret_cast<uint4*>(hist + gl * 8 + 4) = b;
        }
    }
    const int bstar = __shfl_sync(0xffffffffu, myb, gbase + L);
    const uint32_t m = __shfl_sync(0xffffffffu, mym, gbase + L); // candidates in buckets <= b*
    if (st == PCP_FD_OK && (bstar >= PCP_FD_NB - 1 || m > (uint32_t)CAP)) st = PCP_FD_BEYOND;
    const bool live = st == PCP_FD_OK;
    __syncwarp();
    // ---- pass 2: scatter the candidates of buckets <= b* into the list, bucket by bucket ----
    const int lim = (key0 + min(bstar, PCP_FD_NB - 2) + 1) << 21;
    for (int p0 = 0; p0 < Tpad / 2; p0 += G) {
        const float4 a = cpa[p0], bb = cpb[p0];
        const float2 dx = __fadd2_rn(make_float2(a.x, a.y), nqx), dy = __fadd2_rn(make_float2(a.z, a.w), nqy),
                     dz = __fadd2_rn(make_float2(bb.x, bb.y), nqz);
        const float2 d2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
        // bucket(d2) <= b*  <=>  the bits of d2 lie below the upper edge of b* (one compare for the ~90 % left out;
        // the padding lies beyond every edge)
        if (live && min(__float_as_int(d2.x), __float_as_int(d2.y)) < lim) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const float d = h ? d2.y : d2.x;
                if (__float_as_int(d) < lim) {
                    const int b = pcp_fd_bucket(d, key0);
                    const uint32_t at = atomicAdd(&hist[b], 1u);
                    ld2[at] = d;
                    lpos[at] = (uint16_t)(4 * (p0 + gl) + 2 + h);
                }
            }
        }
    }
    __syncwarp();
    // ---- rank inside the bucket (hist[b] is now the END of bucket b) ----
    const float edge = pcp_fk_edge(min(bstar, PCP_FD_NB - 2), key0);
    const float delta = 4.0f * sqrtf(edge) * e1 + 4.0f * e1 * e1 + 1e-6f * edge;
    const float amb2 = 2.0f * delta;
    bool ambig = false;
    float dkl = -1.0f;
    if constexpr (CAP <= 64) {
        // The list, G lanes x SL entries (lane gl: entries gl SL .. gl SL + SL - 1, empty ones at +inf), sorted by a
        // bitonic network on the bit patterns of d2 (non-negative floats order like their bits) with the candidate
        // as payload: strides below SL inside the lane, the others by shuffle.  After it the entry's place IS its
        // rank, and the order is the exact one iff every neighbouring pair up to (kk-1, kk) lies farther apart than
        // the error bound: then no pair at all up to rank kk does, and everything behind rank kk lies behind the
        // k-th.  (Before: every entry counted the smaller and the near ones of its bucket -- twice the instructions,
        // and pairs behind rank kk counted as ambiguous.)
        int key[SL];
        uint32_t pay[SL];
        {
            const int4* kp = reinterpret_cast<const int4*>(ld2 + gl * SL);
            const uint2* pp = reinterpret_cast<const uint2*>(lpos + gl * SL);
#pragma unroll
            for (int v = 0; v < SL / 4; v++) {
                const int4 kv = kp[v];
                const uint2 pv = pp[v];
                key[4 * v + 0] = kv.x, key[4 * v + 1] = kv.y, key[4 * v + 2] = kv.z, key[4 * v + 3] = kv.w;
                pay[4 * v + 0] = pv.x & 0xffffu, pay[4 * v + 1] = pv.x >> 16, pay[4 * v + 2] = pv.y & 0xffffu, pay[4 * v + 3] = pv.y >> 16;
            }
#pragma unroll
            for (int t = 0; t < SL; t++)
                if (!live || gl * SL + t >= (int)m) key[t] = 0x7f800000, pay[t] = 0;
        }
        pcp_fd_bitonic<G, SL>(key, pay, gl);
        // neighbouring pairs: the next entry is the next slot, or the next lane's first
        const int nextlane = __shfl_down_sync(0xffffffffu, key[0], 1, G);
#pragma unroll
        for (int t = 0; t < SL; t++) {
            const int r = gl * SL + t;
            const int nk = t + 1 < SL ? key[t + 1] : (gl == G - 1 ? 0x7f800000 : nextlane);
            if (r < kk) {
                if (__int_as_float(nk) - __int_as_float(key[t]) <= amb2) ambig = true;
                if (live) nbr_base[nbr_off + (uint32_t)r * (uint32_t)nbr_stride] = S.pos_at(pay[t]);
                if (r == kk - 1) dkl = __int_as_float(key[t]);
            }
        }
        if (!live) dkl = -1.0f, ambig = false;
    } else {
#pragma unroll 1 // code size: the body holds a loop of its own, and the kernel has to stay inside the instruction cache
        for (int s = 0; s < SL; s++) {
            const int e = s * G + gl;
            if (live && e < (int)m) {
                const float d = ld2[e];
                const int b = pcp_fd_bucket(d, key0);
                const uint32_t lo = b ? hist[b - 1] : 0u, hi = hist[b];
                // bucket edges: an entry this close to one may belong to the other side
                if (b > 0 && d - pcp_fk_edge(b - 1, key0) <= amb2 && lo > 0) ambig = true;
                if (b < bstar && pcp_fk_edge(b, key0) - d <= amb2) ambig = true;
                // rank = entries of the earlier buckets + smaller entries of this one; `near` = entries within the error
                // bound of d, the entry itself included.  Counted on the bit patterns (non-negative floats order like
                // their bits; a difference of two such patterns, or of |x|'s and amb2's, cannot overflow): the sign bit
                // of the difference is the increment, two instructions per count instead of compare + select + move.
                const int dbits = __float_as_int(d), abits = __float_as_int(amb2) + 1;
                uint32_t less = 0, near = 0;
                const float* pj = ld2 + lo;
#pragma unroll 1 // a bucket holds a handful of entries: unrolled by 8 the remainder code is what runs
                for (const float* const pe = ld2 + hi; pj != pe; pj++) {
                    const float dj = *pj;
                    less = pcp_add_sign(less, __float_as_int(dj) - dbits);            // dj < d
                    near = pcp_add_sign(near, __float_as_int(fabsf(dj - d)) - abits); // |dj - d| <= amb2
                }
                const uint32_t r = lo + less;
                if (near > 1u) ambig = true;
                if (r < (uint32_t)kk) nbr_base[nbr_off + r * (uint32_t)nbr_stride] = S.pos_at(lpos[e]);
                if (r == (uint32_t)kk - 1u) dkl = d;
            }
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) dkl = fmaxf(dkl, __shfl_xor_sync(0xffffffffu, dkl, o));
    ambig = ((__ballot_sync(0xffffffffu, ambig) >> gbase) & gmask) != 0u;
    // FP32 could not order the list (two entries closer than the error bound, an entry at a bucket edge): the list
    // -- EVERY staged candidate below the upper edge of bucket b* -- goes to the re-rank list with the thresholds the
    // exact k-th distance has to clear, and knn_rerank_kernel orders it on exact distances.  (Done here, inside the
    // search, the doubles and the call cost the 64-register kernel 0.9 ms to save the far kernel 1.1.)
    if constexpr (DEFER && CAP <= 64) {
        if (live && ambig && rr.list) { // uniform over the group
            uint32_t slot = 0;
            if (gl == 0) slot = atomicAdd(rr.count, 1u);
            slot = __shfl_sync(gmask << gbase, slot, gbase);
            if (slot < rr.cap) {
                uint32_t* const ent = rr.list + (size_t)slot * (CAP + PCP_RR_HEAD);
                if (gl == 0) {
                    ent[0] = qpos, ent[1] = m, ent[2] = __float_as_uint(edge), ent[3] = __float_as_uint(amb2);
                    ent[4] = __float_as_uint(g2), ent[5] = __float_as_uint(delta);
                }
#pragma unroll
                for (int s = 0; s < SL; s++)
                    if (s * G + gl < (int)m) ent[PCP_RR_HEAD + s * G + gl] = S.pos_at(lpos[s * G + gl]);
                st = PCP_FD_RERANK;
            }
        }
    }
    // the k-th against bucket b*+1 and against the certified radius
    dku = ((dkl >= 0.0f && !ambig) ? dkl : edge) + delta;
    if (live && st != PCP_FD_RERANK) {
        if (ambig || dkl < 0.0f || edge - dkl <= amb2 || !(dkl + amb2 < g2)) st = PCP_FD_AMBIG;
    }
    __syncwarp();
    return st;
}

// resident blocks per SM the kernel is compiled for = what its shared memory allows (deferred rows: 26 KB per block
// at 320 candidates -> 8 blocks, 64 registers; 14.1 ms against 15.7 at 6 blocks on the 25 M-point shard)
__host__ __device__ constexpr int pcp_fd_min_blocks(int cand, bool defer)
{
    return !defer ? (cand > 480 ? 3 : PCP_FK_MINB) : cand <= 320 ? 8 : cand <= 480 ? 5 : cand <= 800 ? 4 : 3;
}

// CAP: list capacity per query (>= k + room for the boundary bucket); GS: lanes per query.
// Settles every query whose k nearest provably lie inside the 3x3x3 block of base cells around its cell; the
// others are appended to the "far" list with the best known upper bound of their k-th squared distance
// (+inf if none) for knn_far_kernel.
template <int CAP, int GS, int CAND, bool FLAT, bool DEFER = false>
static __global__ void __launch_bounds__(32 * PCP_FK_WARPS, pcp_fd_min_blocks(CAND, DEFER)) knn_fast_kernel(pcp_fknn_params P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int NSET = 32 / GS;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    using Smem = pcp_fd_smem<CAP, GS, CAND, DEFER>;
    const Smem S(smem_raw + Smem::bytes(P.k) * warp, P.k);
    const int set = lane / GS;
    if constexpr (!DEFER) S.sqpos[lane] = 0xffffffffu;
    __syncwarp();
    int fill = 0;
    const uint32_t nitem = P.nitem_ptr ? *P.nitem_ptr : P.nitem;
    while (true) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(P.work_counter, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= nitem) break;
        if (P.item_list) item = __ldg(P.item_list + item);
        const uint32_t cell = __ldg(P.item_cell + item);
        const uint32_t first0 = __ldg(P.item_first + item);
        const uint32_t meta = __ldg(P.item_meta + item);
        // the item is an octree leaf: a cell of edge h0 = h * 2^level holding the nq queries first..first+nq-1
        const int level0 = (int)(meta >> 8);
        const int nq0 = (int)(meta & 0xffu);
        // Sub-items (the second pass only): a leaf whose 3x3x3 block holds more points than even this pass stages --
        // a sparse leaf beside a dense surface -- is cut into its child cells (one level down: an eighth of the block's
        // volume), each searched as an item of its own.  Its queries' neighbours are mostly the surface's points, close
        // by, and the far kernel, which would take them otherwise, spends 72 k cycles on each.
        uint32_t first = first0;
        int nq = nq0, level = level0;
        int ccx, ccy, ccz;
        pcp_demorton(__ldg(P.ckey + cell) >> (3 * level0), ccx, ccy, ccz);
        int scx = 0, scy = 0, scz = 0; // lane j: the child cell of the leaf's query j
        int split_lo = -1;             // -1: the leaf itself; else the child that starts at the leaf's query split_lo
        bool split_start = false;
        auto next_sub = [&]() -> bool {
            if constexpr (CAND <= 480) return false;
            if (split_start) split_start = false, split_lo = 0;
            else if (split_lo < 0) return false;
            else split_lo += nq;
            if (split_lo >= nq0) return false;
            ccx = __shfl_sync(0xffffffffu, scx, split_lo), ccy = __shfl_sync(0xffffffffu, scy, split_lo),
            ccz = __shfl_sync(0xffffffffu, scz, split_lo);
            // the points of a leaf are in Morton order: a child's queries are consecutive
            const unsigned grp = __ballot_sync(0xffffffffu, lane >= split_lo && lane < nq0 && scx == ccx && scy == ccy && scz == ccz);
            const unsigned run = ~(grp >> split_lo);
            nq = run ? __ffs(run) - 1 : 32;
            first = first0 + (uint32_t)split_lo;
            level = level0 - 1;
            return true;
        };
        for (bool more = true; more; more = next_sub()) {
        const double h0 = P.g.h * (double)(1 << level);
        // |float - true| of a block-relative coordinate difference: operands up to 2h and 3h in magnitude, each
        // rounded once (2^-24 relative), plus the rounding of the double subtraction p - origin
        const float e1 = (float)(8.0 * 5.97e-8 * h0 + 2.0 * P.epsd) * 1.0001f;
        const float top0 = (float)(2.25 * h0 * h0); // nothing farther than 1.5 h can be certified in this block
        // lane j holds query j of the item (halo points are candidates only)
        const pcp_rec myq = pcp_load_rec(P.rec, first + (lane < nq ? lane : 0));
        const unsigned want = __ballot_sync(0xffffffffu, lane < nq && myq.idx < P.query_limit);
        if (!want) { // no query in the leaf (halo points only): their rows stay "not written"
            if constexpr (DEFER) {
                if (lane < nq) P.nbr_out[(first + (uint32_t)lane) * (uint32_t)P.k] = 0xffffffffu;
            }
            continue;
        }
        const int nwant = __popc(want);
        if constexpr (!DEFER) {
            if (fill + nwant > 32) {
                pcp_fk_flush<Smem, false>(P, S, lane);
                fill = 0;
            }
        }
        const int cx = ccx, cy = ccy, cz = ccz;
        const int dimx = ((P.g.dims[0] - 1) >> level) + 1, dimy = ((P.g.dims[1] - 1) >> level) + 1,
                  dimz = ((P.g.dims[2] - 1) >> level) + 1;
        const double ox = P.g.org[0] + (double)cx * h0, oy = P.g.org[1] + (double)cy * h0, oz = P.g.org[2] + (double)cz * h0;
        const float myfx = (float)(myq.x - ox), myfy = (float)(myq.y - oy), myfz = (float)(myq.z - oz);

        // ---- the 27 neighbour cells (hash) and the staged candidate block ----
        uint32_t nb_start = 0, nb_count = 0;
        if (lane < 27) {
            int dx, dy, dz;
            pcp_nb_offset(lane, dx, dy, dz);
            const int nx = cx + dx, ny = cy + dy, nz = cz + dz;
            if (nx >= 0 && ny >= 0 && nz >= 0 && nx < dimx && ny < dimy && nz < dimz) {
                if (level == 0) {
                    if (!pcp_hash_find(P.htab, P.hmask, pcp_morton(nx, ny, nz), nb_start, nb_count)) nb_count = 0;
                } else {
                    uint32_t a = 0, cn = 0;
                    if (pcp_hash_find(P.ltab, P.lmask, PCP_LKEY(level, pcp_morton(nx, ny, nz)), a, cn)) {
                        nb_start = __ldg(P.cstart + a);
                        nb_count = __ldg(P.cstart + a + cn) - nb_start;
                    }
                }
            }
        }
        uint32_t pre = nb_count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, pre, o);
            if (lane >= o) pre += v;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
        pre -= nb_count;
        if (P.dense_list && total > (uint32_t)CAND) { // left to the second pass
            if (lane == 0) P.dense_list[atomicAdd(P.dense_count, 1u)] = item;
            continue;
        }
        if constexpr (CAND > 480) {
            if (!P.dense_list && total > (uint32_t)CAND && level > 0 && split_lo < 0) {
                const int sh = level - 1;
                scx = min(max(pcp_cell_coord(myq.x, P.g.org[0], P.g.inv_h, 0), 0), P.g.dims[0] - 1) >> sh;
                scy = min(max(pcp_cell_coord(myq.y, P.g.org[1], P.g.inv_h, 0), 0), P.g.dims[1] - 1) >> sh;
                scz = min(max(pcp_cell_coord(myq.z, P.g.org[2], P.g.inv_h, 0), 0), P.g.dims[2] - 1) >> sh;
                split_start = true;
                continue; // next_sub() sets up the first child
            }
        }
        const bool staged = total <= (uint32_t)CAND && total >= (uint32_t)P.kk;
        PCP_FK_DBG(0, 1);
        PCP_FK_DBG(1, staged ? 1 : 0);
        PCP_FK_DBG(2, nwant);
        PCP_FK_DBG(3, total);
        PCP_FK_DBG(4, total > (uint32_t)CAND);
        PCP_FK_DBG(8, total > (uint32_t)CAND ? nwant : 0);
        PCP_FK_DBG(9, total < (uint32_t)P.kk ? nwant : 0);
        unsigned okmask = 0, rrmask = 0; // settled; left to knn_rerank_kernel
        // lane j: upper bound of query j's k-th squared distance (+inf: none; -1: none, block too dense to stage)
        // (PCP_FAR_HINT: no bound, but a level to start from -- the block of this leaf's level holds too many points
        // to stage, so the level below it is the first worth trying; or too few, so the level above it is)
        float myU = (total > (uint32_t)CAND)  ? PCP_FAR_HINT(level > 0 ? level - 1 : 0)
                    : (total < (uint32_t)P.kk) ? PCP_FAR_HINT(level + 1)
                                               : __int_as_float(0x7f800000);
        if (staged) {
            if constexpr (FLAT) {
                // the runs' table, then the block itself, flat over the candidates with PCP_FK_STAGE_U loads in
                // flight per lane
                S.tab[lane] = lane < 27 ? pre : 0xffffffffu;
                S.tab[32 + lane] = nb_start;
                __syncwarp();
#pragma unroll 1
                for (uint32_t f0 = 0; f0 < total; f0 += 32 * PCP_FK_STAGE_U) {
                    pcp_rec r[PCP_FK_STAGE_U];
                    uint32_t ps[PCP_FK_STAGE_U];
#pragma unroll
                    for (int u = 0; u < PCP_FK_STAGE_U; u++) {
                        const uint32_t f = f0 + 32 * u + lane;
                        ps[u] = 0;
                        if (f < total) {
                            ps[u] = S.position(f);
                            r[u] = pcp_load_rec(P.rec, ps[u]);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < PCP_FK_STAGE_U; u++) {
                        const uint32_t f = f0 + 32 * u + lane;
                        if (f < total)
                            S.put(f, (float)(r[u].x - ox), (float)(r[u].y - oy), (float)(r[u].z - oz), ps[u]);
                    }
                }
            } else {
                for (int nb = 0; nb < 27; nb++) {
                    const uint32_t c = __shfl_sync(0xffffffffu, nb_count, nb);
                    if (c == 0) continue;
                    const uint32_t s = __shfl_sync(0xffffffffu, nb_start, nb);
                    const uint32_t p = __shfl_sync(0xffffffffu, pre, nb);
                    for (uint32_t t = lane; t < c; t += 32) {
                        const pcp_rec r = pcp_load_rec(P.rec, s + t);
                        S.put(p + t, (float)(r.x - ox), (float)(r.y - oy), (float)(r.z - oz), s + t);
                    }
                }
            }
            // padding up to a multiple of 2 GS: points no query can reach (pcp_fd_search)
            if (lane < (int)(((total + 2 * GS - 1) & ~(uint32_t)(2 * GS - 1)) - total))
                S.put(total + lane, 1e18f, 1e18f, 1e18f, 0u);
            __syncwarp();
            for (unsigned rest = want; rest;) {
                // group g takes the g-th remaining query
                int jj = -1;
#pragma unroll
                for (int g = 0; g < NSET; g++) {
                    const int b = rest ? __ffs(rest) - 1 : -1;
                    if (g == set) jj = b;
                    rest &= rest - 1; // 0 stays 0
                }
                const bool active = jj >= 0;
                const int sj = active ? jj : 0;
                const float qfx = __shfl_sync(0xffffffffu, myfx, sj), qfy = __shfl_sync(0xffffffffu, myfy, sj),
                            qfz = __shfl_sync(0xffffffffu, myfz, sj);
                // certified radius: distance from the query to the faces of the 3x3x3 block (faces on the grid
                // boundary do not count), in cell-relative FP32 with a safety margin
                float dout = 3.0e38f;
                const float hf = (float)h0;
                if (cx > 0) dout = fminf(dout, qfx + hf);
                if (cx + 1 < dimx) dout = fminf(dout, 2.0f * hf - qfx);
                if (cy > 0) dout = fminf(dout, qfy + hf);
                if (cy + 1 < dimy) dout = fminf(dout, 2.0f * hf - qfy);
                if (cz > 0) dout = fminf(dout, qfz + hf);
                if (cz + 1 < dimz) dout = fminf(dout, 2.0f * hf - qfz);
                dout = fmaxf(dout - (4.0f * e1 + 1e-6f * hf), 0.0f);
                const float g2 = dout * dout * 0.999999f;
                // DEFER: the neighbour row of the query in global memory; a row whose query is not settled here is
                // marked unwritten again (the far kernel writes it whole)
                uint32_t* nbr_base;
                uint32_t nbr_off;
                int nbr_stride;
                if constexpr (DEFER) {
                    nbr_base = P.nbr_out, nbr_off = (first + (uint32_t)sj) * (uint32_t)P.k, nbr_stride = 1; // < 2^32: host
                } else {
                    nbr_base = S.nbr, nbr_off = (uint32_t)(fill + __popc(want & ((1u << sj) - 1u))), nbr_stride = 32;
                }
                float dku;
                const int st = pcp_fd_search<GS, CAP, CAND, DEFER>((int)total, S, lane, active, nbr_base, nbr_off, nbr_stride, P.kk,
                                                                   qfx, qfy, qfz, top0, g2, e1, dku, P.rr, first + (uint32_t)sj);
                // collect the group results as query bits
#pragma unroll
                for (int g = 0; g < NSET; g++) {
                    const int jg = __shfl_sync(0xffffffffu, jj, g * GS);
                    const int sg = __shfl_sync(0xffffffffu, st, g * GS);
                    const float dg = __shfl_sync(0xffffffffu, dku, g * GS);
                    if (jg >= 0) {
                        PCP_FK_DBG(6, sg == PCP_FD_BEYOND);
                        PCP_FK_DBG(7, sg == PCP_FD_AMBIG);
                        if (sg == PCP_FD_OK) okmask |= 1u << jg;
                        if (sg == PCP_FD_RERANK) rrmask |= 1u << jg;
                        if ((sg == PCP_FD_AMBIG || sg == PCP_FD_OK) && lane == jg) myU = dg * 1.000001f;
                    }
                }
            }
            PCP_FK_DBG(5, __popc(okmask));
        }
        // ---- park the finished queries; the rest go to the far list ----
        if constexpr (DEFER) {
            // Every row of the leaf that holds no answer is marked "not written" here -- open queries (the search
            // wrote part of a list into them), the ones left to the re-rank kernel, halo points -- so that no pass
            // over all rows is needed before the search (0.37 ms per 25 M rows).
            const unsigned rows = nq >= 32 ? 0xffffffffu : ((1u << nq) - 1u);
            if (((rows & ~okmask) >> lane) & 1u) P.nbr_out[(first + (uint32_t)lane) * (uint32_t)P.k] = 0xffffffffu;
        }
        if constexpr (!DEFER) {
            if ((okmask >> lane) & 1u) {
                const int sl = fill + __popc(want & ((1u << lane) - 1u));
                S.sqpos[sl] = first + lane;
                S.skd[sl] = myU;
            }
            fill += nwant;
        }
        {
            const unsigned fm = want & ~okmask & ~rrmask;
            if (fm) {
                uint32_t basei = 0;
                // (a record for the third pass if one of the open queries has a bound: see knn_wide.cuh)
                const unsigned bounded = __ballot_sync(0xffffffffu, myU >= 0.f && myU < 3.0e38f) & fm;
                if (lane == 0) {
                    basei = atomicAdd(P.fail_count, (uint32_t)__popc(fm));
                    if (P.wide_list && bounded)
                        P.wide_list[atomicAdd(P.wide_count, 1u)] = make_uint2(basei, (uint32_t)__popc(fm) | ((uint32_t)level << 8));
                }
                basei = __shfl_sync(0xffffffffu, basei, 0);
                if ((fm >> lane) & 1u) {
                    const uint32_t at = basei + __popc(fm & ((1u << lane) - 1u));
                    P.fail_list[at] = first + lane;
                    P.fail_u[at] = myU;
                }
            }
        }
        } // sub-items
    }
    if constexpr (!DEFER) pcp_fk_flush<Smem, false>(P, S, lane);
}


// ---- exact re-rank of the lists the FP32 search could not order ------------------------------------------------
// One warp per entry: squared distances in double from the records, rank by (d2, original index) ascending -- the far
// kernel's rule, the order of the reference's result set -- and the row is written.  The exact k-th distance then has
// to clear the same thresholds as in pcp_fd_search (the upper edge of bucket b*, beyond which candidates were left
// out of the list, and the certified radius); a query that does not joins the far list with its bound.
template <int CAP>
static __global__ void __launch_bounds__(128) knn_rerank_kernel(pcp_fknn_params P)
{
    constexpr int SL = CAP / 32;
    const int lane = threadIdx.x & 31;
    const uint32_t n = min(*P.rr.count, P.rr.cap);
    for (uint32_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < n; w += gridDim.x * (blockDim.x >> 5)) {
        const uint32_t* ent = P.rr.list + (size_t)w * (CAP + PCP_RR_HEAD);
        const uint32_t qpos = ent[0], m = ent[1];
        const float edge = __uint_as_float(ent[2]), amb2 = __uint_as_float(ent[3]), g2 = __uint_as_float(ent[4]),
                    delta = __uint_as_float(ent[5]);
        const pcp_rec q = pcp_load_rec(P.rec, qpos);
        double d2s[SL];
        int ids[SL], rank[SL];
        uint32_t ps[SL];
#pragma unroll
        for (int s = 0; s < SL; s++) {
            const uint32_t e = s * 32 + lane;
            d2s[s] = DBL_MAX, ids[s] = 0x7fffffff, ps[s] = 0, rank[s] = 0;
            if (e < m) {
                ps[s] = ent[PCP_RR_HEAD + e];
                const pcp_rec r = pcp_load_rec(P.rec, ps[s]);
                d2s[s] = pcp_dist2(q.x, q.y, q.z, r.x, r.y, r.z);
                ids[s] = (int)r.idx;
            }
        }
#pragma unroll
        for (int s2 = 0; s2 < SL; s2++) {
            const int mt = min(32, (int)m - 32 * s2);
#pragma unroll 1
            for (int t = 0; t < mt; t++) {
                const double dj = __shfl_sync(0xffffffffu, d2s[s2], t);
                const int ij = __shfl_sync(0xffffffffu, ids[s2], t);
#pragma unroll
                for (int s = 0; s < SL; s++) rank[s] += (dj < d2s[s] || (dj == d2s[s] && ij < ids[s])) ? 1 : 0;
            }
        }
        double dk = -1.0;
        uint32_t* row = P.nbr_out + (size_t)qpos * P.k;
#pragma unroll
        for (int s = 0; s < SL; s++)
            if ((uint32_t)(s * 32 + lane) < m && rank[s] < P.kk) {
                row[rank[s]] = ps[s];
                if (rank[s] == P.kk - 1) dk = d2s[s];
            }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dk = fmax(dk, __shfl_xor_sync(0xffffffffu, dk, o));
        const float dkl = dk < 0.0 ? -1.0f : __double2float_ru(dk);
        const bool ok = dkl >= 0.0f && !(edge - dkl <= amb2) && dkl + amb2 < g2;
        __syncwarp();
        if (!ok && lane == 0) {
            row[0] = 0xffffffffu; // not settled: the far kernel writes the row whole
            const uint32_t at = atomicAdd(P.fail_count, 1u);
            P.fail_list[at] = qpos;
            P.fail_u[at] = ((dkl >= 0.0f ? dkl : edge) + delta) * 1.000001f;
        }
    }
}
