// knn_wide.cuh -- the third dense pass: the leaves whose queries the 3x3x3 block found but could not certify.
//
// More than half of the queries the dense passes leave open have their k nearest staged already: only the k-th
// distance reaches a little beyond the distance to the nearest face of the block (88 % of them by less than a quarter
// of it), so all they bring to the far list is a bound.  The far kernel takes such a query alone -- one warp, its own
// cell enumeration and corner distances.  Here the leaf takes its bounded queries together (1.5 of them on average):
// the box around their bound spheres, the cells of the leaf's own level that the box touches (at most 5 x 5 x 5, one
// hash lookup each), the points of those cells staged with everything outside the box left out (67 of 336 on the
// 25 M-point city shard), and the search of the dense kernel (pcp_fd_search, 8 lanes per query) over that.  The
// certified radius of a query is its distance to the faces of the BOX: everything inside the box is staged, and the
// box holds the query's bound sphere, so 96 % of them settle.  A settled query is struck from the far list
// (0xffffffff); one that is not keeps its entry, with the tighter bound if one was found.  Queries without a bound
// (isolated points, blocks with fewer than k points) are not taken: their box would be the whole ring, which costs a
// leaf more to stage than the far kernel spends on them (measured: 6.6 ms against 2.0, for 0.9 ms less far kernel).
// Exactness is the dense kernel's: FP32 relative to the leaf's origin with the error bound scaled to the wider
// range, provable order or no answer (kd_tree.h:814-845 of the reference: exact k nearest, ascending).
#pragma once
#include "knn_fast.cuh"

#define PCP_WIDE_RUNS 128     // the cells of a 5 x 5 x 5 block, padded to a power of two
#define PCP_WIDE_MAXLOAD 4096 // points in the touched cells beyond which the leaf is left to the far kernel
#define PCP_WIDE_U 4          // loads in flight per lane while staging

template <int CAP, int GS, int CAND> struct pcp_wide_layout {
    using Smem = pcp_fd_smem<CAP, GS, CAND, true>;
    static __host__ __device__ constexpr size_t bytes() { return Smem::bytes(0) + (size_t)PCP_WIDE_RUNS * 8; }
};

#define PCP_WIDE_DBG(i, v)                                                                                   \
    do {                                                                                                     \
        if (P.dbg && lane == 0) atomicAdd(P.dbg + (i), (unsigned long long)(v));                             \
    } while (0)

template <int CAP, int GS, int CAND>
static __global__ void __launch_bounds__(32 * PCP_FK_WARPS, CAND <= 256 ? 6 : 3) knn_wide_kernel(pcp_fknn_params P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using L = pcp_wide_layout<CAP, GS, CAND>;
    using Smem = typename L::Smem;
    constexpr int NSET = 32 / GS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int set = lane / GS;
    unsigned char* const base = smem_raw + L::bytes() * warp;
    const Smem S(base, P.k);
    uint32_t* const wpre = reinterpret_cast<uint32_t*>(base + Smem::bytes(0)); // [128] exclusive prefix of the run lengths
    uint32_t* const wstart = wpre + PCP_WIDE_RUNS;                             // [128] first point of each run
    const uint32_t nrec = *P.wide_count;
    const float INF = __int_as_float(0x7f800000);
    while (true) {
        uint32_t w = 0;
        if (lane == 0) w = atomicAdd(P.wide_work, 1u);
        w = __shfl_sync(0xffffffffu, w, 0);
        if (w >= nrec) break;
        const uint2 wr = P.wide_list[w];
        const int cnt = (int)(wr.y & 0xffu), level = (int)(wr.y >> 8);
        const uint32_t qpos = P.fail_list[wr.x + (lane < cnt ? lane : 0)];
        const float U = lane < cnt ? P.fail_u[wr.x + lane] : INF;
        // the queries with a bound: the others (too few points near them) would ask for the whole 5x5x5 ring, which
        // costs a leaf more to stage than the far kernel spends on them
        const bool mine = lane < cnt && U >= 0.f && U < INF;
        const unsigned want = __ballot_sync(0xffffffffu, mine);
        PCP_WIDE_DBG(24, 1);
        PCP_WIDE_DBG(28, __popc(want));
        if (!want) {
            PCP_WIDE_DBG(25, 1);
            continue;
        }
        const pcp_rec myq = pcp_load_rec(P.rec, qpos);
        const double h0 = P.g.h * (double)(1 << level);
        const float hf = (float)h0;
        // the leaf's cell (sub-items of the second pass: the child cell), from its first open query
        int cx = min(max(pcp_cell_coord(myq.x, P.g.org[0], P.g.inv_h, 0), 0), P.g.dims[0] - 1) >> level;
        int cy = min(max(pcp_cell_coord(myq.y, P.g.org[1], P.g.inv_h, 0), 0), P.g.dims[1] - 1) >> level;
        int cz = min(max(pcp_cell_coord(myq.z, P.g.org[2], P.g.inv_h, 0), 0), P.g.dims[2] - 1) >> level;
        cx = __shfl_sync(0xffffffffu, cx, 0), cy = __shfl_sync(0xffffffffu, cy, 0), cz = __shfl_sync(0xffffffffu, cz, 0);
        // (every query of the record lies in that cell, bounded or not)
        const int dimx = ((P.g.dims[0] - 1) >> level) + 1, dimy = ((P.g.dims[1] - 1) >> level) + 1,
                  dimz = ((P.g.dims[2] - 1) >> level) + 1;
        const double ox = P.g.org[0] + (double)cx * h0, oy = P.g.org[1] + (double)cy * h0, oz = P.g.org[2] + (double)cz * h0;
        const float myfx = (float)(myq.x - ox), myfy = (float)(myq.y - oy), myfz = (float)(myq.z - oz);
        // coordinates between -2 h0 and 3 h0: twice the rounding of the 3x3x3 block's
        const float e1 = (float)(16.0 * 5.97e-8 * h0 + 2.0 * P.epsd) * 1.0001f;
        const float top0 = (float)(6.25 * h0 * h0); // nothing farther than 2.5 h can be certified in a 5x5x5 block
        // ---- the box of the bound spheres ----
        const float r = mine ? sqrtf(U) * 1.001f + 16.0f * e1 : 0.f;
        float lox = mine ? myfx - r : 3e38f, loy = mine ? myfy - r : 3e38f, loz = mine ? myfz - r : 3e38f;
        float hix = mine ? myfx + r : -3e38f, hiy = mine ? myfy + r : -3e38f, hiz = mine ? myfz + r : -3e38f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lox = fminf(lox, __shfl_xor_sync(0xffffffffu, lox, o));
            loy = fminf(loy, __shfl_xor_sync(0xffffffffu, loy, o));
            loz = fminf(loz, __shfl_xor_sync(0xffffffffu, loz, o));
            hix = fmaxf(hix, __shfl_xor_sync(0xffffffffu, hix, o));
            hiy = fmaxf(hiy, __shfl_xor_sync(0xffffffffu, hiy, o));
            hiz = fmaxf(hiz, __shfl_xor_sync(0xffffffffu, hiz, o));
        }
        // per axis: the cells c0..c1 (relative to the leaf's) and the faces flo/fhi up to which everything is staged.
        // A face beyond the 5-cell ring is cut to the ring; a face beyond the grid does not exist (nothing lies
        // there).  The cell of a face is taken with a margin: one cell too many is harmless, one too few is not.
        int c0x, c1x, c0y, c1y, c0z, c1z;
        float flx, fhx, fly, fhy, flz, fhz;
        {
            const float eps = 1e-5f * hf;
            auto axis = [&](float lo, float hi, int c, int dim, int& c0, int& c1, float& flo, float& fhi) {
                flo = lo, fhi = hi;
                c0 = (int)floorf((fmaxf(lo, -2.5f * hf) - eps) / hf);
                c1 = (int)floorf((fminf(hi, 3.5f * hf) + eps) / hf);
                if (c0 < -2) c0 = -2, flo = -2.0f * hf;
                if (c1 > 2) c1 = 2, fhi = 3.0f * hf;
                if (c + c0 < 0) c0 = -c, flo = -INF;
                if (c + c1 > dim - 1) c1 = dim - 1 - c, fhi = INF;
            };
            axis(lox, hix, cx, dimx, c0x, c1x, flx, fhx);
            axis(loy, hiy, cy, dimy, c0y, c1y, fly, fhy);
            axis(loz, hiz, cz, dimz, c0z, c1z, flz, fhz);
        }
        const int ncx = c1x - c0x + 1, ncy = c1y - c0y + 1, ncz = c1z - c0z + 1;
        const int NC = ncx * ncy * ncz; // 1 .. 125
        // ---- the runs: lane owns cells 4 lane .. 4 lane + 3 ----
        uint32_t rs[4], rc[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int t = 4 * lane + i;
            rs[i] = 0, rc[i] = 0;
            if (t < NC) {
                const int iz = t / (ncx * ncy), rem = t - iz * (ncx * ncy), iy = rem / ncx, ix = rem - iy * ncx;
                const int nx = cx + c0x + ix, ny = cy + c0y + iy, nz = cz + c0z + iz;
                if (level == 0) {
                    if (!pcp_hash_find(P.htab, P.hmask, pcp_morton(nx, ny, nz), rs[i], rc[i])) rc[i] = 0;
                } else {
                    uint32_t a = 0, cn = 0;
                    if (pcp_hash_find(P.ltab, P.lmask, PCP_LKEY(level, pcp_morton(nx, ny, nz)), a, cn)) {
                        rs[i] = __ldg(P.cstart + a);
                        rc[i] = __ldg(P.cstart + a + cn) - rs[i];
                    }
                }
            }
        }
        const uint32_t mysum = rc[0] + rc[1] + rc[2] + rc[3];
        uint32_t pre = mysum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, pre, o);
            if (lane >= o) pre += v;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
        pre -= mysum;
        PCP_WIDE_DBG(31, total);
        if (total < (uint32_t)P.kk || total > PCP_WIDE_MAXLOAD) {
            PCP_WIDE_DBG(26, 1);
            continue;
        }
        __syncwarp();
        {
            uint32_t c = pre;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                wpre[4 * lane + i] = (4 * lane + i < NC) ? c : 0xffffffffu;
                wstart[4 * lane + i] = rs[i];
                c += rc[i];
            }
        }
        __syncwarp();
        // ---- stage the points inside the box ----
        uint32_t nst = 0;
#pragma unroll 1
        for (uint32_t f0 = 0; f0 < total; f0 += 32 * PCP_WIDE_U) {
            pcp_rec rr[PCP_WIDE_U];
            uint32_t ps[PCP_WIDE_U];
#pragma unroll
            for (int u = 0; u < PCP_WIDE_U; u++) {
                const uint32_t f = f0 + 32 * u + lane;
                ps[u] = 0;
                if (f < total) {
                    int t = 0;
#pragma unroll
                    for (int step = PCP_WIDE_RUNS / 2; step; step >>= 1)
                        if (wpre[t + step] <= f) t += step;
                    ps[u] = wstart[t] + (f - wpre[t]);
                    rr[u] = pcp_load_rec(P.rec, ps[u]);
                }
            }
#pragma unroll
            for (int u = 0; u < PCP_WIDE_U; u++) {
                const uint32_t f = f0 + 32 * u + lane;
                bool in = false;
                float x = 0.f, y = 0.f, z = 0.f;
                if (f < total) {
                    x = (float)(rr[u].x - ox), y = (float)(rr[u].y - oy), z = (float)(rr[u].z - oz);
                    in = x >= flx && x <= fhx && y >= fly && y <= fhy && z >= flz && z <= fhz;
                }
                const unsigned mask = __ballot_sync(0xffffffffu, in);
                const uint32_t at = nst + (uint32_t)__popc(mask & ((1u << lane) - 1u));
                if (in && at < (uint32_t)(CAND - 2 * GS)) S.put(at, x, y, z, ps[u]);
                nst += (uint32_t)__popc(mask);
            }
        }
        PCP_WIDE_DBG(30, nst);
        if (nst > (uint32_t)(CAND - 2 * GS) || nst < (uint32_t)P.kk) {
            PCP_WIDE_DBG(27, 1);
            __syncwarp();
            continue;
        }
        if (lane < (int)(((nst + 2 * GS - 1) & ~(uint32_t)(2 * GS - 1)) - nst)) S.put(nst + lane, 1e18f, 1e18f, 1e18f, 0u);
        __syncwarp();
        // ---- the dense kernel's search, the open queries four at a time ----
        unsigned okmask = 0, rrmask = 0;
        float newU = INF;
        for (unsigned rest = want; rest;) {
            int jj = -1;
#pragma unroll
            for (int g = 0; g < NSET; g++) {
                const int b = rest ? __ffs(rest) - 1 : -1;
                if (g == set) jj = b;
                rest &= rest - 1;
            }
            const bool active = jj >= 0;
            const int sj = active ? jj : 0;
            const float qfx = __shfl_sync(0xffffffffu, myfx, sj), qfy = __shfl_sync(0xffffffffu, myfy, sj),
                        qfz = __shfl_sync(0xffffffffu, myfz, sj);
            const uint32_t qp = __shfl_sync(0xffffffffu, qpos, sj);
            // certified radius: the distance to the faces of the box
            float dout = fminf(fminf(fminf(qfx - flx, fhx - qfx), fminf(qfy - fly, fhy - qfy)), fminf(qfz - flz, fhz - qfz));
            dout = fmaxf(fminf(dout, 3.0e38f) - (4.0f * e1 + 1e-6f * hf), 0.0f);
            const float g2 = dout * dout * 0.999999f;
            uint32_t* const nbr_base = P.nbr_out;
            const uint32_t nbr_off = qp * (uint32_t)P.k; // < 2^32: host
            float dku;
            const int st = pcp_fd_search<GS, CAP, CAND, true>((int)nst, S, lane, active, nbr_base, nbr_off, 1, P.kk, qfx, qfy, qfz,
                                                              top0, g2, e1, dku, P.rr, qp);
            if (active && st != PCP_FD_OK && (lane & (GS - 1)) == 0) nbr_base[nbr_off] = 0xffffffffu;
#pragma unroll
            for (int g = 0; g < NSET; g++) {
                const int jg = __shfl_sync(0xffffffffu, jj, g * GS);
                const int sg = __shfl_sync(0xffffffffu, st, g * GS);
                const float dg = __shfl_sync(0xffffffffu, dku, g * GS);
                if (jg >= 0) {
                    if (sg == PCP_FD_OK) okmask |= 1u << jg;
                    if (sg == PCP_FD_RERANK) rrmask |= 1u << jg;
                    if (sg == PCP_FD_AMBIG && lane == jg) newU = dg * 1.000001f;
                }
            }
        }
        PCP_WIDE_DBG(29, __popc(okmask));
        if (mine) {
            if (((okmask | rrmask) >> lane) & 1u) P.fail_list[wr.x + lane] = 0xffffffffu;
            else if (newU < U) P.fail_u[wr.x + lane] = newU;
        }
        __syncwarp();
    }
}
