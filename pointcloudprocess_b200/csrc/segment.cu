// segment.cu -- TreeExtration::region_growning (extraction_tree.cpp:170-272): plane segments grown over the k = 50
// neighbour graph from seeds taken in ascending point index.
//
// The reference is a sequential loop: the lowest index still in UnSegment seeds a region (if its curvature is below
// 0.005), a FIFO pops members, searches the 50 nearest neighbours of each and admits the unsegmented ones whose
// normal is within acos(cosfaT) of the FIRST seed's and which lie within distanceT of the popped member's plane;
// a region of <= 5 members is dissolved again (its members stay out of UnSegment).  Which region a point joins, the
// labels and the member order all depend on that order, so it is kept:
//
//   * the neighbour lists do not depend on the segmentation state: one batched exact kNN (the dense / far kernels
//     of knn.cu, k = 50) builds the whole graph first, instead of one tree search per popped member;
//   * ONE persistent kernel (cooperative launch, one block per SM) then runs the seed loop.  Block 0 owns the control
//     flow: it scans the pre-compacted list of low-curvature points for the next seed and expands small frontiers on
//     its own (block-level barriers only); a frontier of more than RG_SOLO_PAR members is expanded by ALL blocks: the
//     pass's (member, neighbour rank) slots are cut into contiguous ranges, one per block;
//   * a pass is exact, not just a BFS level: every admissible slot s = member * k + rank bids atomicMin(claim[j], s)
//     for its neighbour j; after a barrier the lowest bidder of each j wins, which is the slot the sequential loop
//     would have admitted j from; an ordered block scan (+ a scan over the per-block counts) appends the winners in
//     slot order, which is the FIFO order of the reference.  Members admitted in a pass are expanded in a later one,
//     exactly as the FIFO would.
//
// Memory written during the kernel by one SM and read by another (seg, claim, the member queue, the control block)
// is read with ld.cg (L2) and published with __threadfence; the static inputs (graph, normals, float coordinates)
// use the normal cached path.
#include "index.cuh"
#include "knn_cell.cuh"
#include "sort.cuh"

#include <climits>
#include <cstdlib>

namespace {

constexpr int RG_T = 512;              // threads per block
constexpr int RG_ITERS = 8;            // slots per thread per pass
constexpr int RG_SLOTS = RG_T * RG_ITERS;
constexpr unsigned RG_NONE = 0xffffffffu;
constexpr unsigned RG_CMD_PASS = 1, RG_CMD_EXIT = 2;

struct RgControl {
    unsigned cmd_gen, cmd;
    unsigned head, npar, tail;
    int label;
    float snx, sny, snz, sn_n;
    unsigned bar_count, bar_gen;
    int error;
    unsigned long long nseg, total, attempts, solo_passes, wide_passes, scanned;
};

struct RgParams {
    const int32_t* nbr; // n x K neighbour lists, ascending (d2, index), -1 padded
    int K;
    const int32_t* row; // graph row of every record's search point (PointProperty[i].PointID, :220-221)
    const float4* nrm;  // normal_x, normal_y, normal_z, (float)Distance
    const float4* ptf;  // coordinates truncated to float (extraction_tree.cpp:51-53)
    int32_t* seg;
    unsigned char* seedable;
    unsigned* claim;
    const uint32_t* cand; // ascending indices of the points with curvature < curv_t
    uint32_t ncand;
    int32_t* queue; // = seg_points
    long long* seg_offsets;
    float* seg_plane;
    unsigned long long seg_cap;
    double distanceT, cosfaT;
    int min_points;
    RgControl* G;
    unsigned* blockcnt;
    uint32_t n;
};

template <class T> __device__ __forceinline__ T rg_ld(const T* p) { return __ldcg(p); }

// all blocks; sense-reversing counter barrier (the kernel is launched cooperatively: every block is resident)
__device__ __forceinline__ void rg_grid_barrier(RgControl* G, unsigned nb)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned gen = *reinterpret_cast<volatile unsigned*>(&G->bar_gen);
        __threadfence();
        if (atomicAdd(&G->bar_count, 1u) == nb - 1) {
            G->bar_count = 0;
            __threadfence();
            atomicAdd(&G->bar_gen, 1u);
        } else {
            while (*reinterpret_cast<volatile unsigned*>(&G->bar_gen) == gen) {
            }
        }
        __threadfence();
    }
    __syncthreads();
}

// compute_included_angle_between_vector (extraction_tree.cpp:25-33) with the seed's length hoisted
__device__ __forceinline__ float rg_cos(float sx, float sy, float sz, float sn_n, float x, float y, float z)
{
    const float n_n1 = __fadd_rn(__fadd_rn(__fmul_rn(sx, x), __fmul_rn(sy, y)), __fmul_rn(sz, z));
    const float n1_n1 = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
    return fabsf(__fdiv_rn(n_n1, __fmul_rn(sn_n, n1_n1)));
}

// compute_distance_from_point_to_plane (extraction_tree.cpp:47-64)
__device__ __forceinline__ float rg_dist(const float4 pt, const float4 pl)
{
    const double g = (double)__fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(pl.x, pl.x), __fmul_rn(pl.y, pl.y)), __fmul_rn(pl.z, pl.z)));
    const double f1 = (double)__fmul_rn(pl.x, pt.x), f2 = (double)__fmul_rn(pl.y, pt.y), f3 = (double)__fmul_rn(pl.z, pt.z);
    const double f = fabs(__dadd_rn(__dadd_rn(__dadd_rn(f1, f2), f3), (double)pl.w));
    return (float)(f / g);
}

struct RgShared {
    unsigned cnt[RG_ITERS * (RG_T / 32)]; // winners per (iteration, warp), then their exclusive prefix
    unsigned block_total, block_off, grand_total;
    unsigned first;
    // control state of block 0
    unsigned cursor, head, tail, start;
    int label, active;
    unsigned long long nseg;
    float snx, sny, snz, sn_n, snd;
    // command seen by this block
    unsigned cmd, c_head, c_npar, c_tail;
    int c_label;
    float c_snx, c_sny, c_snz, c_sn_n;
};

// One pass over the slots of members queue[head .. head+npar): returns the number of members admitted (every thread
// of every participating block gets the grand total).  WIDE: nb blocks share the slots and synchronise through the
// grid barrier; otherwise block `bid` = 0 works alone.
template <bool WIDE>
__device__ __forceinline__ unsigned rg_pass(const RgParams& P, RgShared& S, unsigned head, unsigned npar, unsigned tail,
                                            int label, float snx, float sny, float snz, float sn_n, unsigned nb,
                                            unsigned bid)
{
    const unsigned K = (unsigned)P.K;
    const unsigned nslot = npar * K;
    const unsigned per_block = WIDE ? (nslot + nb - 1) / nb : nslot; // <= RG_SLOTS by the caller's choice of npar
    const unsigned b0 = bid * per_block;
    const unsigned b1 = min(nslot, b0 + per_block);
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned elig = 0;
    int32_t myj[RG_ITERS];
#pragma unroll
    for (int it = 0; it < RG_ITERS; it++) {
        myj[it] = -1;
        const unsigned s = b0 + (unsigned)it * RG_T + tid;
        if (s < b1) {
            const unsigned pi = s / K, e = s - pi * K;
            const int32_t p = rg_ld(P.queue + head + pi);
            const int32_t r = P.row[p];
            const int32_t j = P.nbr[(size_t)r * K + e];
            if (j >= 0 && rg_ld(P.seg + j) == -1) { // :225
                const float4 nj = P.nrm[j];
                const float cosn = rg_cos(snx, sny, snz, sn_n, nj.x, nj.y, nj.z);
                if ((double)cosn > P.cosfaT) {
                    const float dis = rg_dist(P.ptf[j], P.nrm[p]);
                    if ((double)dis < P.distanceT) { // :233
                        atomicMin(P.claim + j, s);
                        elig |= 1u << it;
                        myj[it] = j;
                    }
                }
            }
        }
    }
    if (WIDE) rg_grid_barrier(P.G, nb);
    else __syncthreads();
    // the lowest slot that bid for j is the one the sequential loop admits j from
    unsigned win = 0;
#pragma unroll
    for (int it = 0; it < RG_ITERS; it++) {
        const unsigned s = b0 + (unsigned)it * RG_T + tid;
        const bool w = ((elig >> it) & 1u) && rg_ld(P.claim + myj[it]) == s;
        const unsigned ballot = __ballot_sync(0xffffffffu, w);
        if (w) win |= 1u << it;
        if (lane == 0) S.cnt[it * (RG_T / 32) + warp] = __popc(ballot);
    }
    __syncthreads();
    if (warp == 0) { // exclusive scan of the RG_ITERS * 16 = 128 counts, four per lane
        unsigned v[4], sum = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) v[c] = S.cnt[lane * 4 + c], sum += v[c];
        unsigned incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        unsigned run = incl - sum;
#pragma unroll
        for (int c = 0; c < 4; c++) S.cnt[lane * 4 + c] = run, run += v[c];
        if (lane == 31) {
            S.block_total = incl;
            if (WIDE) P.blockcnt[bid] = incl;
        }
    }
    if (WIDE) {
        rg_grid_barrier(P.G, nb);
        if (warp == 0) {
            unsigned off = 0, tot = 0;
            for (unsigned b = lane; b < nb; b += 32) {
                const unsigned c = rg_ld(P.blockcnt + b);
                tot += c;
                if (b < bid) off += c;
            }
#pragma unroll
            for (int d = 16; d; d >>= 1) off += __shfl_xor_sync(0xffffffffu, off, d), tot += __shfl_xor_sync(0xffffffffu, tot, d);
            if (lane == 0) S.block_off = off, S.grand_total = tot;
        }
        __syncthreads();
    } else {
        __syncthreads();
        if (tid == 0) S.block_off = 0, S.grand_total = S.block_total;
        __syncthreads();
    }
    const unsigned base = tail + S.block_off;
#pragma unroll
    for (int it = 0; it < RG_ITERS; it++) {
        const bool w = (win >> it) & 1u;
        const unsigned ballot = __ballot_sync(0xffffffffu, w);
        if (w) {
            const unsigned pos = base + S.cnt[it * (RG_T / 32) + warp] + __popc(ballot & ((1u << lane) - 1u));
            const int32_t j = myj[it];
            P.queue[pos] = j;
            P.seg[j] = label;
            P.seedable[j] = 0;
            P.claim[j] = RG_NONE;
        }
    }
    const unsigned total = S.grand_total;
    if (WIDE) rg_grid_barrier(P.G, nb);
    else __syncthreads();
    return total;
}

__global__ void __launch_bounds__(RG_T, 1) rg_kernel(const RgParams P)
{
    __shared__ RgShared S;
    const unsigned nb = gridDim.x, bid = blockIdx.x, tid = threadIdx.x;
    const unsigned K = (unsigned)P.K;
    const unsigned solo_par = RG_SLOTS / K;                                     // members one block expands per pass
    const unsigned wide_par = (unsigned)min((unsigned long long)nb * RG_SLOTS / K, 0x7fffffffull / K); // ... all blocks
    unsigned seen_gen = 0;
    if (bid == 0 && tid == 0) {
        S.cursor = 0, S.head = S.tail = S.start = 0, S.label = 0, S.active = 0, S.nseg = 0;
        P.seg_offsets[0] = 0;
    }
    __syncthreads();
    for (;;) {
        if (bid == 0) {
            // ---- control: runs until a wide pass is needed or the seeds are exhausted
            for (;;) {
                if (S.active && S.tail - S.head > 2 * solo_par && nb > 1) {
                    if (tid == 0) {
                        RgControl* G = P.G;
                        G->cmd = RG_CMD_PASS;
                        G->head = S.head, G->npar = min(S.tail - S.head, wide_par), G->tail = S.tail, G->label = S.label;
                        G->snx = S.snx, G->sny = S.sny, G->snz = S.snz, G->sn_n = S.sn_n;
                        G->wide_passes++;
                    }
                    break;
                }
                if (S.active && S.head < S.tail) {
                    const unsigned head = S.head, tail = S.tail, npar = min(tail - head, solo_par);
                    const unsigned got = rg_pass<false>(P, S, head, npar, tail, S.label, S.snx, S.sny, S.snz, S.sn_n, 1, 0);
                    if (tid == 0) S.head = head + npar, S.tail = tail + got, P.G->solo_passes++;
                    __syncthreads();
                    continue;
                }
                if (S.active) { // the FIFO ran dry: keep or dissolve the region (:249-262)
                    const unsigned size = S.tail - S.start;
                    const bool keep = (long long)size > (long long)P.min_points;
                    if (!keep)
                        for (unsigned q = tid; q < size; q += RG_T) P.seg[rg_ld(P.queue + S.start + q)] = -1;
                    __syncthreads();
                    if (tid == 0) {
                        if (keep) {
                            if (S.nseg >= P.seg_cap) {
                                P.G->error = 1;
                                S.cursor = P.ncand; // stop
                            } else {
                                float* pl = P.seg_plane + 4 * S.nseg;
                                pl[0] = S.snx, pl[1] = S.sny, pl[2] = S.snz, pl[3] = S.snd;
                                S.nseg++;
                                P.seg_offsets[S.nseg] = (long long)S.tail;
                                S.label++;
                                S.start = S.tail;
                            }
                        }
                        S.head = S.tail = S.start;
                        S.active = 0;
                    }
                    __syncthreads();
                    continue;
                }
                // ---- next seed: the first candidate at or after the cursor that is still in UnSegment (:190-194)
                if (S.cursor >= P.ncand) {
                    if (tid == 0) {
                        RgControl* G = P.G;
                        G->cmd = RG_CMD_EXIT;
                        G->nseg = S.nseg, G->total = S.start;
                    }
                    break;
                }
                if (tid == 0) S.first = RG_T;
                __syncthreads();
                const unsigned c = S.cursor + tid;
                const bool ok = c < P.ncand && rg_ld(P.seedable + P.cand[c]) != 0;
                if (ok) atomicMin(&S.first, tid);
                __syncthreads();
                if (tid == 0) {
                    P.G->scanned += RG_T;
                    if (S.first == RG_T) {
                        S.cursor += RG_T;
                    } else {
                        const uint32_t r = P.cand[S.cursor + S.first];
                        S.cursor += S.first + 1;
                        const float4 nr = P.nrm[r];
                        S.snx = nr.x, S.sny = nr.y, S.snz = nr.z, S.snd = nr.w;
                        S.sn_n = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(nr.x, nr.x), __fmul_rn(nr.y, nr.y)), __fmul_rn(nr.z, nr.z)));
                        P.seg[r] = S.label;
                        P.seedable[r] = 0;
                        P.queue[S.start] = (int32_t)r;
                        S.head = S.start, S.tail = S.start + 1;
                        S.active = 1;
                        P.G->attempts++;
                    }
                }
                __syncthreads();
            }
            // publish the command
            __syncthreads();
            if (tid == 0) {
                __threadfence();
                atomicAdd(&P.G->cmd_gen, 1u);
            }
        }
        // ---- every block: wait for the command
        if (tid == 0) {
            seen_gen++;
            while ((int)(*reinterpret_cast<volatile unsigned*>(&P.G->cmd_gen) - seen_gen) < 0) {
            }
            __threadfence();
            const RgControl* G = P.G;
            S.cmd = rg_ld(&G->cmd);
            S.c_head = rg_ld(&G->head), S.c_npar = rg_ld(&G->npar), S.c_tail = rg_ld(&G->tail), S.c_label = rg_ld(&G->label);
            S.c_snx = rg_ld(&G->snx), S.c_sny = rg_ld(&G->sny), S.c_snz = rg_ld(&G->snz), S.c_sn_n = rg_ld(&G->sn_n);
        }
        __syncthreads();
        if (S.cmd == RG_CMD_EXIT) return;
        const unsigned head = S.c_head, npar = S.c_npar, tail = S.c_tail;
        const unsigned got = rg_pass<true>(P, S, head, npar, tail, S.c_label, S.c_snx, S.c_sny, S.c_snz, S.c_sn_n, nb, bid);
        if (bid == 0) {
            if (tid == 0) S.head = head + npar, S.tail = tail + got;
            __syncthreads();
        }
    }
}

// normals + float coordinates in original order, candidate flags, validity of the PointID rows
__global__ void __launch_bounds__(256) rg_prepare_kernel(const unsigned char* __restrict__ props, uint32_t n, double curv_t,
                                                         float4* nrm, int32_t* row, uint32_t* flag, int32_t* seg,
                                                         unsigned char* seedable, unsigned* claim, int* bad_row)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i > n) return;
    if (i == n) {
        flag[n] = 0;
        return;
    }
    const unsigned char* r = props + (size_t)i * 48;
    const float4 a = *reinterpret_cast<const float4*>(r);            // normal_x, normal_y, normal_z, pad
    const double2 b = *reinterpret_cast<const double2*>(r + 16);     // Distance, curvature
    const int32_t pid = *reinterpret_cast<const int32_t*>(r + 32);   // PointID
    nrm[i] = make_float4(a.x, a.y, a.z, (float)b.x);
    row[i] = pid;
    if (pid < 0 || (uint32_t)pid >= n) *bad_row = 1;
    flag[i] = (b.y < curv_t) ? 1u : 0u; // :194
    seg[i] = -1;
    seedable[i] = 1;
    claim[i] = RG_NONE;
}

__global__ void __launch_bounds__(256) rg_points_kernel(const pcp_rec* __restrict__ rec, uint32_t nv, float4* ptf)
{
    const uint32_t p = blockIdx.x * 256u + threadIdx.x;
    if (p >= nv) return;
    const pcp_rec r = pcp_load_rec(rec, p);
    ptf[(size_t)r.idx] = make_float4((float)r.x, (float)r.y, (float)r.z, 0.f);
}

__global__ void __launch_bounds__(256) rg_compact_kernel(const uint32_t* __restrict__ flag, const uint32_t* __restrict__ off,
                                                         uint32_t n, uint32_t* cand)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i < n && flag[i]) cand[off[i]] = i;
}

__global__ void __launch_bounds__(256) rg_writeback_kernel(unsigned char* props, const int32_t* __restrict__ seg, uint32_t n)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i < n) *reinterpret_cast<int32_t*>(props + (size_t)i * 48 + 36) = seg[i];
}

} // namespace

extern "C" int pcp_region_growing(const pcp_index* ix, void* props, int k, double distanceT, double cosfaT, double curv_t,
                                  int min_points, int64_t* seg_offsets, size_t seg_cap, int32_t* seg_points,
                                  float* seg_plane, size_t* nseg_out, uint64_t* stats)
{
    if (!ix || !props || !seg_offsets || !seg_points || !seg_plane || !nseg_out || k < 1 || seg_cap < 1)
        return pcp_fail(ix ? ix->ctx : nullptr, PCP_ERR_INVALID, "pcp_region_growing: bad argument");
    pcp_ctx* ctx = ix->ctx;
    if (k > 128) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_region_growing: k > 128 is not supported");
    const size_t n0 = ix->n_orig;
    if (n0 >= 0x7fffffffull) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_region_growing: cloud too large");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    *nseg_out = 0;
    cudaStream_t st = ctx->stream;
    OutStage o_props, o_off, o_pts, o_plane;
    PCP_TRY(o_off.open(ctx, seg_offsets, (seg_cap + 1) * 8));
    if (n0 == 0) {
        PCP_CUDA_TRY(ctx, cudaMemsetAsync(o_off.dptr, 0, 8, st));
        return o_off.commit(8);
    }
    PCP_TRY(o_props.open(ctx, props, n0 * 48, true));
    PCP_TRY(o_pts.open(ctx, seg_points, n0 * 4));
    PCP_TRY(o_plane.open(ctx, seg_plane, seg_cap * 16));
    DevBuf<int32_t> nbr, row, seg;
    DevBuf<unsigned char> scratch_props, seedable;
    DevBuf<float4> nrm, ptf;
    DevBuf<uint32_t> flag, off, cand;
    DevBuf<unsigned> claim, blockcnt;
    DevBuf<RgControl> G;
    DevBuf<int> bad;
    PCP_TRY(nbr.alloc(ctx, n0 * (size_t)k));
    PCP_TRY(scratch_props.alloc(ctx, n0 * 48));
    // the neighbour graph: nearestKSearch(point, k) of every point (:222), ascending (d2, index)
    PCP_TRY(pcp_features_knn_ex(ix, k, -1, scratch_props.p, nullptr, nullptr, nbr.p, nullptr));
    scratch_props.release();
    PCP_TRY(row.alloc(ctx, n0));
    PCP_TRY(seg.alloc(ctx, n0));
    PCP_TRY(seedable.alloc(ctx, n0));
    PCP_TRY(nrm.alloc(ctx, n0));
    PCP_TRY(ptf.alloc(ctx, n0));
    PCP_TRY(flag.alloc(ctx, n0 + 1));
    PCP_TRY(off.alloc(ctx, n0 + 1));
    PCP_TRY(claim.alloc(ctx, n0));
    PCP_TRY(G.alloc(ctx, 1));
    PCP_TRY(bad.alloc(ctx, 1));
    uint32_t ncand = 0;
    {
        StageTimer t(ctx, "segment.prepare", (double)n0 * 100);
        PCP_CUDA_TRY(ctx, cudaMemsetAsync(G.p, 0, sizeof(RgControl), st));
        PCP_CUDA_TRY(ctx, cudaMemsetAsync(bad.p, 0, sizeof(int), st));
        PCP_CUDA_TRY(ctx, cudaMemsetAsync(ptf.p, 0, n0 * sizeof(float4), st));
        rg_prepare_kernel<<<pcp_blocks_for(n0 + 1, 256), 256, 0, st>>>(static_cast<const unsigned char*>(o_props.dptr),
                                                                       (uint32_t)n0, curv_t, nrm.p, row.p, flag.p, seg.p,
                                                                       seedable.p, claim.p, bad.p);
        PCP_LAUNCH_CHECK(ctx);
        if (ix->n) {
            rg_points_kernel<<<pcp_blocks_for(ix->n, 256), 256, 0, st>>>(ix->rec, (uint32_t)ix->n, ptf.p);
            PCP_LAUNCH_CHECK(ctx);
        }
        PCP_TRY(pcp_exclusive_scan_u32(ctx, flag.p, off.p, n0 + 1));
        struct {
            uint32_t ncand;
            int bad;
        } rb;
        PCP_TRY(pcp_readback(ctx, &rb.ncand, off.p + n0, 4));
        PCP_TRY(pcp_readback(ctx, &rb.bad, bad.p, 4));
        if (rb.bad) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_region_growing: a PointID lies outside the cloud");
        ncand = rb.ncand;
        PCP_TRY(cand.alloc(ctx, ncand));
        rg_compact_kernel<<<pcp_blocks_for(n0, 256), 256, 0, st>>>(flag.p, off.p, (uint32_t)n0, cand.p);
        PCP_LAUNCH_CHECK(ctx);
    }
    RgControl res;
    {
        StageTimer t(ctx, "segment.grow", (double)n0 * ((double)k * 4 + 48));
        int nb = ctx->sm_count, per_sm = 0;
        PCP_CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rg_kernel, RG_T, 0));
        if (per_sm < 1) return pcp_fail(ctx, PCP_ERR_CUDA, "pcp_region_growing: the kernel does not fit an SM");
        if (const char* e = getenv("PCP_RG_BLOCKS")) nb = atoi(e) > 0 && atoi(e) < nb ? atoi(e) : nb;
        PCP_TRY(blockcnt.alloc(ctx, (size_t)nb));
        RgParams P;
        P.nbr = nbr.p, P.K = k, P.row = row.p, P.nrm = nrm.p, P.ptf = ptf.p;
        P.seg = seg.p, P.seedable = seedable.p, P.claim = claim.p;
        P.cand = cand.p, P.ncand = ncand;
        P.queue = static_cast<int32_t*>(o_pts.dptr);
        P.seg_offsets = static_cast<long long*>(o_off.dptr);
        P.seg_plane = static_cast<float*>(o_plane.dptr);
        P.seg_cap = seg_cap;
        P.distanceT = distanceT, P.cosfaT = cosfaT, P.min_points = min_points;
        P.G = G.p, P.blockcnt = blockcnt.p, P.n = (uint32_t)n0;
        void* args[] = {&P};
        PCP_CUDA_TRY(ctx, cudaLaunchCooperativeKernel((const void*)rg_kernel, dim3((unsigned)nb), dim3(RG_T), args, 0, st));
        PCP_LAUNCH_CHECK(ctx);
        rg_writeback_kernel<<<pcp_blocks_for(n0, 256), 256, 0, st>>>(static_cast<unsigned char*>(o_props.dptr), seg.p, (uint32_t)n0);
        PCP_LAUNCH_CHECK(ctx);
        PCP_TRY(pcp_readback(ctx, &res, G.p, sizeof(RgControl)));
    }
    if (res.error) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_region_growing: more than seg_cap = %zu regions", seg_cap);
    *nseg_out = (size_t)res.nseg;
    if (stats) {
        stats[0] = res.attempts, stats[1] = res.solo_passes, stats[2] = res.wide_passes, stats[3] = res.scanned;
        stats[4] = res.total, stats[5] = ncand;
    }
    PCP_TRY(o_off.commit((size_t)(res.nseg + 1) * 8));
    PCP_TRY(o_pts.commit((size_t)res.total * 4));
    PCP_TRY(o_plane.commit((size_t)res.nseg * 16));
    return o_props.commit();
}
