// sort.cu -- the key sort that rebuilds the voxel order (voxel_grid.h:948, std::sort of the cell indices) and the
// Morton search structure (kd_tree.h:795-797, FLANN's tree build), hand-written for sm_100a: a stable LSD radix sort
// of (key, u32 value) pairs in the "onesweep" shape -- ONE read of the keys for the digit histograms of every pass,
// then one read + one write of the pairs per pass with a chained (decoupled look-back) scan of the per-tile digit
// counts -- plus the exclusive prefix sums the callers need.
//
// What it knows about its keys (cell keys and Morton codes):
//   * the caller passes the number of significant key bits; the passes split them EVENLY (42 bits -> 6 x 7, not
//     5 x 8 + 2), so every pass ranks against as few bins as possible.  PCP_SORT_DIGIT9=1 allows 9-bit digits where
//     that saves a whole pass (42 bits -> 5 passes): measured at 25 M pairs it gains 10 % on clustered keys (Morton
//     codes of points in voxel order) and loses 8 % on keys in random order (shorter runs per bin), so it is off;
//   * a pass whose histogram is a single bin (the high cell / Morton bits of a slab of a larger cloud, a flat cloud's
//     z bits) permutes nothing and is skipped outright -- the histograms of all passes are known before the first one;
//   * values are the point indices 0..n-1 on entry of every caller, so the first executed pass generates them instead
//     of reading them (4 of 24 bytes per pair and pass).
// A pass: 512 threads rank a tile of 4096 pairs.  Keys are loaded warp-striped, every warp counts its digits in its
// own shared histogram, thread d turns column d into warp offsets, publishes the tile's count of digit d and looks
// back over the earlier tiles' counts (32-bit words: 2 flag bits + 30 value bits, hence n < 2^30); ranks come from
// __match_any_sync in load order (stable); pairs are exchanged through shared memory so that the global writes are
// runs of consecutive addresses per digit.  Tiles are handed out by an atomic counter, so a tile only ever waits for
// tiles that are already running.
//
// PCP_SORT_CUB=1 routes the same calls to CUB (sort_cub.cu), kept as the A/B reference of tests/test_gpu_sort.py.
#include "sort.cuh"
#include <cstdlib>

int pcp_cub_sort_pairs_u64(pcp_ctx* ctx, unsigned long long** keys, unsigned long long** keys_alt, uint32_t** vals,
                           uint32_t** vals_alt, size_t n, int bits);
int pcp_cub_sort_pairs_u32(pcp_ctx* ctx, uint32_t** keys, uint32_t** keys_alt, uint32_t** vals, uint32_t** vals_alt, size_t n,
                           int bits);

namespace {

constexpr int RS_THREADS = 512;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 8;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS; // 4096 pairs
constexpr int RS_MAXPASS = 8;
constexpr int RS_NBMAX = 512; // bins of the widest digit (9 bits); histograms are laid out with this stride
constexpr uint32_t RS_AGG = 1u << 30, RS_PRE = 2u << 30, RS_VAL = (1u << 30) - 1u;

struct RsPlan {
    int npass;
    int shift[RS_MAXPASS];
    int bits[RS_MAXPASS];
};

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" : : "l"(p), "r"(v) : "memory");
}

// ---- digit histograms of every pass: one read of the keys ---------------------------------------------------
// Shared-memory atomics cost 2.7 cycles per warp instruction when the 32 lanes hit different words and 32 when they hit
// one (measured, tools/ubench/match_atoms.cu).  Equal digits across a warp are the rule for the high digits of cell
// keys and Morton codes (sorted or nearly sorted input, constant top bits), so the warp first ORs together the bits in
// which its keys differ from lane 0's (two redux instructions per key, whatever the number of passes): a pass whose
// digit bits do not differ at all is counted by one lane.
template <typename K>
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const K* __restrict__ keys, uint32_t n, RsPlan plan,
                                                             uint32_t* __restrict__ ghist)
{
    __shared__ uint32_t sh[RS_MAXPASS * RS_NBMAX];
    for (int i = threadIdx.x; i < plan.npass * RS_NBMAX; i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    constexpr int U = 4;
    const uint32_t stride = gridDim.x * RS_THREADS;
    const uint32_t nround = (n + 31u) & ~31u; // whole warps stay in the loop together
    for (uint32_t i0 = blockIdx.x * RS_THREADS + threadIdx.x; i0 < nround; i0 += U * stride) {
        K key[U];
        bool valid[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint32_t i = i0 + u * stride;
            valid[u] = i < n;
            key[u] = valid[u] ? keys[i] : (K)0;
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (i0 + u * stride - lane >= n) break; // uniform per warp: lane 0 is beyond the input
            // (the valid lanes are a prefix of the warp; the others take lane 0's key and count nothing)
            const K k0 = __shfl_sync(0xffffffffu, key[u], 0);
            const K x = valid[u] ? (key[u] ^ k0) : (K)0;
            K diff = (K)__reduce_or_sync(0xffffffffu, (uint32_t)x);
            if constexpr (sizeof(K) == 8) diff |= (K)__reduce_or_sync(0xffffffffu, (uint32_t)(x >> 32)) << 32;
            const uint32_t nvalid = (uint32_t)__popc(__ballot_sync(0xffffffffu, valid[u]));
#pragma unroll
            for (int p = 0; p < RS_MAXPASS; p++) { // unrolled: the plan stays in the parameter bank
                if (p >= plan.npass) break;
                const uint32_t m = (1u << plan.bits[p]) - 1u;
                const uint32_t d = (uint32_t)(key[u] >> plan.shift[p]) & m;
                if (((uint32_t)(diff >> plan.shift[p]) & m) == 0u) {
                    if (lane == 0) atomicAdd(&sh[p * RS_NBMAX + d], nvalid);
                } else if (valid[u]) {
                    atomicAdd(&sh[p * RS_NBMAX + d], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < plan.npass * RS_NBMAX; i += RS_THREADS)
        if (sh[i]) atomicAdd(&ghist[i], sh[i]);
}

// exclusive scan of each pass's bins (in place) and the mask of the passes that a single bin makes trivial
__global__ void __launch_bounds__(RS_NBMAX) rs_scan_kernel(uint32_t* __restrict__ ghist, int npass, uint32_t n,
                                                           uint32_t* __restrict__ skipmask)
{
    __shared__ uint32_t ws[RS_NBMAX / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t mask = 0;
    for (int p = 0; p < npass; p++) {
        const uint32_t c = ghist[p * RS_NBMAX + threadIdx.x];
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += v;
        }
        __syncthreads(); // ws of the pass before is read no more
        if (lane == 31) ws[warp] = incl;
        __syncthreads();
        uint32_t wpre = 0;
        for (uint32_t w = 0; w < warp; w++) wpre += ws[w];
        ghist[p * RS_NBMAX + threadIdx.x] = wpre + incl - c;
        if (__syncthreads_or(c == n)) mask |= 1u << p;
    }
    if (threadIdx.x == 0) *skipmask = mask;
}

// ---- one pass --------------------------------------------------------------------------------------------------
// 512 threads rank a tile of 4096 pairs, 8 per thread, warp-striped (item i of lane l = element warp*256 + 32*i + l of
// the tile, so ranking in (warp, item, lane) order is ranking in input order: stable).
//   counts   every warp counts its digits in its own row of s_cnt;
//   offsets  thread d (256 of them) turns column d into the warps' starts inside bin d, publishes the tile's count of
//            digit d and looks back over the earlier tiles for the bin's global start;
//   ranks    item by item: the lanes with the same digit ("peers") take consecutive places after the warp's running
//            offset.  How the peers are found is chosen per warp from the data (measured on B200, cycles per warp
//            instruction and SM: match.any 2.4 for one distinct value, 15 for 8, 62 for 32 random ones; a shared-memory
//            atomic 2.7 for 32 distinct addresses, 32 for one): __match_any_sync when the warp's digits are few
//            (nearly sorted input, high digits), otherwise an atomicOr of the lane bits into a per-digit mask word;
//   exchange keys and values go through shared memory in rank order, so the global writes are runs per digit.
// IOTA: the values are 0..n-1 in order (not read).
constexpr int RS_MATCH_MAX_DISTINCT = 8;
constexpr int RS_LB = 8; // look-back words in flight per bin

template <typename K, int NB> struct RsSmem {
    static constexpr size_t CNT = 0;                                  // [RS_WARPS][NB] u32
    static constexpr size_t GOFF = CNT + (size_t)RS_WARPS * NB * 4;  // [NB] u32
    static constexpr size_t WSUM = GOFF + NB * 4;                    // [RS_WARPS] u32 + tile
    static constexpr size_t EX = WSUM + 128;                          // keys [RS_TILE] K, then values [RS_TILE] u32
    static constexpr size_t MASK_BYTES = (size_t)2 * RS_WARPS * NB * 4; // peer masks, double-buffered: alias EX
    static constexpr size_t EX_BYTES = (size_t)RS_TILE * (sizeof(K) + 4);
    static constexpr size_t BYTES = EX + (EX_BYTES > MASK_BYTES ? EX_BYTES : MASK_BYTES);
};

template <typename K, bool IOTA, int NB>
__global__ void __launch_bounds__(RS_THREADS, 2)
    rs_pass_kernel(const K* __restrict__ kin, const uint32_t* __restrict__ vin, K* __restrict__ kout,
                   uint32_t* __restrict__ vout, uint32_t n, int shift, int bits, const uint32_t* __restrict__ digit_base,
                   uint32_t* __restrict__ lookback, uint32_t* __restrict__ tile_counter)
{
    extern __shared__ __align__(16) unsigned char rs_smem[];
    using L = RsSmem<K, NB>;
    uint32_t(*s_cnt)[NB] = reinterpret_cast<uint32_t(*)[NB]>(rs_smem + L::CNT);
    uint32_t* s_goff = reinterpret_cast<uint32_t*>(rs_smem + L::GOFF);
    uint32_t* s_wsum = reinterpret_cast<uint32_t*>(rs_smem + L::WSUM);
    K* s_ex = reinterpret_cast<K*>(rs_smem + L::EX);
    uint32_t* s_v = reinterpret_cast<uint32_t*>(rs_smem + L::EX + (size_t)RS_TILE * sizeof(K));
    uint32_t* s_mask = reinterpret_cast<uint32_t*>(rs_smem + L::EX); // [2][RS_WARPS][NB], dead before the exchange
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_wsum[RS_WARPS] = atomicAdd(tile_counter, 1u);
    {
        const uint4 z = make_uint4(0, 0, 0, 0);
        uint4* c4 = reinterpret_cast<uint4*>(rs_smem + L::CNT);
        for (uint32_t i = tid; i < RS_WARPS * NB / 4; i += RS_THREADS) c4[i] = z;
        uint4* m4 = reinterpret_cast<uint4*>(s_mask);
        for (uint32_t i = tid; i < 2 * RS_WARPS * NB / 4; i += RS_THREADS) m4[i] = z;
    }
    __syncthreads();
    const uint32_t tile = s_wsum[RS_WARPS];
    const uint32_t base = tile * (uint32_t)RS_TILE;
    const uint32_t cnt = min((uint32_t)RS_TILE, n - base);
    const uint32_t dmask = (1u << bits) - 1u;
    // Elements beyond the input are all-ones keys: they fall into the last bin behind every real key (the ranking is
    // stable), i.e. at tile positions >= cnt of the digit-sorted tile, and are never written.
    K key[RS_ITEMS];
    uint32_t val[RS_ITEMS];
    const uint32_t t0 = warp * (32u * RS_ITEMS) + lane;
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
        const uint32_t t = t0 + 32u * i;
        key[i] = t < cnt ? kin[base + t] : ~(K)0;
    }
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
        const uint32_t t = t0 + 32u * i;
        if constexpr (IOTA) val[i] = base + t;
        else val[i] = t < cnt ? vin[base + t] : 0u;
    }
    // ---- counts; the first and the last item tell how many distinct digits the warp's instructions see ----
    const uint32_t ltmask = (1u << lane) - 1u;
    unsigned peers[RS_ITEMS];
    bool use_match;
    {
        const uint32_t d0 = (uint32_t)(key[0] >> shift) & dmask, d1 = (uint32_t)(key[RS_ITEMS - 1] >> shift) & dmask;
        peers[0] = __match_any_sync(0xffffffffu, d0);
        peers[RS_ITEMS - 1] = __match_any_sync(0xffffffffu, d1);
        const int nd0 = __popc(__ballot_sync(0xffffffffu, (peers[0] & ltmask) == 0u));
        const int nd1 = __popc(__ballot_sync(0xffffffffu, (peers[RS_ITEMS - 1] & ltmask) == 0u));
        use_match = max(nd0, nd1) <= RS_MATCH_MAX_DISTINCT;
    }
    if (use_match) {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; i++) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
            if (i != 0 && i != RS_ITEMS - 1) peers[i] = __match_any_sync(0xffffffffu, d);
            if ((peers[i] & ltmask) == 0u) atomicAdd(&s_cnt[warp][d], (uint32_t)__popc(peers[i])); // one lane per digit
        }
    } else {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; i++) atomicAdd(&s_cnt[warp][(uint32_t)(key[i] >> shift) & dmask], 1u);
    }
    __syncthreads();
    uint32_t pub = 0, bin_start = 0;
    if (tid < NB) {
        // thread = bin: offsets of the warps inside the bin, the tile's count (published at once), the bins' starts
        uint32_t tot = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) {
            const uint32_t v = s_cnt[w][tid];
            s_cnt[w][tid] = tot;
            tot += v;
        }
        pub = tot - (tid == dmask ? (uint32_t)RS_TILE - cnt : 0u); // the padding is not data
        st_relaxed(lookback + (size_t)tile * NB + tid, pub | (tile ? RS_AGG : RS_PRE));
        uint32_t incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += v;
        }
        if (lane == 31) s_wsum[warp] = incl;
        if constexpr (NB == RS_THREADS) __syncthreads();
        else asm volatile("bar.sync 1, %0;" ::"n"(NB) : "memory"); // the warps of the bins only
        uint32_t wpre = 0;
        for (uint32_t w = 0; w < warp; w++) wpre += s_wsum[w];
        bin_start = wpre + incl - tot;
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) s_cnt[w][tid] += bin_start;
    }
    __syncthreads();
    // ---- ranks in input order ----
    uint32_t rank[RS_ITEMS];
    if (use_match) {
#pragma unroll
        for (int i = 0; i < RS_ITEMS; i++) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
            const uint32_t b = s_cnt[warp][d];
            __syncwarp();
            rank[i] = b + (uint32_t)__popc(peers[i] & ltmask);
            if (lane == (uint32_t)(31 - __clz(peers[i]))) s_cnt[warp][d] = b + (uint32_t)__popc(peers[i]);
            __syncwarp();
        }
    } else {
        // peers by atomicOr into the digit's mask word; two buffers in turn, so the leader's clearing of item i's word
        // is ordered before its next use (item i + 2) by item i + 1's barriers
        uint32_t* mk = s_mask + warp * NB;
#pragma unroll
        for (int i = 0; i < RS_ITEMS; i++) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & dmask;
            uint32_t* m = mk + (i & 1) * (RS_WARPS * NB) + d;
            atomicOr(m, 1u << lane);
            __syncwarp();
            const uint32_t pm = *m;
            const uint32_t b = s_cnt[warp][d];
            __syncwarp();
            rank[i] = b + (uint32_t)__popc(pm & ltmask);
            if (lane == (uint32_t)(31 - __clz(pm))) {
                s_cnt[warp][d] = b + (uint32_t)__popc(pm);
                *m = 0u;
            }
        }
    }
    __syncthreads(); // the masks are dead: their memory takes the keys
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
        s_ex[rank[i]] = key[i];
        s_v[rank[i]] = val[i];
    }
    if (tid < NB) {
        // The look-back, as late as possible (the earlier tiles have had the whole ranking to publish) and RS_LB words
        // at a time: a walk over the counts of the tiles before this one, back to the nearest tile that already knows
        // its prefix, is a chain of dependent L2 round trips, and one tile per trip made the first wave of tiles
        // (two per SM, all started together) the longest part of a pass.
        uint32_t ex = 0;
        if (tile) {
            int t = (int)tile - 1;
            bool done = false;
            while (!done) {
                uint32_t v[RS_LB];
#pragma unroll
                for (int j = 0; j < RS_LB; j++) v[j] = t - j >= 0 ? ld_relaxed(lookback + (size_t)(t - j) * NB + tid) : RS_PRE;
#pragma unroll
                for (int j = 0; j < RS_LB; j++) {
                    if (done || (v[j] >> 30) == 0u) break; // not published yet: load again from there
                    ex += v[j] & RS_VAL;
                    t--;
                    done = (v[j] & RS_PRE) != 0u; // tile 0 always publishes a prefix; t < 0 reads as an empty prefix
                }
            }
            st_relaxed(lookback + (size_t)tile * NB + tid, (ex + pub) | RS_PRE);
        }
        s_goff[tid] = digit_base[tid] + ex - bin_start; // + tile position = global position (mod 2^32 arithmetic)
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < RS_ITEMS; j++) {
        const uint32_t p = (uint32_t)j * RS_THREADS + tid;
        const K kk = s_ex[p];
        const uint32_t g = s_goff[(uint32_t)(kk >> shift) & dmask] + p;
        if (p < cnt) {
            kout[g] = kk;
            vout[g] = s_v[p];
        }
    }
}

__global__ void __launch_bounds__(256) rs_iota_kernel(uint32_t* __restrict__ v, uint32_t n)
{
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i < n) v[i] = i;
}

template <typename K, int NB>
int launch_pass(pcp_ctx* ctx, bool iota, const K* kin, const uint32_t* vin, K* kout, uint32_t* vout, uint32_t nn, uint32_t ntile,
                int shift, int bits, const uint32_t* digit_base, uint32_t* lookback, uint32_t* counter)
{
    constexpr size_t smem = RsSmem<K, NB>::BYTES;
    if (iota) {
        PCP_CUDA_TRY(ctx, cudaFuncSetAttribute(rs_pass_kernel<K, true, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rs_pass_kernel<K, true, NB><<<ntile, RS_THREADS, smem, ctx->stream>>>(kin, vin, kout, vout, nn, shift, bits, digit_base,
                                                                            lookback, counter);
    } else {
        PCP_CUDA_TRY(ctx, cudaFuncSetAttribute(rs_pass_kernel<K, false, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rs_pass_kernel<K, false, NB><<<ntile, RS_THREADS, smem, ctx->stream>>>(kin, vin, kout, vout, nn, shift, bits, digit_base,
                                                                             lookback, counter);
    }
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

template <typename K>
int sort_pairs(pcp_ctx* ctx, K** keys, K** keys_alt, uint32_t** vals, uint32_t** vals_alt, size_t n, int bits, bool iota)
{
    if (n == 0) return PCP_OK;
    if (n >= ((size_t)1 << 30)) return pcp_fail(ctx, PCP_ERR_RANGE, "sort: 2^30 or more items in one call");
    if (bits < 1) bits = 1;
    if (bits > (int)(8 * sizeof(K))) bits = (int)(8 * sizeof(K));
    // digits of at most 8 bits (9 with PCP_SORT_DIGIT9, see the header); the bits are split evenly, the first
    // (bits % npass) passes take one bit more
    static const int max_digit = getenv("PCP_SORT_DIGIT9") ? 9 : 8;
    RsPlan plan;
    plan.npass = (bits + 7) / 8;
    if ((bits + max_digit - 1) / max_digit < plan.npass) plan.npass = (bits + max_digit - 1) / max_digit;
    for (int p = 0, at = 0; p < plan.npass; p++) {
        plan.bits[p] = bits / plan.npass + (p < bits % plan.npass ? 1 : 0);
        plan.shift[p] = at;
        at += plan.bits[p];
    }
    const uint32_t nn = (uint32_t)n;
    const uint32_t ntile = (nn + RS_TILE - 1) / RS_TILE;
    // [npass][512] histograms, [8] tile counters, [8] skip mask and spare, [npass][ntile][512] look-back words
    const size_t head = (size_t)plan.npass * RS_NBMAX + 16;
    const size_t words = head + (size_t)plan.npass * ntile * RS_NBMAX;
    DevBuf<uint32_t> work;
    PCP_TRY(work.alloc(ctx, words));
    PCP_CUDA_TRY(ctx, cudaMemsetAsync(work.p, 0, words * 4, ctx->stream));
    uint32_t* ghist = work.p;
    uint32_t* counters = work.p + (size_t)plan.npass * RS_NBMAX;
    uint32_t* skip_dev = counters + 8;
    uint32_t* lookback = work.p + head;
    int hb = ctx->sm_count * 4;
    const int need = (int)((nn + RS_THREADS * 4 - 1) / (RS_THREADS * 4));
    if (hb > need) hb = need;
    rs_hist_kernel<K><<<hb, RS_THREADS, 0, ctx->stream>>>(*keys, nn, plan, ghist);
    PCP_LAUNCH_CHECK(ctx);
    rs_scan_kernel<<<1, RS_NBMAX, 0, ctx->stream>>>(ghist, plan.npass, nn, skip_dev);
    PCP_LAUNCH_CHECK(ctx);
    uint32_t skip = 0;
    PCP_TRY(pcp_readback(ctx, &skip, skip_dev, 4));
    K *kin = *keys, *kout = *keys_alt;
    uint32_t *vin = *vals, *vout = *vals_alt;
    bool first = true;
    for (int p = 0; p < plan.npass; p++) {
        if (skip & (1u << p)) continue;
        uint32_t* lb = lookback + (size_t)p * ntile * RS_NBMAX;
        if (plan.bits[p] > 8)
            PCP_TRY((launch_pass<K, 512>(ctx, first && iota, kin, vin, kout, vout, nn, ntile, plan.shift[p], plan.bits[p],
                                         ghist + p * RS_NBMAX, lb, counters + p)));
        else
            PCP_TRY((launch_pass<K, 256>(ctx, first && iota, kin, vin, kout, vout, nn, ntile, plan.shift[p], plan.bits[p],
                                         ghist + p * RS_NBMAX, lb, counters + p)));
        first = false;
        K* tk = kin;
        kin = kout, kout = tk;
        uint32_t* tv = vin;
        vin = vout, vout = tv;
    }
    if (first && iota) { // nothing moved (every pass trivial): the values still have to exist
        rs_iota_kernel<<<pcp_blocks_for(nn, 256), 256, 0, ctx->stream>>>(vin, nn);
        PCP_LAUNCH_CHECK(ctx);
    }
    *keys = kin, *keys_alt = kout, *vals = vin, *vals_alt = vout;
    return PCP_OK;
}

// ---- exclusive prefix sums: per-tile sums, one block over the tile sums, per-tile scan ---------------------------
constexpr int SC_THREADS = 256;
constexpr int SC_ITEMS = 16;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;

template <typename T> __device__ __forceinline__ T sc_shfl_up(T v, int o) { return __shfl_up_sync(0xffffffffu, v, o); }
template <typename T> __device__ __forceinline__ T sc_shfl(T v, int l) { return __shfl_sync(0xffffffffu, v, l); }

// exclusive scan over the block of one value per thread; total returned to every thread
template <typename T> __device__ __forceinline__ T sc_block_exclusive(T v, T* s_w /*[8]*/, T& total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const T u = sc_shfl_up(incl, o);
        if (lane >= o) incl += u;
    }
    __syncthreads(); // s_w may still be read from an earlier call
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    T wpre = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < SC_THREADS / 32; w++) {
        const T u = s_w[w];
        if (w < warp) wpre += u;
        tot += u;
    }
    total = tot;
    return wpre + incl - v;
}

template <typename T>
__global__ void __launch_bounds__(SC_THREADS) sc_reduce_kernel(const T* __restrict__ in, size_t n, T* __restrict__ sums)
{
    __shared__ T s_w[SC_THREADS / 32];
    const size_t b0 = (size_t)blockIdx.x * SC_TILE;
    T acc = 0;
#pragma unroll
    for (int j = 0; j < SC_ITEMS; j++) {
        const size_t i = b0 + (size_t)j * SC_THREADS + threadIdx.x;
        if (i < n) acc += in[i];
    }
    T total;
    sc_block_exclusive<T>(acc, s_w, total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

template <typename T> __global__ void __launch_bounds__(SC_THREADS) sc_spine_kernel(T* __restrict__ sums, size_t nb)
{
    __shared__ T s_w[SC_THREADS / 32];
    T carry = 0;
    for (size_t c0 = 0; c0 < nb; c0 += SC_THREADS) {
        const size_t i = c0 + threadIdx.x;
        const T v = i < nb ? sums[i] : (T)0;
        T total;
        const T ex = sc_block_exclusive<T>(v, s_w, total);
        if (i < nb) sums[i] = carry + ex;
        carry += total;
    }
}

__device__ __forceinline__ int sc_pad(int p) { return p + (p >> 4); } // 16 consecutive items per thread: stride 17

template <typename T>
__global__ void __launch_bounds__(SC_THREADS) sc_apply_kernel(const T* __restrict__ in, T* __restrict__ out, size_t n,
                                                              const T* __restrict__ offs)
{
    __shared__ T s_w[SC_THREADS / 32];
    __shared__ T s_x[SC_TILE + SC_TILE / 16];
    const size_t b0 = (size_t)blockIdx.x * SC_TILE;
    // coalesced load into shared memory, then thread t scans its 16 consecutive items
#pragma unroll
    for (int j = 0; j < SC_ITEMS; j++) {
        const size_t i = b0 + (size_t)j * SC_THREADS + threadIdx.x;
        s_x[sc_pad(j * SC_THREADS + (int)threadIdx.x)] = i < n ? in[i] : (T)0;
    }
    __syncthreads();
    T v[SC_ITEMS], acc = 0;
#pragma unroll
    for (int j = 0; j < SC_ITEMS; j++) {
        v[j] = s_x[sc_pad((int)threadIdx.x * SC_ITEMS + j)];
        acc += v[j];
    }
    T total;
    T ex = sc_block_exclusive<T>(acc, s_w, total) + offs[blockIdx.x];
#pragma unroll
    for (int j = 0; j < SC_ITEMS; j++) {
        s_x[sc_pad((int)threadIdx.x * SC_ITEMS + j)] = ex; // a thread rewrites only the slots it read
        ex += v[j];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < SC_ITEMS; j++) {
        const size_t i = b0 + (size_t)j * SC_THREADS + threadIdx.x;
        if (i < n) out[i] = s_x[sc_pad(j * SC_THREADS + (int)threadIdx.x)];
    }
}

template <typename T> int excl_scan(pcp_ctx* ctx, const T* in, T* out, size_t n)
{
    if (n == 0) return PCP_OK;
    const size_t nb = (n + SC_TILE - 1) / SC_TILE;
    if (nb > 0x7fffffffull) return pcp_fail(ctx, PCP_ERR_RANGE, "scan: too many items");
    DevBuf<T> sums;
    PCP_TRY(sums.alloc(ctx, nb));
    sc_reduce_kernel<T><<<(unsigned)nb, SC_THREADS, 0, ctx->stream>>>(in, n, sums.p);
    PCP_LAUNCH_CHECK(ctx);
    sc_spine_kernel<T><<<1, SC_THREADS, 0, ctx->stream>>>(sums.p, nb);
    PCP_LAUNCH_CHECK(ctx);
    sc_apply_kernel<T><<<(unsigned)nb, SC_THREADS, 0, ctx->stream>>>(in, out, n, sums.p);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

bool use_cub()
{
    static const bool v = getenv("PCP_SORT_CUB") != nullptr;
    return v;
}

} // namespace

int pcp_sort_pairs_u64(pcp_ctx* ctx, unsigned long long** keys, unsigned long long** keys_alt, uint32_t** vals,
                       uint32_t** vals_alt, size_t n, int bits, bool iota_vals)
{
    if (use_cub()) {
        if (iota_vals && n) {
            rs_iota_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(*vals, (uint32_t)n);
            PCP_LAUNCH_CHECK(ctx);
        }
        return pcp_cub_sort_pairs_u64(ctx, keys, keys_alt, vals, vals_alt, n, bits);
    }
    return sort_pairs<unsigned long long>(ctx, keys, keys_alt, vals, vals_alt, n, bits, iota_vals);
}
int pcp_sort_pairs_u32(pcp_ctx* ctx, uint32_t** keys, uint32_t** keys_alt, uint32_t** vals, uint32_t** vals_alt, size_t n,
                       int bits, bool iota_vals)
{
    if (use_cub()) {
        if (iota_vals && n) {
            rs_iota_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(*vals, (uint32_t)n);
            PCP_LAUNCH_CHECK(ctx);
        }
        return pcp_cub_sort_pairs_u32(ctx, keys, keys_alt, vals, vals_alt, n, bits);
    }
    return sort_pairs<uint32_t>(ctx, keys, keys_alt, vals, vals_alt, n, bits, iota_vals);
}

int pcp_exclusive_scan_u32(pcp_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n) { return excl_scan(ctx, in, out, n); }
int pcp_exclusive_scan_u64(pcp_ctx* ctx, const unsigned long long* in, unsigned long long* out, size_t n)
{
    return excl_scan(ctx, in, out, n);
}
// (tree order: used for predictions only, never for results -- see serial_sum.cu)
int pcp_exclusive_scan_f64(pcp_ctx* ctx, const double* in, double* out, size_t n) { return excl_scan(ctx, in, out, n); }

// In-place sort of n (key, value) pairs by the low `bits` bits of the key, stable; host or device arrays.
// key_bytes: 4 or 8.  vals == NULL sorts (key, index) pairs and drops the indices.
extern "C" int pcp_sort_pairs(pcp_ctx* ctx, void* keys, uint32_t* vals, size_t n, int key_bytes, int bits)
{
    if (!ctx || (n && !keys) || (key_bytes != 4 && key_bytes != 8))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_sort_pairs: bad argument");
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    OutStage k, v;
    PCP_TRY(k.open(ctx, keys, n * (size_t)key_bytes, true));
    PCP_TRY(v.open(ctx, vals, n * 4, true));
    DevBuf<char> kalt;
    DevBuf<uint32_t> valt, vtmp;
    PCP_TRY(kalt.alloc(ctx, n * (size_t)key_bytes));
    PCP_TRY(valt.alloc(ctx, n));
    uint32_t* v0 = static_cast<uint32_t*>(v.dptr);
    if (!v0) {
        PCP_TRY(vtmp.alloc(ctx, n));
        v0 = vtmp.p;
    }
    uint32_t* v1 = valt.p;
    StageTimer t(ctx, "sort.pairs", (double)n * (key_bytes + 4) * 2);
    if (key_bytes == 8) {
        unsigned long long *k0 = static_cast<unsigned long long*>(k.dptr), *k1 = reinterpret_cast<unsigned long long*>(kalt.p);
        PCP_TRY(pcp_sort_pairs_u64(ctx, &k0, &k1, &v0, &v1, n, bits, vals == nullptr));
        if (k0 != k.dptr) PCP_CUDA_TRY(ctx, cudaMemcpyAsync(k.dptr, k0, n * 8, cudaMemcpyDeviceToDevice, ctx->stream));
    } else {
        uint32_t *k0 = static_cast<uint32_t*>(k.dptr), *k1 = reinterpret_cast<uint32_t*>(kalt.p);
        PCP_TRY(pcp_sort_pairs_u32(ctx, &k0, &k1, &v0, &v1, n, bits, vals == nullptr));
        if (k0 != k.dptr) PCP_CUDA_TRY(ctx, cudaMemcpyAsync(k.dptr, k0, n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    if (v.dptr && v0 != v.dptr) PCP_CUDA_TRY(ctx, cudaMemcpyAsync(v.dptr, v0, n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    PCP_TRY(k.commit());
    PCP_TRY(v.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

// Exclusive prefix sums of n uint32 (host or device arrays; out may equal in).
extern "C" int pcp_exclusive_sum_u32(pcp_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n)
{
    if (!ctx || (n && (!in || !out))) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_exclusive_sum_u32: bad argument");
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage i;
    OutStage o;
    PCP_TRY(i.open(ctx, in, n * 4));
    PCP_TRY(o.open(ctx, out, n * 4));
    PCP_TRY(pcp_exclusive_scan_u32(ctx, static_cast<const uint32_t*>(i.dptr), static_cast<uint32_t*>(o.dptr), n));
    PCP_TRY(o.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}

// ---- the prefix sum for callers that must not allocate -------------------------------------------------------------
// pcp_exclusive_sum_u32_noalloc: the same three kernels as pcp_exclusive_sum_u32, with the per-tile sums (the scan's
// only scratch) in a fixed array in device memory instead of the context's arena.  No counterpart in the reference;
// added for callers that cannot call cudaMalloc (CUDA graph capture, tight
// serving loops).  Every GPU that loads the library holds the array (64 KiB) until the process exits.
//   * PCP_NOALLOC_SCAN_MAX_N items at most (one scratch word per tile of 4096 items);
//   * in / out must be DEVICE pointers on the context's GPU (a host array would need a staging buffer);
//   * nothing in the array has to start at zero: sc_reduce_kernel writes every word [0, tiles) that the later kernels
//     read, so the contents a previous call left behind are never seen;
//   * every call on a device shares the one array: calls that overlap on the GPU (two streams, two contexts on the
//     same device) race on it.  Calls on one stream are ordered and safe.
static __device__ __align__(256) uint32_t g_noalloc_scan_sums[PCP_NOALLOC_SCAN_MAX_N / SC_TILE];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_ptr[NOALLOC_MAX_DEVICES]; // device address of the array, per device (NULL: not looked up yet)
} // namespace

extern "C" int pcp_exclusive_sum_u32_noalloc(pcp_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n)
{
    static_assert(PCP_NOALLOC_SCAN_MAX_N % SC_TILE == 0, "pcp_exclusive_sum_u32_noalloc: the maximum is a whole number of tiles");
    if (!ctx || (n && (!in || !out))) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_exclusive_sum_u32_noalloc: bad argument");
    if (n > (size_t)PCP_NOALLOC_SCAN_MAX_N)
        return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_exclusive_sum_u32_noalloc: %zu items exceed the fixed scratch's maximum of %d "
                                            "(PCP_NOALLOC_SCAN_MAX_N); use pcp_exclusive_sum_u32", n, PCP_NOALLOC_SCAN_MAX_N);
    if (n == 0) return PCP_OK;
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
        return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_exclusive_sum_u32_noalloc: device %d beyond the %d supported", ctx->device,
                        NOALLOC_MAX_DEVICES);
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (pcp_classify(ctx, in) != PCP_PTR_DEVICE || pcp_classify(ctx, out) != PCP_PTR_DEVICE)
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_exclusive_sum_u32_noalloc: in and out must be device pointers on the "
                                              "context's GPU (staging a host array would allocate)");
    uint32_t* sums = nullptr;
    {
        // the array's name is a host-side stand-in here: its device address is looked up once per device, with that
        // device current (set above)
        std::lock_guard<std::mutex> g(g_noalloc_mu);
        if (!g_noalloc_ptr[ctx->device]) {
            void* p = nullptr;
            PCP_CUDA_TRY(ctx, cudaGetSymbolAddress(&p, g_noalloc_scan_sums));
            if (!p) return pcp_fail(ctx, PCP_ERR_CUDA, "pcp_exclusive_sum_u32_noalloc: cudaGetSymbolAddress returned NULL");
            g_noalloc_ptr[ctx->device] = static_cast<uint32_t*>(p);
        }
        sums = g_noalloc_ptr[ctx->device];
    }
    const size_t nb = (n + SC_TILE - 1) / SC_TILE;
    sc_reduce_kernel<uint32_t><<<(unsigned)nb, SC_THREADS, 0, ctx->stream>>>(in, n, sums);
    PCP_LAUNCH_CHECK(ctx);
    sc_spine_kernel<uint32_t><<<1, SC_THREADS, 0, ctx->stream>>>(sums, nb);
    PCP_LAUNCH_CHECK(ctx);
    sc_apply_kernel<uint32_t><<<(unsigned)nb, SC_THREADS, 0, ctx->stream>>>(in, out, n, sums);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK; // asynchronous on the context's stream, like pcp_memcpy_async: pcp_synchronize to wait
}
