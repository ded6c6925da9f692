// index.cu -- builds the Morton-sorted multi-level grid (see index.cuh).  Replaces
// KdTreeFLANN::setInputCloud / convertCloudToArray (kd_tree.h:772-798, 928-956): non-finite points are
// dropped, every record keeps its ORIGINAL cloud index (index_mapping_).
#include "index.cuh"
#include "reduce.cuh"
#include "sort.cuh"
#include <cfloat>
#include <cmath>
#include <algorithm>
#include <cstdlib>

namespace {

__global__ void __launch_bounds__(256) morton_key_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                         pcp_grid g, unsigned long long* __restrict__ keys,
                                                         uint32_t* __restrict__ vals)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* p = reinterpret_cast<const double*>(pts + i * stride);
    double x = p[0], y = p[1], z = p[2];
    unsigned long long key = PCP_EMPTY_KEY;
    if (isfinite(x) && isfinite(y) && isfinite(z)) {
        int cx = min(max(pcp_cell_coord(x, g.org[0], g.inv_h, g.dims[0]), 0), g.dims[0] - 1);
        int cy = min(max(pcp_cell_coord(y, g.org[1], g.inv_h, g.dims[1]), 0), g.dims[1] - 1);
        int cz = min(max(pcp_cell_coord(z, g.org[2], g.inv_h, g.dims[2]), 0), g.dims[2] - 1);
        key = pcp_morton(cx, cy, cz);
    }
    keys[i] = key;
    if (vals) vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) gather_rec_kernel(const char* __restrict__ pts, size_t stride,
                                                         const uint32_t* __restrict__ vals, size_t nv,
                                                         pcp_rec* __restrict__ rec, float4* __restrict__ fpt,
                                                         double ox, double oy, double oz, int vec)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= nv) return;
    uint32_t o = vals[i];
    const char* src = pts + (size_t)o * stride;
    pcp_d4 v;
    if (vec) { // packed 32-byte records on a 32-byte boundary: one 256-bit load
        v = pcp_ld256(src);
    } else {
        const double* p = reinterpret_cast<const double*>(src);
        v.a = p[0], v.b = p[1], v.c = p[2];
    }
    v.d = __longlong_as_double((long long)o);
    pcp_st256(rec + i, v); // (x, y, z, idx): the record, in one store
    fpt[i] = make_float4((float)(v.a - ox), (float)(v.b - oy), (float)(v.c - oz), 0.f);
}

__global__ void __launch_bounds__(256) cell_flag_kernel(const unsigned long long* __restrict__ keys, size_t nv,
                                                        uint32_t* __restrict__ flag)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= nv) return;
    flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}

__global__ void __launch_bounds__(256) cell_fill_kernel(const unsigned long long* __restrict__ keys,
                                                        const uint32_t* __restrict__ flag,
                                                        const uint32_t* __restrict__ scan, size_t nv,
                                                        unsigned long long* __restrict__ ckey,
                                                        uint32_t* __restrict__ cstart, uint32_t* __restrict__ ncell_out)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= nv) return;
    if (flag[i]) {
        ckey[scan[i]] = keys[i];
        cstart[scan[i]] = (uint32_t)i;
    }
    if (i == nv - 1) {
        uint32_t m = scan[i] + flag[i];
        cstart[m] = (uint32_t)nv;
        *ncell_out = m;
    }
}

// point-weighted occupancy histogram: bin floor(log2(count)) += count (block-local bins, one atomic per bin)
__global__ void __launch_bounds__(256) occupancy_kernel(const uint32_t* __restrict__ cstart, uint32_t ncell,
                                                        unsigned long long* __restrict__ bins)
{
    __shared__ unsigned long long sb[32];
    if (threadIdx.x < 32) sb[threadIdx.x] = 0ull;
    __syncthreads();
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < ncell; c += gridDim.x * blockDim.x) {
        const uint32_t cnt = cstart[c + 1] - cstart[c];
        atomicAdd(&sb[31 - __clz(cnt)], (unsigned long long)cnt);
    }
    __syncthreads();
    if (threadIdx.x < 32 && sb[threadIdx.x]) atomicAdd(&bins[threadIdx.x], sb[threadIdx.x]);
}

__global__ void __launch_bounds__(256) hash_insert_kernel(const unsigned long long* __restrict__ ckey,
                                                          const uint32_t* __restrict__ cstart, uint32_t ncell,
                                                          pcp_hslot* __restrict__ htab, uint32_t hmask)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    unsigned long long key = ckey[c];
    uint32_t s = pcp_hash64(key) & hmask;
    while (true) {
        unsigned long long prev = atomicCAS(&htab[s].key, PCP_EMPTY_KEY, key);
        if (prev == PCP_EMPTY_KEY) {
            htab[s].start = cstart[c];
            htab[s].count = cstart[c + 1] - cstart[c];
            return;
        }
        s = (s + 1) & hmask;
    }
}

// ---- level table: every cell of levels 0..PCP_HLEVELS-1 as a run of base cells ----
__device__ __forceinline__ bool lvl_differs(const unsigned long long* __restrict__ ckey, uint32_t a, uint32_t b, int l)
{
    return (ckey[a] >> (3 * l)) != (ckey[b] >> (3 * l));
}

__global__ void __launch_bounds__(256) lhash_count_kernel(const unsigned long long* __restrict__ ckey, uint32_t ncell,
                                                          unsigned long long* __restrict__ total)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned n = 0;
    if (c < ncell)
        for (int l = 0; l < PCP_HLEVELS; l++) n += (c == 0 || lvl_differs(ckey, c - 1, c, l)) ? 1u : 0u;
    n = __reduce_add_sync(0xffffffffu, n);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(total, (unsigned long long)n);
}

__global__ void __launch_bounds__(256) lhash_heads_kernel(const unsigned long long* __restrict__ ckey, uint32_t ncell,
                                                          pcp_hslot* __restrict__ tab, uint32_t mask)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    for (int l = 0; l < PCP_HLEVELS; l++) {
        if (!(c == 0 || lvl_differs(ckey, c - 1, c, l))) continue;
        const unsigned long long key = PCP_LKEY(l, ckey[c] >> (3 * l));
        uint32_t s = pcp_hash64(key) & mask;
        while (true) {
            const unsigned long long prev = atomicCAS(&tab[s].key, PCP_EMPTY_KEY, key);
            if (prev == PCP_EMPTY_KEY) {
                tab[s].start = c;
                tab[s].count = 0;
                break;
            }
            s = (s + 1) & mask;
        }
    }
}

__global__ void __launch_bounds__(256) lhash_tails_kernel(const unsigned long long* __restrict__ ckey, uint32_t ncell,
                                                          pcp_hslot* __restrict__ tab, uint32_t mask)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    for (int l = 0; l < PCP_HLEVELS; l++) {
        if (!(c == ncell - 1 || lvl_differs(ckey, c, c + 1, l))) continue;
        const unsigned long long key = PCP_LKEY(l, ckey[c] >> (3 * l));
        uint32_t s = pcp_hash64(key) & mask;
        while (tab[s].key != key) s = (s + 1) & mask;
        tab[s].count = c + 1 - tab[s].start;
    }
}

__global__ void __launch_bounds__(256) cinfo_kernel(const unsigned long long* __restrict__ ckey,
                                                    const uint32_t* __restrict__ cstart, uint32_t ncell,
                                                    uint4* __restrict__ cinfo)
{
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    int x, y, z;
    pcp_demorton(ckey[c], x, y, z);
    cinfo[c] = make_uint4((unsigned)x, (unsigned)y, (unsigned)z, cstart[c]);
}

// ---- work items of the dense kNN kernel: the leaves of the implicit octree with capacity cmax ----
// A base cell with more than cmax points is cut into chunks of 32 queries (level 0).  Sparser cells are merged
// upwards: the leaf of base cell c is its largest ancestor (level <= PCP_ITEM_LMAX) that still holds at most
// cmax points; the first base cell of the leaf emits the one item that carries all of the leaf's points, so
// that the 3x3x3 block staged for them grows with the local point spacing (knn_fast.cuh).
// PCP_ITEM_CMAX / PCP_ITEM_CMIN (experiments, tools/sweep_items.sh): cmin > 0 also makes a node of at most 32 points an
// item when one of its children holds fewer than cmin -- measured: no better than a larger cmax (38.9-39.1 ms).
__device__ __forceinline__ int leaf_of_cell(const unsigned long long* __restrict__ ckey,
                                            const uint32_t* __restrict__ cstart, const pcp_hslot* __restrict__ ltab,
                                            uint32_t lmask, uint32_t c, uint32_t cmax_cmin, uint32_t& ca, uint32_t& npts)
{
    // A node is an item when it holds at most cmax points, or at most 32 with a child of fewer than cmin; a base cell's
    // leaf is its LARGEST such ancestor (every base cell below a node sees the same ancestors: consistent).
    const uint32_t cmax = cmax_cmin & 0xffffu, cmin = cmax_cmin >> 16;
    int level = 0;
    ca = c;
    npts = cstart[c + 1] - cstart[c];
    if (npts > cmax) return 0;
    const unsigned long long key = ckey[c];
    for (int l = 1; l <= PCP_ITEM_LMAX; l++) {
        uint32_t a = 0, cn = 0;
        if (!pcp_hash_find(ltab, lmask, PCP_LKEY(l, key >> (3 * l)), a, cn)) break;
        const uint32_t np = cstart[a + cn] - cstart[a];
        if (np > 32u || (np > cmax && cmin == 0u)) break;
        bool ok = np <= cmax;
        if (!ok) { // does a child (level l - 1) hold fewer than cmin points?
            uint32_t run = 0;
            unsigned long long ck = ckey[a] >> (3 * (l - 1));
            for (uint32_t b = a; b < a + cn && !ok; b++) {
                const unsigned long long k2 = ckey[b] >> (3 * (l - 1));
                if (k2 != ck) {
                    ok = run < cmin;
                    ck = k2, run = 0;
                }
                run += cstart[b + 1] - cstart[b];
            }
            if (!ok) ok = run < cmin;
        }
        if (ok) level = l, ca = a, npts = np;
    }
    return level;
}

__global__ void __launch_bounds__(256) item_count_kernel(const unsigned long long* __restrict__ ckey,
                                                         const uint32_t* __restrict__ cstart,
                                                         const pcp_hslot* __restrict__ ltab, uint32_t lmask,
                                                         uint32_t ncell, uint32_t cmax, uint32_t* __restrict__ nchunk)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    uint32_t ca, npts;
    const int level = leaf_of_cell(ckey, cstart, ltab, lmask, c, cmax, ca, npts);
    nchunk[c] = level == 0 ? (npts + 31) >> 5 : (ca == c ? 1u : 0u);
}

__global__ void __launch_bounds__(256) item_fill_kernel(const unsigned long long* __restrict__ ckey,
                                                        const uint32_t* __restrict__ cstart,
                                                        const pcp_hslot* __restrict__ ltab, uint32_t lmask,
                                                        const uint32_t* __restrict__ nchunk,
                                                        const uint32_t* __restrict__ off, uint32_t ncell, uint32_t cmax,
                                                        uint32_t* __restrict__ item_cell,
                                                        uint32_t* __restrict__ item_first,
                                                        uint32_t* __restrict__ item_meta, uint32_t* __restrict__ nitem)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncell) return;
    const uint32_t o = off[c], k = nchunk[c], s = cstart[c];
    if (k) {
        uint32_t ca, npts;
        const int level = leaf_of_cell(ckey, cstart, ltab, lmask, c, cmax, ca, npts);
        for (uint32_t j = 0; j < k; j++) {
            item_cell[o + j] = c;
            item_first[o + j] = s + 32 * j;
            item_meta[o + j] = ((uint32_t)level << 8) | min(32u, npts - 32 * j);
        }
    }
    if (c == ncell - 1) *nitem = o + k;
}

int ilog2_ceil(int v)
{
    int b = 0;
    while ((1 << b) < v) b++;
    return b;
}

void grid_from_h(pcp_grid& g, const pcp_bbox& bb, double h)
{
    double ext[3];
    double emax = 0;
    for (int d = 0; d < 3; d++) {
        ext[d] = bb.mx[d] - bb.mn[d];
        if (!(ext[d] >= 0)) ext[d] = 0;
        emax = fmax(emax, ext[d]);
    }
    // at most 2^21 cells per axis (63-bit Morton codes)
    double hmin = emax / 2097150.0;
    if (!(h > hmin)) h = hmin > 0 ? hmin : 1.0;
    int dmax = 1;
    for (int d = 0; d < 3; d++) {
        g.org[d] = bb.mn[d];
        g.dims[d] = (int)(ext[d] / h) + 1;
        dmax = g.dims[d] > dmax ? g.dims[d] : dmax;
    }
    g.h = h;
    g.inv_h = 1.0 / h;
    g.nbits = ilog2_ceil(dmax);
    if (g.nbits < 1) g.nbits = 1;
    g.top_level = g.nbits; // at level nbits every coordinate >> level is 0
}

} // namespace

static void index_release(pcp_index* ix)
{
    if (!ix) return;
    pcp_ctx* c = ix->ctx;
    pcp_arena_free(c, ix->rec);
    pcp_arena_free(c, ix->fpt);
    pcp_arena_free(c, ix->ckey);
    pcp_arena_free(c, ix->cstart);
    pcp_arena_free(c, ix->htab);
    pcp_arena_free(c, ix->ltab);
    pcp_arena_free(c, ix->cinfo);
    pcp_arena_free(c, ix->item_cell);
    pcp_arena_free(c, ix->item_first);
    pcp_arena_free(c, ix->item_meta);
    delete ix;
}

int pcp_index_build_items(pcp_ctx* ctx, pcp_index* ix, uint32_t cmax)
{
    if (cmax < 1) cmax = 1;
    if (cmax > 32) cmax = 32;
    {
        static const char* e1 = getenv("PCP_ITEM_CMAX");
        static const char* e2 = getenv("PCP_ITEM_CMIN");
        if (e1 && atoi(e1) > 0) cmax = (uint32_t)std::min(atoi(e1), 32);
        if (e2 && atoi(e2) > 0) cmax |= (uint32_t)std::min(atoi(e2), 32) << 16;
    }
    if (ix->item_cell && ix->item_cmax == cmax) return PCP_OK;
    StageTimer t(ctx, "index.items");
    const uint32_t ncell = ix->ncell;
    pcp_arena_free(ctx, ix->item_cell);
    pcp_arena_free(ctx, ix->item_first);
    pcp_arena_free(ctx, ix->item_meta);
    ix->item_cell = ix->item_first = ix->item_meta = nullptr;
    ix->nitem = 0;
    DevBuf<uint32_t> nchunk, off, cnt;
    PCP_TRY(nchunk.alloc(ctx, ncell));
    PCP_TRY(off.alloc(ctx, ncell));
    PCP_TRY(cnt.alloc(ctx, 1));
    item_count_kernel<<<pcp_blocks_for(ncell, 256), 256, 0, ctx->stream>>>(ix->ckey, ix->cstart, ix->ltab, ix->lmask,
                                                                         ncell, cmax, nchunk.p);
    ctx->launches++;
    PCP_CUDA_TRY(ctx, cudaGetLastError());
    PCP_TRY(pcp_exclusive_scan_u32(ctx, nchunk.p, off.p, ncell));
    // nitem <= ncell + n/32
    const size_t cap_items = (size_t)ncell + ix->n / 32 + 1;
    void* q = nullptr;
    PCP_TRY(pcp_arena_alloc(ctx, cap_items * 4, &q));
    ix->item_cell = (uint32_t*)q;
    PCP_TRY(pcp_arena_alloc(ctx, cap_items * 4, &q));
    ix->item_first = (uint32_t*)q;
    PCP_TRY(pcp_arena_alloc(ctx, cap_items * 4, &q));
    ix->item_meta = (uint32_t*)q;
    item_fill_kernel<<<pcp_blocks_for(ncell, 256), 256, 0, ctx->stream>>>(ix->ckey, ix->cstart, ix->ltab, ix->lmask,
                                                                        nchunk.p, off.p, ncell, cmax, ix->item_cell,
                                                                        ix->item_first, ix->item_meta, cnt.p);
    ctx->launches++;
    PCP_CUDA_TRY(ctx, cudaGetLastError());
    uint32_t ni = 0;
    PCP_TRY(pcp_readback(ctx, &ni, cnt.p, 4));
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ix->nitem = ni;
    ix->item_cmax = cmax;
    return PCP_OK;
}

int pcp_index_build_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, double cell_size, pcp_index** out)
{
    *out = nullptr;
    if (n > (size_t)0x7fffffff) return pcp_fail(ctx, PCP_ERR_RANGE, "index: more than 2^31-1 points");
    pcp_index* ix = new pcp_index();
    ix->ctx = ctx;
    ix->n_orig = n;
    pcp_bbox bb;
    memset(&bb, 0, sizeof(bb));
    if (n) {
        StageTimer t(ctx, "index.bbox", 24.0 * n);
        int s = pcp_bbox_host(ctx, dpts, n, stride, &bb);
        if (s != PCP_OK) {
            delete ix;
            return s;
        }
    }
    ix->n = (size_t)bb.count;
    size_t nv = ix->n;
    if (nv == 0) {
        memset(&ix->grid, 0, sizeof(ix->grid));
        ix->grid.h = ix->grid.inv_h = 1.0;
        *out = ix;
        return PCP_OK;
    }
    // initial guess: points on a surface spanning the two largest extents, ~16 points per cell
    double e[3] = {bb.mx[0] - bb.mn[0], bb.mx[1] - bb.mn[1], bb.mx[2] - bb.mn[2]};
    double a = fmax(e[0] * e[1], fmax(e[0] * e[2], e[1] * e[2]));
    double h = cell_size > 0 ? cell_size : sqrt(16.0 * a / (double)nv);
    if (!(h > 0) || !std::isfinite(h)) h = 1.0;
    // Warm start: the cell size the data-driven search below settled on for the previous cloud of this context,
    // scaled by the change of the areal density.  Successive clouds of a job (tiles, shards, passes) are alike, so
    // the first grid usually passes the occupancy test and the second sort is saved; the test itself is unchanged.
    const double areal = a > 0 ? (double)nv / a : 0.0;
    if (!(cell_size > 0) && ctx->hint_h > 0 && areal > 0 && ctx->hint_areal > 0) {
        const double ratio = ctx->hint_areal / areal;
        if (ratio > 0.25 && ratio < 4.0) h = ctx->hint_h * sqrt(ratio);
    }

    DevBuf<unsigned long long> k0, k1, bins;
    DevBuf<uint32_t> v0, v1, flag, scan, ncell_dev;
    int st = PCP_OK;
#define IX_TRY(expr)                                                                                         \
    do {                                                                                                     \
        st = (expr);                                                                                         \
        if (st != PCP_OK) {                                                                                  \
            index_release(ix);                                                                               \
            return st;                                                                                       \
        }                                                                                                    \
    } while (0)
#define IX_CUDA(expr)                                                                                        \
    do {                                                                                                     \
        cudaError_t _e = (expr);                                                                             \
        if (_e != cudaSuccess) {                                                                             \
            index_release(ix);                                                                               \
            return pcp_fail(ctx, PCP_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e));              \
        }                                                                                                    \
    } while (0)
    IX_TRY(k0.alloc(ctx, n));
    IX_TRY(k1.alloc(ctx, n));
    IX_TRY(v0.alloc(ctx, n));
    IX_TRY(v1.alloc(ctx, n));
    IX_TRY(flag.alloc(ctx, nv));
    IX_TRY(scan.alloc(ctx, nv));
    IX_TRY(ncell_dev.alloc(ctx, 2));
    IX_TRY(bins.alloc(ctx, 32));
    unsigned long long *keys = k0.p, *keys_alt = k1.p;
    uint32_t *vals = v0.p, *vals_alt = v1.p;
    uint32_t ncell = 0;
    const int max_iter = cell_size > 0 ? 1 : 6;
    for (int iter = 0; iter < max_iter; iter++) {
        grid_from_h(ix->grid, bb, h);
        h = ix->grid.h;
        keys = k0.p, keys_alt = k1.p, vals = v0.p, vals_alt = v1.p;
        {
            StageTimer t(ctx, "index.keys", 24.0 * n);
            morton_key_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(dpts), n,
                                                                             stride, ix->grid, keys, nullptr); // values: the sort's iota
            ctx->launches++;
            IX_CUDA(cudaGetLastError());
        }
        {
            StageTimer t(ctx, "index.sort");
            int bits = 3 * ix->grid.nbits + (nv < n ? 1 : 0);
            if (nv < n && bits < 64) bits = 64; // the all-ones key of dropped points must sort last
            IX_TRY(pcp_sort_pairs_u64(ctx, &keys, &keys_alt, &vals, &vals_alt, n, bits > 64 ? 64 : bits, true));
        }
        {
            StageTimer t(ctx, "index.cells");
            cell_flag_kernel<<<pcp_blocks_for(nv, 256), 256, 0, ctx->stream>>>(keys, nv, flag.p);
            ctx->launches++;
            IX_CUDA(cudaGetLastError());
            IX_TRY(pcp_exclusive_scan_u32(ctx, flag.p, scan.p, nv));
            // ncell = scan[nv-1] + flag[nv-1]
            uint32_t tail[2];
            IX_TRY(pcp_readback(ctx, &tail[0], scan.p + (nv - 1), 4));
            IX_TRY(pcp_readback(ctx, &tail[1], flag.p + (nv - 1), 4));
            IX_CUDA(cudaStreamSynchronize(ctx->stream));
            ncell = tail[0] + tail[1];
        }
        pcp_arena_free(ctx, ix->ckey);
        pcp_arena_free(ctx, ix->cstart);
        ix->ckey = nullptr;
        ix->cstart = nullptr;
        void* q = nullptr;
        IX_TRY(pcp_arena_alloc(ctx, (size_t)ncell * 8, &q));
        ix->ckey = (unsigned long long*)q;
        IX_TRY(pcp_arena_alloc(ctx, ((size_t)ncell + 1) * 4, &q));
        ix->cstart = (uint32_t*)q;
        {
            StageTimer t(ctx, "index.cells");
            cell_fill_kernel<<<pcp_blocks_for(nv, 256), 256, 0, ctx->stream>>>(keys, flag.p, scan.p, nv, ix->ckey,
                                                                             ix->cstart, ncell_dev.p);
            ctx->launches++;
            IX_CUDA(cudaGetLastError());
        }
        ix->ncell = ncell;
        if (iter + 1 >= max_iter) break;
        // data-driven cell size: aim the point-weighted median occupancy at ~16 points per cell
        unsigned long long hb[32];
        IX_CUDA(cudaMemsetAsync(bins.p, 0, 32 * 8, ctx->stream));
        occupancy_kernel<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(ix->cstart, ncell, bins.p);
        ctx->launches++;
        IX_CUDA(cudaGetLastError());
        IX_TRY(pcp_readback(ctx, hb, bins.p, 32 * 8));
        IX_CUDA(cudaStreamSynchronize(ctx->stream));
        unsigned long long acc = 0;
        int mb = 0;
        for (int b = 0; b < 32; b++) {
            acc += hb[b];
            if (2 * acc >= nv) {
                mb = b;
                break;
            }
        }
        double occ = 1.5 * (double)(1u << mb); // centre of [2^b, 2^(b+1))
        if (nv <= 64) break;                    // tiny clouds: any grid works
        if (mb == 0) {
            h *= 2.5;
        } else if (occ < 10.0 || occ >= 40.0) {
            h *= pow(16.0 / occ, 1.0 / 2.5);
        } else {
            break;
        }
    }
    if (!(cell_size > 0) && nv > 64) ctx->hint_h = ix->grid.h, ctx->hint_areal = areal;
    void* q = nullptr;
    IX_TRY(pcp_arena_alloc(ctx, nv * sizeof(pcp_rec), &q));
    ix->rec = (pcp_rec*)q;
    IX_TRY(pcp_arena_alloc(ctx, nv * sizeof(float4), &q));
    ix->fpt = (float4*)q;
    {
        // |float - true| <= half an ulp of the largest origin-relative coordinate, plus the rounding of the
        // double subtraction p - origin
        double emax = 0, cmax = 0;
        for (int d = 0; d < 3; d++) {
            emax = fmax(emax, bb.mx[d] - bb.mn[d]);
            cmax = fmax(cmax, fmax(fabs(bb.mx[d]), fabs(bb.mn[d])));
        }
        ix->eps0 = emax * 5.97e-8 + cmax * 2.3e-16;
        ix->epsd = cmax * 2.3e-16;
        StageTimer t(ctx, "index.gather", 72.0 * nv);
        gather_rec_kernel<<<pcp_blocks_for(nv, 256), 256, 0, ctx->stream>>>(static_cast<const char*>(dpts), stride,
                                                                          vals, nv, ix->rec, ix->fpt, ix->grid.org[0],
                                                                          ix->grid.org[1], ix->grid.org[2],
                                                                          (stride == 32 && ((uintptr_t)dpts & 31) == 0) ? 1 : 0);
        ctx->launches++;
        IX_CUDA(cudaGetLastError());
    }
    {
        StageTimer t(ctx, "index.hash");
        uint32_t cap = 64;
        while (cap < 2u * ncell && cap < 0x80000000u) cap <<= 1;
        ix->hmask = cap - 1;
        IX_TRY(pcp_arena_alloc(ctx, (size_t)cap * sizeof(pcp_hslot), &q));
        ix->htab = (pcp_hslot*)q;
        IX_CUDA(cudaMemsetAsync(ix->htab, 0xff, (size_t)cap * sizeof(pcp_hslot), ctx->stream));
        hash_insert_kernel<<<pcp_blocks_for(ncell, 256), 256, 0, ctx->stream>>>(ix->ckey, ix->cstart, ncell, ix->htab,
                                                                              ix->hmask);
        ctx->launches++;
        IX_CUDA(cudaGetLastError());
    }
    {
        StageTimer t(ctx, "index.levels");
        DevBuf<unsigned long long> tot;
        IX_TRY(tot.alloc(ctx, 1));
        IX_CUDA(cudaMemsetAsync(tot.p, 0, 8, ctx->stream));
        lhash_count_kernel<<<pcp_blocks_for(ncell, 256), 256, 0, ctx->stream>>>(ix->ckey, ncell, tot.p);
        ctx->launches++;
        IX_CUDA(cudaGetLastError());
        unsigned long long ne = 0;
        IX_TRY(pcp_readback(ctx, &ne, tot.p, 8));
        IX_CUDA(cudaStreamSynchronize(ctx->stream));
        uint32_t cap = 64;
        while ((unsigned long long)cap < 2ull * ne && cap < 0x80000000u) cap <<= 1;
        ix->lmask = cap - 1;
        IX_TRY(pcp_arena_alloc(ctx, (size_t)cap * sizeof(pcp_hslot), &q));
        ix->ltab = (pcp_hslot*)q;
        IX_CUDA(cudaMemsetAsync(ix->ltab, 0xff, (size_t)cap * sizeof(pcp_hslot), ctx->stream));
        lhash_heads_kernel<<<pcp_blocks_for(ncell, 256), 256, 0, ctx->stream>>>(ix->ckey, ncell, ix->ltab, ix->lmask);
        ctx->launches++;
        IX_CUDA(cudaGetLastError());
        lhash_tails_kernel<<<pcp_blocks_for(ncell, 256), 256, 0, ctx->stream>>>(ix->ckey, ncell, ix->ltab, ix->lmask);
        ctx->launches++;
        IX_CUDA(cudaGetLastError());
        IX_TRY(pcp_arena_alloc(ctx, (size_t)ncell * sizeof(uint4), &q));
        ix->cinfo = (uint4*)q;
        cinfo_kernel<<<pcp_blocks_for(ncell, 256), 256, 0, ctx->stream>>>(ix->ckey, ix->cstart, ncell, ix->cinfo);
        ctx->launches++;
        IX_CUDA(cudaGetLastError());
    }
    IX_TRY(pcp_index_build_items(ctx, ix, pcp_item_cmax(20)));
#undef IX_TRY
#undef IX_CUDA
    *out = ix;
    return PCP_OK;
}

extern "C" {

int pcp_index_build(pcp_ctx* ctx, const void* pts, size_t n, size_t stride, double cell_size, pcp_index** out)
{
    if (!ctx || !out || stride < 24 || (n && !pts)) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_index_build: bad argument");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    InStage in;
    PCP_TRY(in.open(ctx, pts, n * stride));
    return pcp_index_build_device(ctx, in.dptr, n, stride, cell_size, out);
}

void pcp_index_free(pcp_index* index)
{
    if (!index) return;
    std::lock_guard<std::recursive_mutex> lock(index->ctx->mu);
    cudaSetDevice(index->ctx->device);
    index_release(index);
}

int64_t pcp_index_size(const pcp_index* index) { return index ? (int64_t)index->n : 0; }
double pcp_index_cell_size(const pcp_index* index) { return index ? index->grid.h : 0.0; }

} // extern "C"
