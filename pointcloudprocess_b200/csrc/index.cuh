// index.cuh -- the search structure that replaces the FLANN kd-tree behind KdTreeFLANN::setInputCloud
// (kd_tree.h:772-798): points sorted along a Morton curve over a uniform grid of base cell h, plus
//   * `rec`   : the sorted points as 32-byte records {x, y, z, original index}  (two 16-byte loads)
//   * `ckey`  : the sorted distinct base-cell Morton codes, `cstart` their first point (ncell+1)
//   * `htab`  : open-addressing hash  Morton code -> (first point, count)  for the base level
// A cell of level l (edge h*2^l) is the contiguous run of base cells sharing the code prefix code>>3l,
// so one sort serves every level (an implicit linear octree / LBVH); coarse levels are located by binary
// search in `ckey`.
#pragma once
#include "common.cuh"

struct __align__(16) pcp_rec {
    double x, y, z;
    long long idx; // original cloud index
};

struct __align__(16) pcp_hslot {
    unsigned long long key; // Morton code, ~0 = empty
    uint32_t start;
    uint32_t count;
};

struct pcp_grid {
    double org[3];   // grid origin (bounding-box minimum)
    double h;        // base cell edge
    double inv_h;
    int dims[3];     // base cells per axis
    int nbits;       // bits per axis in the Morton code
    int top_level;   // first level at which the whole grid is one cell
};

struct pcp_index {
    pcp_ctx* ctx = nullptr;
    size_t n_orig = 0;  // points handed to pcp_index_build
    size_t n = 0;       // valid (finite) points == total_nr_points_
    pcp_grid grid;
    pcp_rec* rec = nullptr;
    float4* fpt = nullptr;  // (float)(p - grid origin) of the sorted points, for the FP32 candidate filter
    double eps0 = 0;        // bound on the error of those floats (see knn_fast.cuh)
    double epsd = 0;        // bound on the rounding error of a double subtraction p - origin
    unsigned long long* ckey = nullptr;
    uint32_t* cstart = nullptr;
    uint32_t ncell = 0;
    pcp_hslot* htab = nullptr;
    uint32_t hmask = 0;
    // cells of levels 0..PCP_HLEVELS-1: key (level<<58 | Morton prefix) -> (first base cell, base cells)
    pcp_hslot* ltab = nullptr;
    uint32_t lmask = 0;
    uint4* cinfo = nullptr; // per base cell: integer coordinates x, y, z and the first point
    // per-cell work items for the cell-cooperative kernels: item -> (cell, chunk of 32 queries)
    uint32_t* item_cell = nullptr;
    uint32_t* item_first = nullptr; // first sorted point of the chunk
    uint32_t* item_meta = nullptr;  // level << 8 | queries (index.cu: octree leaves of capacity item_cmax)
    uint32_t nitem = 0;
    uint32_t item_cmax = 0;
};

#define PCP_EMPTY_KEY 0xffffffffffffffffull
#define PCP_HLEVELS 5
#define PCP_ITEM_LMAX 3           // coarsest level of a dense-kernel work item
// leaf capacity for k neighbours: a planar patch of cmax/4..cmax points per leaf keeps the k-th neighbour inside
// the 3x3x3 block of leaves for most queries while the block stays within the staging capacity.  Swept on the 25 M-point
// city shard (tools/sweep_items.sh; whole step): 1.15 k = 23 -> 39.2 ms, 26 -> 39.1, 29 -> 38.8, 32 -> 39.1: larger leaves
// leave fewer queries to the far kernel (1.36 M -> 1.18 M) and more blocks to the second dense pass; 1.45 k it is.
static inline uint32_t pcp_item_cmax(int k)
{
    const int c = (int)(1.45 * k + 0.5);
    return (uint32_t)(c < 8 ? 8 : (c > 32 ? 32 : c));
}
#define PCP_LKEY(level, prefix) (((unsigned long long)(level) << 58) | (prefix))

__host__ __device__ __forceinline__ unsigned long long pcp_spread3(unsigned long long v)
{
    v &= 0x1fffffull;
    v = (v | (v << 32)) & 0x1f00000000ffffull;
    v = (v | (v << 16)) & 0x1f0000ff0000ffull;
    v = (v | (v << 8)) & 0x100f00f00f00f00full;
    v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}
__host__ __device__ __forceinline__ unsigned long long pcp_compact3(unsigned long long v)
{
    v &= 0x1249249249249249ull;
    v = (v | (v >> 2)) & 0x10c30c30c30c30c3ull;
    v = (v | (v >> 4)) & 0x100f00f00f00f00full;
    v = (v | (v >> 8)) & 0x1f0000ff0000ffull;
    v = (v | (v >> 16)) & 0x1f00000000ffffull;
    v = (v | (v >> 32)) & 0x1fffffull;
    return v;
}
__host__ __device__ __forceinline__ unsigned long long pcp_morton(int cx, int cy, int cz)
{
    return pcp_spread3((unsigned)cx) | (pcp_spread3((unsigned)cy) << 1) | (pcp_spread3((unsigned)cz) << 2);
}
__host__ __device__ __forceinline__ void pcp_demorton(unsigned long long m, int& cx, int& cy, int& cz)
{
    cx = (int)pcp_compact3(m);
    cy = (int)pcp_compact3(m >> 1);
    cz = (int)pcp_compact3(m >> 2);
}
__host__ __device__ __forceinline__ uint32_t pcp_hash64(unsigned long long k)
{
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdull;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ull;
    k ^= k >> 33;
    return (uint32_t)k;
}

// base-cell coordinate of a coordinate value (same expression at build and query time)
__device__ __forceinline__ int pcp_cell_coord(double v, double org, double inv_h, int dim)
{
    double t = __dmul_rn(__dsub_rn(v, org), inv_h);
    int c = (t >= 0.0) ? (int)t : -1;
    return c;
}

#ifdef __CUDACC__
__device__ __forceinline__ bool pcp_hash_find(const pcp_hslot* __restrict__ htab, uint32_t hmask,
                                              unsigned long long key, uint32_t& start, uint32_t& count)
{
    uint32_t s = pcp_hash64(key) & hmask;
    while (true) {
        // one 16-byte load per probe
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>(htab + s));
        unsigned long long k = ((unsigned long long)raw.y << 32) | raw.x;
        if (k == key) {
            start = raw.z;
            count = raw.w;
            return true;
        }
        if (k == PCP_EMPTY_KEY) return false;
        s = (s + 1) & hmask;
    }
}

// points [lo, hi) of the level-l cell with coordinates (cx,cy,cz) at that level (binary search in ckey)
__device__ __forceinline__ void pcp_level_range(const unsigned long long* __restrict__ ckey,
                                                const uint32_t* __restrict__ cstart, uint32_t ncell, int level,
                                                int cx, int cy, int cz, uint32_t& lo, uint32_t& hi)
{
    unsigned long long pre = pcp_morton(cx, cy, cz);
    int sh = 3 * level;
    unsigned long long klo = sh >= 64 ? 0ull : (pre << sh);
    unsigned long long khi = sh >= 64 ? PCP_EMPTY_KEY : ((pre + 1) << sh);
    uint32_t a = 0, b = ncell;
    while (a < b) { // first cell with key >= klo
        uint32_t m = (a + b) >> 1;
        if (__ldg(ckey + m) < klo) a = m + 1; else b = m;
    }
    uint32_t first = a;
    b = ncell;
    while (a < b) { // first cell with key >= khi
        uint32_t m = (a + b) >> 1;
        if (__ldg(ckey + m) < khi) a = m + 1; else b = m;
    }
    lo = __ldg(cstart + first);
    hi = __ldg(cstart + a);
}
#endif

int pcp_index_build_items(pcp_ctx* ctx, pcp_index* ix, uint32_t cmax);
int pcp_index_build_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, double cell_size, pcp_index** out);
