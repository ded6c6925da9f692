// knn_cell.cuh -- the cell-cooperative exact kNN + eigen-feature kernel (the hot kernel).
//
// One warp owns up to 32 query points of one base cell (contiguous in the Morton order); each lane is one
// query.  All lanes share the candidate set -- the 3x3x3 block of base cells around the query cell -- so the
// candidate stream is read with warp-uniform 16-byte loads (one L1 wavefront serves 32 queries) and the
// loop has no divergence.  Selection is a two-pass histogram select, not a per-thread heap:
//   pass 1  every lane bins the squared distance of every candidate into NB equal-width buckets kept in
//           shared memory (hist[bucket][lane], one bank per lane: conflict-free);
//   scan    the lane finds the first bucket b* at which its cumulative count reaches k;
//   pass 2  candidates with bucket <= b* (k + a few) are written to the lane's shared-memory list at
//           offsets taken from the bucket prefix sums -- the list comes out bucket-sorted;
//   sort    an insertion sort over the nearly sorted list orders it by (d2, original index).
// A query is exact iff bucket edge(b*+1) <= (distance to the nearest face of the scanned block)^2; queries
// that fail that test (or overflow the list) go to the fall-back list for the warp-per-query kernel.
// The winning k neighbours feed calculate_plan_parameter_h_points (plane_fit.cuh) in the same kernel,
// so neighbour lists never travel to HBM unless the caller asks for them.
#pragma once
#include "index.cuh"
#include "plane_fit.cuh"

#define PCP_KNN_NB 32         // distance buckets
#define PCP_KNN_WARPS 4       // warps per block

struct pcp_knn_params {
    const pcp_rec* rec;
    const pcp_hslot* htab;
    uint32_t hmask;
    const unsigned long long* ckey;
    const uint32_t* cstart;
    uint32_t ncell;
    const uint32_t* item_cell;
    const uint32_t* item_first;
    uint32_t nitem;
    pcp_grid g;
    int k;   // requested
    int kk;  // min(k, valid points)
    int cap; // list capacity per query (>= kk)
    void* props;
    double* eig3;
    float* lps3;
    int32_t* knn_idx;
    uint32_t* fail_list; // sorted position | reason<<31
    uint32_t* fail_count;
};

__device__ __forceinline__ pcp_rec pcp_load_rec(const pcp_rec* __restrict__ rec, uint32_t i)
{
    // one 256-bit load (LDG.E.256 on sm_100: half the load instructions and L1 wavefronts of two 128-bit ones; the
    // records are 32 bytes, 32-byte aligned, and read-only for as long as the index lives)
    pcp_rec r;
    double w;
    asm("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(w) : "l"(rec + i));
    r.idx = __double_as_longlong(w);
    return r;
}

__device__ __forceinline__ double pcp_dist2(double qx, double qy, double qz, double x, double y, double z)
{
    // FLANN L2_Simple: diff = a - b; result += diff*diff, in x,y,z order (no FMA: built with -fmad=false)
    double dx = qx - x, dy = qy - y, dz = qz - z;
    double r = dx * dx;
    r = r + dy * dy;
    r = r + dz * dz;
    return r;
}

// squared guaranteed radius of the 3x3x3 block of level-`level` cells around cell (cx,cy,cz) for query q
__device__ __forceinline__ double pcp_guarantee2(const pcp_grid& g, int level, int cx, int cy, int cz, double qx,
                                                 double qy, double qz)
{
    const double H = ldexp(g.h, level);
    double dout = DBL_MAX;
    const double q[3] = {qx, qy, qz};
    const int c[3] = {cx, cy, cz};
#pragma unroll
    for (int a = 0; a < 3; a++) {
        int dim_l = ((g.dims[a] - 1) >> level) + 1;
        if (c[a] - 1 >= 0) dout = fmin(dout, q[a] - (g.org[a] + (double)(c[a] - 1) * H));
        if (c[a] + 1 <= dim_l - 1) dout = fmin(dout, (g.org[a] + (double)(c[a] + 2) * H) - q[a]);
    }
    if (dout == DBL_MAX) return DBL_MAX; // the block covers the whole grid
    double slack = 1e-9 * H + 8e-16 * fmax(fabs(qx), fmax(fabs(qy), fabs(qz)));
    double safe = dout - slack;
    return safe > 0.0 ? safe * safe : 0.0;
}

// neighbour order: centre cell first, then by growing offset, so near candidates come early
__device__ __forceinline__ void pcp_nb_offset(int lane, int& dx, int& dy, int& dz)
{
    // lane 0 -> (0,0,0); others enumerate the remaining 26 offsets
    int t = lane == 0 ? 13 : (lane <= 13 ? lane - 1 : lane);
    dx = t % 3 - 1;
    dy = (t / 3) % 3 - 1;
    dz = t / 9 - 1;
}

static __global__ void __launch_bounds__(32 * PCP_KNN_WARPS) knn_cell_kernel(pcp_knn_params P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const size_t per_warp = (size_t)PCP_KNN_NB * 32 * 4 + (size_t)P.cap * 32 * 12;
    unsigned char* base = smem_raw + per_warp * warp;
    double* list_d2 = reinterpret_cast<double*>(base);                                   // [cap][32]
    uint32_t* list_pos = reinterpret_cast<uint32_t*>(base + (size_t)P.cap * 32 * 8);     // [cap][32]
    uint32_t* hist = reinterpret_cast<uint32_t*>(base + (size_t)P.cap * 32 * 12);        // [NB][32]

    const uint32_t item = blockIdx.x * PCP_KNN_WARPS + warp;
    if (item >= P.nitem) return;
    const uint32_t cell = __ldg(P.item_cell + item);
    const uint32_t first = __ldg(P.item_first + item);
    const uint32_t cend = __ldg(P.cstart + cell + 1);
    const int nq = min(32u, cend - first);
    const bool active = lane < nq;
    const uint32_t qpos = first + (active ? lane : 0);
    const pcp_rec q = pcp_load_rec(P.rec, qpos);

    int cx, cy, cz;
    pcp_demorton(__ldg(P.ckey + cell), cx, cy, cz);

    // 27 neighbour cells: lanes 0..26 look one up each
    uint32_t nb_start = 0, nb_count = 0;
    if (lane < 27) {
        int dx, dy, dz;
        pcp_nb_offset(lane, dx, dy, dz);
        int nx = cx + dx, ny = cy + dy, nz = cz + dz;
        if (nx >= 0 && ny >= 0 && nz >= 0 && nx < P.g.dims[0] && ny < P.g.dims[1] && nz < P.g.dims[2])
            if (!pcp_hash_find(P.htab, P.hmask, pcp_morton(nx, ny, nz), nb_start, nb_count)) nb_count = 0;
    }
    uint32_t total = nb_count;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);

    // bucket width: buckets span [0, range); range follows the local density (k-th distance of a planar
    // patch with `total` points in the 3x3 block is ~ 9 k H^2 / (pi total)), capped by the largest
    // possible guarantee (1.5 H)^2
    const double H = P.g.h;
    double range = 2.25 * H * H;
    double est = 4.0 * (9.0 * (double)P.kk * H * H) / (3.141592653589793 * (double)total);
    if (est < range) range = est;
    const double inv_bw = (double)PCP_KNN_NB / range;
    const double bw = range / (double)PCP_KNN_NB;
    const double g2 = pcp_guarantee2(P.g, 0, cx, cy, cz, q.x, q.y, q.z);

#pragma unroll
    for (int b = 0; b < PCP_KNN_NB; b++) hist[b * 32 + lane] = 0;
    __syncwarp();

    // ---- pass 1: histogram of squared distances ----
    for (int nb = 0; nb < 27; nb++) {
        const uint32_t s = __shfl_sync(0xffffffffu, nb_start, nb);
        const uint32_t c = __shfl_sync(0xffffffffu, nb_count, nb);
#pragma unroll 4
        for (uint32_t t = 0; t < c; t++) {
            const pcp_rec r = pcp_load_rec(P.rec, s + t);
            const double d2 = pcp_dist2(q.x, q.y, q.z, r.x, r.y, r.z);
            const double sb = d2 * inv_bw;
            const int b = sb < (double)(PCP_KNN_NB - 1) ? (int)sb : (PCP_KNN_NB - 1);
            hist[b * 32 + lane] += 1;
        }
    }
    __syncwarp();

    // ---- scan: bucket offsets, b* ----
    int bstar = -1;
    uint32_t cum = 0, m = 0;
#pragma unroll
    for (int b = 0; b < PCP_KNN_NB; b++) {
        uint32_t h = hist[b * 32 + lane];
        hist[b * 32 + lane] = cum; // becomes the write cursor of bucket b
        cum += h;
        if (bstar < 0 && b < PCP_KNN_NB - 1 && cum >= (uint32_t)P.kk) bstar = b, m = cum;
    }
    int reason = 0;
    if (bstar < 0) reason = 1;                                                   // k-th neighbour beyond the buckets
    else if ((double)(bstar + 1) * bw * (1.0 + 1e-12) > g2) reason = 1;          // not provably inside the block
    else if (m > (uint32_t)P.cap) reason = 2;                                    // list overflow (ties / dense bucket)
    const bool ok = active && reason == 0;
    __syncwarp();

    // ---- pass 2: collect candidates of buckets <= b* ----
    const int blim = ok ? bstar : -1;
    for (int nb = 0; nb < 27; nb++) {
        const uint32_t s = __shfl_sync(0xffffffffu, nb_start, nb);
        const uint32_t c = __shfl_sync(0xffffffffu, nb_count, nb);
#pragma unroll 4
        for (uint32_t t = 0; t < c; t++) {
            const pcp_rec r = pcp_load_rec(P.rec, s + t);
            const double d2 = pcp_dist2(q.x, q.y, q.z, r.x, r.y, r.z);
            const double sb = d2 * inv_bw;
            const int b = sb < (double)(PCP_KNN_NB - 1) ? (int)sb : (PCP_KNN_NB - 1);
            if (b <= blim) {
                uint32_t slot = hist[b * 32 + lane];
                hist[b * 32 + lane] = slot + 1;
                list_d2[slot * 32 + lane] = d2;
                list_pos[slot * 32 + lane] = s + t;
            }
        }
    }
    __syncwarp();

    // fall-back queue (warp-aggregated append)
    {
        const bool failq = active && reason != 0;
        unsigned fm = __ballot_sync(0xffffffffu, failq);
        if (fm) {
            uint32_t basei = 0;
            if (lane == __ffs(fm) - 1) basei = atomicAdd(P.fail_count, (uint32_t)__popc(fm));
            basei = __shfl_sync(0xffffffffu, basei, __ffs(fm) - 1);
            if (failq) P.fail_list[basei + __popc(fm & ((1u << lane) - 1))] = qpos | (reason == 1 ? 0x80000000u : 0u);
        }
    }
    if (!ok) return;

    // ---- order the list by (d2, original index): insertion sort over a bucket-sorted list ----
    for (uint32_t i = 1; i < m; i++) {
        const double d = list_d2[i * 32 + lane];
        const uint32_t p = list_pos[i * 32 + lane];
        long long pidx = -1;
        int j = (int)i - 1;
        while (j >= 0) {
            const double dj = list_d2[j * 32 + lane];
            bool gt = dj > d;
            if (dj == d) { // exact tie: order by original index
                if (pidx < 0) pidx = P.rec[p].idx;
                gt = P.rec[list_pos[j * 32 + lane]].idx > pidx;
            }
            if (!gt) break;
            list_d2[(j + 1) * 32 + lane] = dj;
            list_pos[(j + 1) * 32 + lane] = list_pos[j * 32 + lane];
            j--;
        }
        list_d2[(j + 1) * 32 + lane] = d;
        list_pos[(j + 1) * 32 + lane] = p;
    }

    // ---- features on the first kk entries ----
    const size_t jq = (size_t)q.idx;
    const int N = P.kk;
    if (P.knn_idx) {
        for (int e = 0; e < P.k; e++)
            P.knn_idx[jq * P.k + e] = e < N ? (int32_t)P.rec[list_pos[e * 32 + lane]].idx : -1;
    }
    pcp_plane pl;
    const bool fitted = N > 3; // calculate_feature.cpp:238
    if (fitted) {
        double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
        for (int e = 0; e < N; e++) {
            const pcp_rec r = pcp_load_rec(P.rec, list_pos[e * 32 + lane]);
            XAve += r.x;
            YAve += r.y;
            ZAve += r.z;
        }
        XAve /= (double)N;
        YAve /= (double)N;
        ZAve /= (double)N;
        double s00 = 0, s01 = 0, s02 = 0, s11 = 0, s12 = 0, s22 = 0;
        for (int e = 0; e < N; e++) {
            const pcp_rec r = pcp_load_rec(P.rec, list_pos[e * 32 + lane]);
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
}
