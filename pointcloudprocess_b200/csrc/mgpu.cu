// mgpu.cu -- the whole pipeline over several GPUs of ONE process, behind the C ABI (include/pcp_cuda.h: pcp_mgpu_*):
// a caller shaped like the reference's main.cpp (one process, one cloud in host memory) shards without torch, NCCL or
// a launcher.  One host thread per device drives that device's context through the library's own entry points; the
// exchanges are peer-to-peer copies over NVLink (cudaMemcpyPeerAsync) between barriers of the host threads.  The
// choreography is the one of pointcloudprocess_b200/sharded.py (the torch.distributed driver bench.py runs):
//   plan          per-range box / count / tree sums, the serial-order centroid of the whole cloud as a chain over the
//                 ranges (pcp_serial_sum_begin / _end), the grid box of the centred cloud, a histogram of 1 m voxel
//                 columns along the longer horizontal axis, slabs of equal count at least a halo wide;
//   redistribute  stable partition by slab (pcp_partition_columns), every device pulls its slab's pieces;
//   shard         remove_duplicate in the whole cloud's frame, halo pull, index + kNN + plane fit, boundary queries
//                 answered by the neighbour (pcp_knn) and merged (pcp_features_merge).
// Results are those of the single-device pipeline, point for point (tests/test_gpu_mgpu.py); their ORDER is slab by
// slab along the cut axis, each slab in voxel order.
#include "common.cuh"
#include <cfloat>
#include <atomic>
#include <algorithm>
#include <climits>
#include <cmath>
#include <condition_variable>
#include <thread>

struct pcp_mgpu {
    std::vector<pcp_ctx*> ctx;
    std::string err;
};

namespace {

struct Barrier {
    std::mutex mu;
    std::condition_variable cv;
    int n, waiting = 0;
    unsigned long long gen = 0;
    explicit Barrier(int n_) : n(n_) {}
    void wait()
    {
        std::unique_lock<std::mutex> l(mu);
        const unsigned long long g = gen;
        if (++waiting == n) {
            waiting = 0, gen++;
            cv.notify_all();
        } else {
            cv.wait(l, [&] { return gen != g; });
        }
    }
};

// rows of row_words 32-bit words: dst[i] = src[idx[i]] (idx < 0: the fill word), and dst[idx[i]] = src[i]
__global__ void __launch_bounds__(256) rows_gather_kernel(const uint32_t* __restrict__ src, int row_words, const int32_t* __restrict__ idx,
                                                          size_t n, uint32_t* __restrict__ dst, uint32_t fill)
{
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= n * (size_t)row_words) return;
    const size_t i = t / row_words, w = t % row_words;
    const int32_t s = idx[i];
    dst[t] = s < 0 ? fill : src[(size_t)s * row_words + w];
}
__global__ void __launch_bounds__(256) rows_scatter_kernel(const uint32_t* __restrict__ src, int row_words, const int32_t* __restrict__ idx,
                                                           size_t n, uint32_t* __restrict__ dst)
{
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= n * (size_t)row_words) return;
    const size_t i = t / row_words, w = t % row_words;
    dst[(size_t)idx[i] * row_words + w] = src[t];
}
// out[i] = x, y, z of record idx[i] (zeros for idx < 0: a padded list entry, its distance is +inf)
__global__ void __launch_bounds__(256) xyz_gather_kernel(const char* __restrict__ pts, size_t stride, const int32_t* __restrict__ idx,
                                                         size_t n, double* __restrict__ out)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t s = idx[i];
    double x = 0, y = 0, z = 0;
    if (s >= 0) {
        const double* p = reinterpret_cast<const double*>(pts + (size_t)s * stride);
        x = p[0], y = p[1], z = p[2];
    }
    out[3 * i] = x, out[3 * i + 1] = y, out[3 * i + 2] = z;
}
__global__ void __launch_bounds__(256) fill_f64_kernel(double* __restrict__ p, size_t n, double v)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
__global__ void __launch_bounds__(256) pointid_offset_kernel(char* __restrict__ props, size_t n, int32_t off)
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) *reinterpret_cast<int32_t*>(props + i * 48 + 32) += off;
}

// plain device memory of one device (cudaMalloc: these buffers are read by peers)
struct Dev {
    void* p = nullptr;
    size_t bytes = 0;
    int alloc(size_t nb)
    {
        release();
        bytes = nb ? nb : 16;
        return cudaMalloc(&p, bytes) == cudaSuccess ? PCP_OK : PCP_ERR_NOMEM;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr, bytes = 0;
    }
    ~Dev() { release(); }
    template <typename T> T* as() const { return static_cast<T*>(p); }
};

constexpr double kDblMin = 2.2250738585072014e-308;

// slab boundaries (voxel columns): equal counts on histogram-bin boundaries, every slab at least `halo` wide
// (sharded.py: cut_slabs)
int cut_slabs(const std::vector<unsigned long long>& hist, int world, int cpb, int div_ax, double leaf_d, double halo,
              std::vector<int>& bounds, std::string& err)
{
    const int nbins = (int)hist.size();
    int min_bins = (int)std::ceil(halo / (cpb * leaf_d) - 1e-9);
    if (min_bins < 1) min_bins = 1;
    std::vector<long long> cum(nbins);
    long long acc = 0;
    for (int i = 0; i < nbins; i++) cum[i] = (acc += (long long)hist[i]);
    const long long total = acc;
    const int last_bin = div_ax / cpb + 1;
    if (world > 1 && world * min_bins > last_bin + 1) {
        err = "cloud too narrow along the cut axis for this many slabs at least a halo wide";
        return PCP_ERR_RANGE;
    }
    std::vector<int> b(1, 0);
    for (int r = 1; r < world; r++) {
        const long long target = (total * r) / world;
        int want = (int)(std::lower_bound(cum.begin(), cum.end(), target) - cum.begin()) + 1;
        const int lo = b.back() + min_bins, hi = last_bin + 1 - (world - r) * min_bins;
        b.push_back(std::min(std::max(want, lo), std::max(hi, lo)));
    }
    b.push_back(nbins);
    bounds.resize(world + 1);
    for (int r = 0; r <= world; r++) bounds[r] = b[r] * cpb;
    return PCP_OK;
}

struct Shared {
    int world = 0;
    const char* pts = nullptr;
    size_t n = 0, stride = 0;
    float leaf = 0;
    int k = 0;
    double halo = 0;
    char* out_pts = nullptr;
    char* out_props = nullptr;
    // per rank
    std::vector<double> stats;              // [world][10]
    double carry[3] = {0, 0, 0};
    int turn = 0;
    std::mutex turn_mu;
    std::condition_variable turn_cv;
    std::vector<std::vector<unsigned long long>> hist;
    std::vector<std::vector<uint64_t>> counts; // [src][dst] records
    std::vector<void*> part;                   // partition buffers (device pointers, per source)
    std::vector<size_t> m;                     // voxelised own points per rank
    std::vector<void*> halo_stage;             // staged halo records per rank: [to left | to right]
    std::vector<size_t> halo_l, halo_r;        // counts
    std::vector<void*> bq[2];                  // boundary queries per rank and side (xyz doubles)
    std::vector<size_t> bqn[2];
    std::vector<void*> ans_xyz[2], ans_d2[2];  // answers per rank (the answering one) and side (the side it answered FROM)
    std::atomic<int> failed{0};
    std::mutex err_mu;
    std::string err;
    int status = PCP_OK;
    void fail(int code, const std::string& what)
    {
        std::lock_guard<std::mutex> l(err_mu);
        if (!failed.load()) status = code, err = what;
        failed.store(1);
    }
};

} // namespace

namespace {

#define MG(expr)                                                                                                      \
    do {                                                                                                              \
        if (!S.failed.load()) {                                                                                       \
            const int _s = (expr);                                                                                    \
            if (_s != PCP_OK) S.fail(_s, std::string(#expr) + ": " + pcp_last_error(ctx));                            \
        }                                                                                                             \
    } while (0)
#define MGC(expr)                                                                                                     \
    do {                                                                                                              \
        if (!S.failed.load()) {                                                                                       \
            const cudaError_t _e = (expr);                                                                            \
            if (_e != cudaSuccess) S.fail(PCP_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));          \
        }                                                                                                             \
    } while (0)

void worker(pcp_mgpu* mg, Shared& S, Barrier& B, int r)
{
    pcp_ctx* ctx = mg->ctx[r];
    const int world = S.world, k = S.k;
    const int dev = ctx->device;
    cudaSetDevice(dev);
    cudaStream_t st = ctx->stream;
    const size_t first = (S.n * (size_t)r) / world, nr = (S.n * (size_t)(r + 1)) / world - first;
    const bool has_l = r > 0, has_r = r < world - 1;
    const double H = S.halo;
    auto blocks = [](size_t n) { return (unsigned)((n + 255) / 256); };
    auto sync = [&]() { MGC(cudaStreamSynchronize(st)); };
    auto peer_pull = [&](void* dst, const void* src, int src_rank, size_t bytes) {
        if (bytes) MGC(cudaMemcpyPeerAsync(dst, dev, src, mg->ctx[src_rank]->device, bytes, st));
    };

    // ---- A. this device's range of the cloud, as packed records; its box, count and tree sums ----
    Dev src, raw;
    MG(src.alloc(nr * 32));
    if (S.stride == 32) {
        if (nr) MGC(cudaMemcpyAsync(src.p, S.pts + first * 32, nr * 32, cudaMemcpyHostToDevice, st));
    } else {
        MG(raw.alloc(nr * 48));
        if (nr) MGC(cudaMemcpyAsync(raw.p, S.pts + first * 48, nr * 48, cudaMemcpyHostToDevice, st));
        MG(pcp_cloud_pack(ctx, raw.p, nr, src.p));
        sync();
        raw.release();
    }
    MG(pcp_cloud_stats(ctx, src.p, nr, 32, &S.stats[(size_t)r * 10]));
    B.wait();
    // ---- B. the serial-order centroid of the whole cloud: prepared in parallel, the running sum walks the ranges ----
    double pred[3] = {0, 0, 0};
    for (int q = 0; q < r; q++)
        for (int a = 0; a < 3; a++) pred[a] += S.stats[(size_t)q * 10 + 6 + a];
    void* hsum = nullptr;
    MG(pcp_serial_sum_begin(ctx, src.p, nr, 32, pred, &hsum));
    {
        std::unique_lock<std::mutex> l(S.turn_mu);
        S.turn_cv.wait(l, [&] { return S.turn == r; });
        if (hsum) {
            const int s2 = pcp_serial_sum_end(hsum, S.failed.load() ? nullptr : S.carry);
            if (s2 != PCP_OK) S.fail(s2, std::string("pcp_serial_sum_end: ") + pcp_last_error(ctx));
        }
        S.turn++;
        S.turn_cv.notify_all();
    }
    B.wait();
    // ---- C. the plan: every thread computes the same from the same reduced data ----
    double count = 0, mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX}, centre[3];
    for (int q = 0; q < world; q++) {
        const double* s = &S.stats[(size_t)q * 10];
        if (s[9] <= 0) continue;
        count += s[9];
        for (int a = 0; a < 3; a++) mn[a] = std::min(mn[a], s[a]), mx[a] = std::max(mx[a], s[3 + a]);
    }
    if (count <= 0) S.fail(PCP_ERR_INVALID, "pcp_mgpu: the cloud holds no finite point");
    const float leaf_f = S.leaf > 0 ? S.leaf : 0.02f; // without a voxel grid the columns only serve the cut
    const double leaf_d = (double)leaf_f, inv = 1.0 / leaf_d;
    int32_t min_b[3] = {0, 0, 0}, div_b[3] = {1, 1, 1};
    for (int a = 0; a < 3; a++) {
        centre[a] = count > 0 ? S.carry[a] / count : 0.0; // compute3DCentroid: sum / count
        const double lo = mn[a] + (-centre[a]), hi = std::max(mx[a] + (-centre[a]), kDblMin);
        min_b[a] = (int32_t)(lo * inv); // truncation, voxel_grid.h:835-840
        div_b[a] = (int32_t)(hi * inv) - min_b[a] + 1;
    }
    const int ax = (mx[0] - mn[0]) >= (mx[1] - mn[1]) ? 0 : 1;
    const int cpb = std::max(1, (int)std::lround(1.0 / leaf_d));
    const int nbins = div_b[ax] / cpb + 2;
    S.hist[r].assign((size_t)std::max(nbins, 1), 0ull);
    MG(pcp_column_histogram(ctx, src.p, nr, 32, ax, leaf_f, centre[ax], min_b[ax], 0, cpb, nbins, reinterpret_cast<uint64_t*>(S.hist[r].data())));
    B.wait();
    std::vector<int> bounds(world + 1, 0);
    if (!S.failed.load()) {
        std::vector<unsigned long long> hist((size_t)nbins, 0ull);
        for (int q = 0; q < world; q++)
            for (int b = 0; b < nbins; b++) hist[b] += S.hist[q][b];
        std::string e;
        const int s2 = cut_slabs(hist, world, cpb, div_b[ax], leaf_d, H, bounds, e);
        if (s2 != PCP_OK) S.fail(s2, e);
    }
    auto face = [&](int col) { return centre[ax] + (double)(col + min_b[ax]) * leaf_d; };
    // ---- D. redistribute: stable partition by slab, every device pulls its slab's pieces in range order ----
    Dev part;
    MG(part.alloc(nr * 32));
    S.counts[r].assign(world, 0);
    MG(pcp_partition_columns(ctx, src.p, nr, 32, ax, leaf_f, centre[ax], min_b[ax], bounds.data(), world, part.p, S.counts[r].data()));
    S.part[r] = part.p;
    sync();
    B.wait();
    size_t n_own = 0;
    for (int q = 0; q < world; q++) n_own += (size_t)S.counts[q][r];
    const size_t cap = n_own + n_own / 4 + ((size_t)1 << 20); // own + halos
    Dev cloud, work;
    MG(cloud.alloc(std::max(n_own, (size_t)1) * 32));
    MG(work.alloc(cap * 32));
    {
        size_t ro = 0;
        for (int q = 0; q < world; q++) {
            size_t so = 0;
            for (int d = 0; d < r; d++) so += (size_t)S.counts[q][d];
            const size_t c = (size_t)S.counts[q][r];
            if (!S.failed.load()) peer_pull(cloud.as<char>() + ro * 32, static_cast<char*>(S.part[q]) + so * 32, q, c * 32);
            ro += c;
        }
    }
    sync();
    B.wait();
    part.release();
    src.release();
    // ---- E. remove_duplicate in the whole cloud's frame (cells, centroids and order of the single-device run) ----
    size_t m = n_own;
    if (S.leaf > 0) {
        MG(pcp_remove_duplicate_to(ctx, cloud.p, n_own, 32, leaf_f, centre, min_b, div_b, work.p, &m));
    } else if (n_own) {
        MGC(cudaMemcpyAsync(work.p, cloud.p, n_own * 32, cudaMemcpyDeviceToDevice, st));
    }
    S.m[r] = m;
    // ---- F. halo: voxelised points within H of the slab faces, staged in `cloud` (free now), pulled by the neighbours ----
    const double xlo = face(bounds[r]), xhi = face(bounds[r + 1]);
    size_t cl = 0, cr = 0;
    if (has_l) MG(pcp_select_range(ctx, work.p, m, 32, ax, -INFINITY, xlo + H, cloud.p, &cl));
    if (has_r) MG(pcp_select_range(ctx, work.p, m, 32, ax, xhi - H, INFINITY, cloud.as<char>() + cl * 32, &cr));
    if (cl + cr > std::max(n_own, (size_t)1)) S.fail(PCP_ERR_RANGE, "pcp_mgpu: halo selection larger than its staging");
    S.halo_stage[r] = cloud.p, S.halo_l[r] = cl, S.halo_r[r] = cr;
    sync();
    B.wait();
    const size_t nl = has_l ? S.halo_r[r - 1] : 0, nrr = has_r ? S.halo_l[r + 1] : 0;
    const size_t ntot = m + nl + nrr;
    if (ntot > cap) S.fail(PCP_ERR_RANGE, "pcp_mgpu: shard + halos exceed the buffer");
    if (has_l && !S.failed.load())
        peer_pull(work.as<char>() + m * 32, static_cast<char*>(S.halo_stage[r - 1]) + S.halo_l[r - 1] * 32, r - 1, nl * 32);
    if (has_r && !S.failed.load()) peer_pull(work.as<char>() + (m + nl) * 32, S.halo_stage[r + 1], r + 1, nrr * 32);
    sync();
    B.wait();
    cloud.release();
    // ---- G. index + kNN + plane fit on own + halo points, features for the own points only ----
    pcp_index* ix = nullptr;
    Dev props, kth;
    MG(props.alloc(std::max(ntot, (size_t)1) * 48));
    MG(kth.alloc(std::max(ntot, (size_t)1) * 4));
    MG(pcp_index_build(ctx, work.p, ntot, 32, 0.0, &ix));
    if (ix) MG(pcp_features_knn_ex(ix, k, (int64_t)m, props.p, nullptr, nullptr, nullptr, kth.as<float>()));
    // ---- H. queries whose k-th neighbour may lie beyond the halo: answered by the neighbour, merged here ----
    size_t nF = 0, csub[2] = {0, 0};
    Dev F, qF, kthF, sub[2], sxyz[2], aq[2], aidx[2], ad2[2], axyz[2];
    if (world > 1) {
        size_t capb = (size_t)1 << 16;
        for (int attempt = 0; attempt < 8 && !S.failed.load(); attempt++) {
            MG(F.alloc(capb * 4));
            MG(qF.alloc(capb * 24));
            if (S.failed.load()) break;
            const int s2 = pcp_select_boundary(ctx, work.p, m, 32, ax, kth.as<float>(), has_l, xlo - H, has_r, xhi + H, capb,
                                               F.as<int32_t>(), qF.as<double>(), &nF);
            if (s2 == PCP_OK) break;
            if (s2 == PCP_ERR_RANGE && nF > capb) {
                capb = nF + nF / 4;
                continue;
            }
            S.fail(s2, std::string("pcp_select_boundary: ") + pcp_last_error(ctx));
        }
        if (nF && !S.failed.load()) {
            MG(kthF.alloc(nF * 4));
            rows_gather_kernel<<<blocks(nF), 256, 0, st>>>(kth.as<uint32_t>(), 1, F.as<int32_t>(), nF, kthF.as<uint32_t>(), 0u);
            for (int s = 0; s < 2; s++) {
                if (!(s == 0 ? has_l : has_r)) continue;
                MG(sub[s].alloc(nF * 4));
                MG(sxyz[s].alloc(nF * 24));
                // a flagged query goes only to the neighbour whose halo it outreaches
                MG(pcp_select_boundary(ctx, qF.p, nF, 24, ax, kthF.as<float>(), s == 0, xlo - H, s == 1, xhi + H, nF, sub[s].as<int32_t>(),
                                       sxyz[s].as<double>(), &csub[s]));
            }
        }
        for (int s = 0; s < 2; s++) S.bq[s][r] = sxyz[s].p, S.bqn[s][r] = csub[s];
        sync();
        B.wait();
        // t = 0: the left neighbour's queries for its RIGHT side (this rank); t = 1: the right neighbour's for its LEFT
        for (int t = 0; t < 2; t++) {
            const int nb = t == 0 ? r - 1 : r + 1;
            S.ans_xyz[t][r] = S.ans_d2[t][r] = nullptr;
            if (nb < 0 || nb >= world || S.failed.load()) continue;
            const size_t nq = S.bqn[1 - t][nb];
            if (!nq) continue;
            MG(aq[t].alloc(nq * 24));
            MG(aidx[t].alloc(nq * k * 4));
            MG(ad2[t].alloc(nq * k * 8));
            MG(axyz[t].alloc(nq * k * 24));
            if (S.failed.load()) continue;
            peer_pull(aq[t].p, S.bq[1 - t][nb], nb, nq * 24);
            MG(pcp_knn(ix, aq[t].p, nq, 24, k, aidx[t].as<int32_t>(), ad2[t].as<double>(), nullptr));
            xyz_gather_kernel<<<blocks(nq * k), 256, 0, st>>>(work.as<char>(), 32, aidx[t].as<int32_t>(), nq * k, axyz[t].as<double>());
            S.ans_xyz[t][r] = axyz[t].p, S.ans_d2[t][r] = ad2[t].p;
        }
        sync();
        B.wait();
        if (nF && !S.failed.load()) {
            int nlists = 1;
            for (int s = 0; s < 2; s++) nlists += csub[s] ? 1 : 0;
            Dev oidx, xyz_all, d2_all, txyz, td2;
            const size_t per = nF * (size_t)k;
            MG(oidx.alloc(per * 4));
            MG(xyz_all.alloc((size_t)nlists * per * 24));
            MG(d2_all.alloc((size_t)nlists * per * 8));
            MG(pcp_knn(ix, qF.p, nF, 24, k, oidx.as<int32_t>(), d2_all.as<double>(), nullptr));
            if (!S.failed.load()) xyz_gather_kernel<<<blocks(per), 256, 0, st>>>(work.as<char>(), 32, oidx.as<int32_t>(), per, xyz_all.as<double>());
            int l = 1;
            for (int s = 0; s < 2 && !S.failed.load(); s++) {
                if (!csub[s]) continue;
                const int nb = s == 0 ? r - 1 : r + 1;
                const size_t c = csub[s];
                MG(txyz.alloc(c * k * 24));
                MG(td2.alloc(c * k * 8));
                if (S.failed.load()) break;
                // the neighbour answered in its slot t = 1 - s (asked from its other side)
                peer_pull(txyz.p, S.ans_xyz[1 - s][nb], nb, c * k * 24);
                peer_pull(td2.p, S.ans_d2[1 - s][nb], nb, c * k * 8);
                double* lx = xyz_all.as<double>() + (size_t)l * per * 3;
                double* ld = d2_all.as<double>() + (size_t)l * per;
                MGC(cudaMemsetAsync(lx, 0, per * 24, st)); // queries not sent to this side: an empty list (d2 = +inf)
                fill_f64_kernel<<<blocks(per), 256, 0, st>>>(ld, per, (double)INFINITY);
                rows_scatter_kernel<<<blocks(c * k * 6), 256, 0, st>>>(txyz.as<uint32_t>(), k * 6, sub[s].as<int32_t>(), c, reinterpret_cast<uint32_t*>(lx));
                rows_scatter_kernel<<<blocks(c * k * 2), 256, 0, st>>>(td2.as<uint32_t>(), k * 2, sub[s].as<int32_t>(), c, reinterpret_cast<uint32_t*>(ld));
                sync(); // txyz / td2 are reused by the other side
                l++;
            }
            MG(pcp_features_merge(ctx, nF, k, nlists, xyz_all.as<double>(), d2_all.as<double>(), F.as<int32_t>(), props.p, nullptr, nullptr));
        }
        sync();
        B.wait(); // the neighbours have pulled their answers: the buffers may go
    }
    // ---- I. results to the host: slab after slab ----
    size_t off = 0;
    for (int q = 0; q < r; q++) off += S.m[q];
    if (m && !S.failed.load()) {
        pointid_offset_kernel<<<blocks(m), 256, 0, st>>>(props.as<char>(), m, (int32_t)off);
        if (S.out_props) MGC(cudaMemcpyAsync(S.out_props + off * 48, props.p, m * 48, cudaMemcpyDeviceToHost, st));
        if (S.out_pts) {
            if (S.stride == 32) {
                MGC(cudaMemcpyAsync(S.out_pts + off * 32, work.p, m * 32, cudaMemcpyDeviceToHost, st));
            } else {
                Dev un;
                MG(un.alloc(m * 48));
                MG(pcp_cloud_unpack(ctx, work.p, m, un.p));
                MGC(cudaMemcpyAsync(S.out_pts + off * 48, un.p, m * 48, cudaMemcpyDeviceToHost, st));
                sync();
            }
        }
    }
    sync();
    if (ix) pcp_index_free(ix);
}

} // namespace

extern "C" {

int pcp_mgpu_create(int ndev, const int* devices, pcp_mgpu** out)
{
    if (!out || ndev < 1 || ndev > 64) return pcp_fail(nullptr, PCP_ERR_INVALID, "pcp_mgpu_create: bad argument");
    *out = nullptr;
    pcp_mgpu* mg = new pcp_mgpu();
    for (int i = 0; i < ndev; i++) {
        pcp_ctx* c = nullptr;
        const int st = pcp_create(&c, devices ? devices[i] : i);
        if (st != PCP_OK) {
            for (pcp_ctx* q : mg->ctx) pcp_destroy(q);
            delete mg;
            return st; // the message is pcp_create's (pcp_last_error(NULL))
        }
        mg->ctx.push_back(c);
    }
    // peer access both ways where the hardware has it (NVLink / NVSwitch on a B200 box); without it the copies still
    // work, staged by the driver
    for (int i = 0; i < ndev; i++)
        for (int j = 0; j < ndev; j++) {
            if (i == j || mg->ctx[i]->device == mg->ctx[j]->device) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, mg->ctx[i]->device, mg->ctx[j]->device);
            if (can) {
                cudaSetDevice(mg->ctx[i]->device);
                const cudaError_t e = cudaDeviceEnablePeerAccess(mg->ctx[j]->device, 0);
                if (e != cudaSuccess) cudaGetLastError(); // already enabled
            }
        }
    *out = mg;
    return PCP_OK;
}

void pcp_mgpu_destroy(pcp_mgpu* mg)
{
    if (!mg) return;
    for (pcp_ctx* c : mg->ctx) pcp_destroy(c);
    delete mg;
}

int pcp_mgpu_device_count(const pcp_mgpu* mg) { return mg ? (int)mg->ctx.size() : 0; }
pcp_ctx* pcp_mgpu_ctx(pcp_mgpu* mg, int i) { return mg && i >= 0 && i < (int)mg->ctx.size() ? mg->ctx[i] : nullptr; }
const char* pcp_mgpu_last_error(const pcp_mgpu* mg) { return mg ? mg->err.c_str() : ""; }

int pcp_mgpu_features_knn(pcp_mgpu* mg, const void* pts, size_t n, size_t stride, float leaf, int k, double halo, void* out_pts,
                          void* props, size_t* m_out)
{
    if (!mg || !m_out || k < 1 || k > 128 || !(halo > 0) || (stride != 32 && stride != 48) || (n && !pts))
        return pcp_fail(nullptr, PCP_ERR_INVALID, "pcp_mgpu_features_knn: bad argument (stride 32 or 48, 1 <= k <= 128, halo > 0)");
    *m_out = 0;
    if (n == 0) return PCP_OK;
    const int world = (int)mg->ctx.size();
    Shared S;
    S.world = world, S.pts = static_cast<const char*>(pts), S.n = n, S.stride = stride, S.leaf = leaf, S.k = k, S.halo = halo;
    S.out_pts = static_cast<char*>(out_pts), S.out_props = static_cast<char*>(props);
    S.stats.assign((size_t)world * 10, 0.0);
    S.hist.resize(world);
    S.counts.resize(world);
    S.part.assign(world, nullptr);
    S.m.assign(world, 0);
    S.halo_stage.assign(world, nullptr);
    S.halo_l.assign(world, 0), S.halo_r.assign(world, 0);
    for (int s = 0; s < 2; s++) {
        S.bq[s].assign(world, nullptr), S.bqn[s].assign(world, 0);
        S.ans_xyz[s].assign(world, nullptr), S.ans_d2[s].assign(world, nullptr);
    }
    Barrier B(world);
    std::vector<std::thread> th;
    for (int r = 0; r < world; r++) th.emplace_back(worker, mg, std::ref(S), std::ref(B), r);
    for (auto& t : th) t.join();
    if (S.failed.load()) {
        mg->err = S.err;
        return pcp_fail(nullptr, S.status, "pcp_mgpu_features_knn: %s", S.err.c_str());
    }
    size_t m = 0;
    for (int r = 0; r < world; r++) m += S.m[r];
    *m_out = m;
    return PCP_OK;
}

} // extern "C"
