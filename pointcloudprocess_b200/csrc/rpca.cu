// rpca.cu -- CalculateFeature::calculate_plan_parameter_rpca (calculate_feature.cpp:208-368), the robust plane
// fit the reference ships: per point, `iterations` 3-point hypotheses; for each, the h = (int)(2/3 N) neighbours
// closest to the hypothesis plane are refitted and the refit with the smallest lambda3 is kept; the neighbours
// that pass a MAD test on their distance to that plane (Rz < 2.5) are fitted once more and that plane is the result.
//
// The reference draws the hypotheses with rand() under OpenMP (seeded by time): its output differs run to run.
// This restatement replaces the generator by a counter-based one (splitmix64 of (seed, point, hypothesis, draw))
// so that the result is a pure function of (cloud, k, Pr, epi, seed) and can be checked bit for bit against the
// CPU restatement the tests use; every other step keeps the reference's arithmetic: the h-subset is refitted through
// the float members of LAS_POINT_PROPERTY_sim (data_struct.h:181-187) and the point-to-plane distance is the float
// one of extraction_tree.cpp:47-64.
//
// Layout: one thread per point, 64 threads per block. The N <= 32 neighbour coordinates of a thread live in shared
// memory transposed (element e of thread t at [e * 64 + t]: conflict-free), because they are indexed by the
// random draws and by rank. Sums run in the order the reference runs them (cvMatMul over the columns), which is
// why a point is not spread over a warp: an ordered FP64 sum does not parallelise.
#include "index.cuh"
#include "knn_cell.cuh"
#include "plane_fit.cuh"

#include <cmath>

namespace {

constexpr int RB = 64; // threads per block

struct RpcaParams {
    const pcp_rec* rec;
    const uint32_t* inv;   // original index -> sorted position
    const int32_t* knn_idx; // n_orig x k, ascending (d2, index), -1 padded
    size_t n_orig;
    int k;
    int iterations;
    unsigned long long seed;
    void* props;
};

__global__ void __launch_bounds__(256) rpca_inverse_kernel(const pcp_rec* __restrict__ rec, uint32_t n, uint32_t* inv)
{
    const uint32_t p = blockIdx.x * 256u + threadIdx.x;
    if (p < n) inv[(size_t)rec[p].idx] = p;
}

__device__ __forceinline__ unsigned long long rpca_mix(unsigned long long x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

__device__ __forceinline__ int rpca_draw(unsigned long long base, int i, int n, int N)
{
    return (int)((rpca_mix(base + (unsigned long long)(i * 3 + n)) >> 33) % (unsigned long long)N);
}

struct RpcaPlane {
    float a, b, c, d;
    double g; // (double)sqrtf(a*a + b*b + c*c), the same for every point of one plane
};

__device__ __forceinline__ void rpca_set_plane(RpcaPlane& q, const pcp_plane& pl)
{
    q.a = pl.nx, q.b = pl.ny, q.c = pl.nz, q.d = pl.Distance;
    const float gg = __fadd_rn(__fadd_rn(__fmul_rn(q.a, q.a), __fmul_rn(q.b, q.b)), __fmul_rn(q.c, q.c));
    q.g = (double)__fsqrt_rn(gg); // sqrt(float): MSVC's float overload, widened (extraction_tree.cpp:57)
}

// TreeExtration::compute_distance_from_point_to_plane (extraction_tree.cpp:47-64): float coordinates, float
// products, double sum, double quotient, float result.
__device__ __forceinline__ float rpca_dist(const double* nb, int n, const RpcaPlane& q)
{
    const float x1 = (float)nb[(3 * n + 0) * RB], y1 = (float)nb[(3 * n + 1) * RB], z1 = (float)nb[(3 * n + 2) * RB];
    const double f1 = (double)__fmul_rn(q.a, x1), f2 = (double)__fmul_rn(q.b, y1), f3 = (double)__fmul_rn(q.c, z1);
    const double f = fabs(__dadd_rn(__dadd_rn(__dadd_rn(f1, f2), f3), (double)q.d));
    return (float)(f / q.g);
}

// calculate_plan_parameter_h_points (:119-206) over the neighbours ord[0..h), optionally through the float members
// of LAS_POINT_PROPERTY_sim (:262-266). Inlined at its three call sites (the Jacobi solver inside stays one shared
// function): as a __noinline__ function the kernel built by ptxas 12.9 -O3 kept, in its MAD inlier test, a point it
// computed as an outlier (6 % of the records differed from the oracle; the same PTX assembled at -O1 was right).
// Observed, not explained: compute-sanitizer (initcheck) is closed on this pool, so neither a compiler fault nor an
// uninitialised read on our side is established; the inlined form is bit-equal to the oracle in every test.
__device__ __forceinline__ void rpca_fit(const double* nb, const unsigned char* ord, int h, int as_float, pcp_plane& out)
{
    double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
    for (int e = 0; e < h; e++) {
        const int n = ord[e * RB];
        double x = nb[(3 * n + 0) * RB], y = nb[(3 * n + 1) * RB], z = nb[(3 * n + 2) * RB];
        if (as_float) x = (double)(float)x, y = (double)(float)y, z = (double)(float)z;
        XAve += x, YAve += y, ZAve += z;
    }
    XAve /= (double)h, YAve /= (double)h, ZAve /= (double)h;
    double s00 = 0, s01 = 0, s02 = 0, s11 = 0, s12 = 0, s22 = 0;
    for (int e = 0; e < h; e++) {
        const int n = ord[e * RB];
        double x = nb[(3 * n + 0) * RB], y = nb[(3 * n + 1) * RB], z = nb[(3 * n + 2) * RB];
        if (as_float) x = (double)(float)x, y = (double)(float)y, z = (double)(float)z;
        const double X = x - XAve, Y = y - YAve, Z = z - ZAve;
        s00 += X * X, s01 += X * Y, s02 += X * Z, s11 += Y * Y, s12 += Y * Z, s22 += Z * Z;
    }
    pcp_plane_from_scatter(s00, s01, s02, s11, s12, s22, XAve, YAve, ZAve, out);
}

// position of element n in the stable ascending order of v[0..N)
__device__ __forceinline__ int rpca_rank(const float* v, int n, int N)
{
    const float x = v[n * RB];
    int r = 0;
    for (int m = 0; m < N; m++) {
        const float y = v[m * RB];
        r += (y < x || (y == x && m < n)) ? 1 : 0;
    }
    return r;
}

// the value at position `at` of the ascending order of v[0..N)
__device__ __forceinline__ float rpca_nth(const float* v, int at, int N)
{
    float out = 0.f;
    for (int n = 0; n < N; n++)
        if (rpca_rank(v, n, N) == at) out = v[n * RB];
    return out;
}

__global__ void __launch_bounds__(RB) rpca_kernel(const RpcaParams P)
{
    extern __shared__ double rpca_sm[];
    const int k = P.k;
    double* nb = rpca_sm + threadIdx.x;                                                       // [3k][RB]
    float* dis = reinterpret_cast<float*>(rpca_sm + (size_t)3 * k * RB) + threadIdx.x;         // [k][RB]
    unsigned char* ord = reinterpret_cast<unsigned char*>(rpca_sm + (size_t)3 * k * RB) + (size_t)4 * k * RB +
                         threadIdx.x;                                                         // [k][RB]
    const size_t j = (size_t)blockIdx.x * RB + threadIdx.x;
    if (j >= P.n_orig) return;

    int N = 0;
    for (int e = 0; e < k; e++) {
        const int32_t id = P.knn_idx[j * (size_t)k + e];
        if (id < 0) break;
        const pcp_rec r = pcp_load_rec(P.rec, P.inv[id]);
        nb[(3 * e + 0) * RB] = r.x, nb[(3 * e + 1) * RB] = r.y, nb[(3 * e + 2) * RB] = r.z;
        N++;
    }
    pcp_plane pl;
    bool fitted = false;
    if (N > 3) { // :238
        const int h_free = (int)(2.0 / 3 * N);
        const unsigned long long base = P.seed * 0x9E3779B97F4A7C15ull + (unsigned long long)j * 0xD1B54A32D192ED03ull;
        RpcaPlane best;
        float best_min = 0.f;
        bool have = false;
        for (int i = 0; i < P.iterations; i++) {
            const int n0 = rpca_draw(base, i, 0, N), n1 = rpca_draw(base, i, 1, N), n2 = rpca_draw(base, i, 2, N);
            if (n0 == n1 || n0 == n2 || n1 == n2) continue; // :249
            ord[0] = (unsigned char)n0, ord[RB] = (unsigned char)n1, ord[2 * RB] = (unsigned char)n2;
            rpca_fit(nb, ord, 3, 0, pl); // calculate_plan_parameter_3points
            RpcaPlane q;
            rpca_set_plane(q, pl);
            for (int n = 0; n < N; n++) dis[n * RB] = rpca_dist(nb, n, q);
            for (int n = 0; n < N; n++) ord[rpca_rank(dis, n, N) * RB] = (unsigned char)n;
            rpca_fit(nb, ord, h_free, 1, pl);
            if (!have || pl.min_value < best_min) {
                rpca_set_plane(best, pl);
                best_min = pl.min_value;
                have = true;
            }
        }
        if (have) {
            for (int n = 0; n < N; n++) dis[n * RB] = rpca_dist(nb, n, best);
            const float median = rpca_nth(dis, N / 2, N);
            for (int n = 0; n < N; n++) dis[n * RB] = fabsf(__fsub_rn(dis[n * RB], median));
            const float MAD = (float)(1.4826 * (double)rpca_nth(dis, N / 2, N));
            int m = 0;
            for (int n = 0; n < N; n++) {
                bool keep = true;
                if (MAD != 0.f) keep = (double)(dis[n * RB] / MAD) < 2.5; // :322-330
                if (keep) ord[(m++) * RB] = (unsigned char)n;
            }
            if (m > 3) {
                rpca_fit(nb, ord, m, 0, pl);
                fitted = true;
            }
        }
    }
    pcp_store_props(P.props, j, fitted, pl);
}

} // namespace

extern "C" int pcp_rpca_iterations(float Pr, float epi)
{
    // compute_iteration_number (calculate_feature.cpp:28-33): log10 of a float over log10 of a double
    return (int)(std::log10(1 - Pr) / std::log10(1 - std::pow((double)(1 - epi), 3)));
}

extern "C" int pcp_features_rpca(const pcp_index* ix, int k, float Pr, float epi, uint64_t seed, void* props)
{
    if (!ix || !props || k < 1) return pcp_fail(ix ? ix->ctx : nullptr, PCP_ERR_INVALID, "pcp_features_rpca: bad argument");
    pcp_ctx* ctx = ix->ctx;
    if (k > 32) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_features_rpca: k > 32 is not supported");
    if (!(Pr > 0.f && Pr < 1.f) || !(epi >= 0.f && epi < 1.f))
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_features_rpca: Pr must lie in (0,1) and epi in [0,1)");
    const int iterations = pcp_rpca_iterations(Pr, epi);
    if (iterations < 0 || iterations > 100000)
        return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_features_rpca: iteration count out of range");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const size_t n0 = ix->n_orig;
    if (n0 == 0) return PCP_OK;
    OutStage o_props;
    PCP_TRY(o_props.open(ctx, props, n0 * 48));
    DevBuf<int32_t> nn;
    DevBuf<uint32_t> inv;
    PCP_TRY(nn.alloc(ctx, n0 * (size_t)k));
    PCP_TRY(inv.alloc(ctx, n0));
    // neighbour lists in ascending (d2, index) order; the plain fit it also writes into props is overwritten below
    PCP_TRY(pcp_features_knn_ex(ix, k, -1, o_props.dptr, nullptr, nullptr, nn.p, nullptr));
    {
        StageTimer t(ctx, "features.rpca", (double)n0 * ((double)k * 36 + 48));
        cudaStream_t st = ctx->stream;
        if (ix->n) {
            rpca_inverse_kernel<<<pcp_blocks_for(ix->n, 256), 256, 0, st>>>(ix->rec, (uint32_t)ix->n, inv.p);
            PCP_LAUNCH_CHECK(ctx);
        }
        RpcaParams P;
        P.rec = ix->rec;
        P.inv = inv.p;
        P.knn_idx = nn.p;
        P.n_orig = n0;
        P.k = k;
        P.iterations = iterations;
        P.seed = seed;
        P.props = o_props.dptr;
        const size_t smem = (size_t)k * RB * (24 + 4 + 1);
        if (smem > 48 * 1024)
            PCP_CUDA_TRY(ctx, cudaFuncSetAttribute(rpca_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rpca_kernel<<<pcp_blocks_for(n0, RB), RB, smem, st>>>(P);
        PCP_LAUNCH_CHECK(ctx);
    }
    return o_props.commit();
}
