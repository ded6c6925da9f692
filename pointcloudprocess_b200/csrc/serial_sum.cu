// serial_sum.cu -- left-to-right double accumulation of x, y, z over the finite points of a cloud,
// bit-identical to the reference's serial loop in compute3DCentroid (point_cloud_helper.h:203-226).
//
// Floating-point addition is not associative, so a tree reduction gives a centroid that differs from the
// reference's in the last bits -- enough to move ~1e-5 of the points across a voxel boundary after
// remove_duplicate's centring (point_cloud_helper.cpp:44-52).  The ORDER of additions is therefore kept,
// but the work is made parallel with one observation: while the running sum s stays inside one binade
// [2^e, 2^(e+1)), fl(s + x) = s + R(x) where R rounds x to a multiple of u = 2^(e-52) -- an INTEGER
// addition in units of u.  The only state the rounding depends on is the parity of s/u, and only when
// x/u has a fractional part of exactly 1/2 (round-half-even); after such a tie the parity is even.
// A run of elements is therefore summarised by a small transfer record (Xfer) that composes
// associatively, so runs are reduced in parallel and only the few runs in which s changes binade
// (<= ~2 per power of two of n) are replayed element by element.
//
//   pass 1  leaf sums (256 points per leaf, plain tree sums) -> exclusive scan = predicted s at each leaf
//   pass 2  per leaf and axis: Xfer under the predicted exponent          (whole GPU, reads xyz once more)
//   pass 3  Xfer of groups of 32 leaves, of super-groups of 32 groups     (tiny)
//   pass 4  one warp per axis walks the super-groups with the TRUE s: a record is applied only if its
//           exponent equals that of s and s + [min prefix, max prefix] stays inside the binade; otherwise
//           the walk descends (group -> leaf -> elements).  Mispredictions cost time, never correctness.
#include "reduce.cuh"
#include "sort.cuh"
#include <climits>

namespace {

#define SS_LEAF 256          // points per leaf (one warp, 8 consecutive points per lane)
#define SS_FAN 32            // children per group / super-group
#define SS_E_DIRTY INT_MIN   // record unusable: replay the children
#define SS_E_ANY INT_MAX     // identity record (no elements)

struct __align__(16) Xfer {
    long long a;     // !tie: total added (units of u).  tie: sum up to and including floor(x/u) of the first tie
    long long post;  // tie: total added after the first tie element (parity after a tie is even, so it is known)
    long long minp;  // min / max over the prefix sums (ties counted as floor; widened by ntie when checked)
    long long maxp;
    int ntie;
    int e;           // exponent of the binade the record was computed for
    long long tie;   // 0 / 1
};

__device__ __forceinline__ Xfer xfer_identity()
{
    Xfer t;
    t.a = t.post = t.minp = t.maxp = 0;
    t.ntie = 0;
    t.e = SS_E_ANY;
    t.tie = 0;
    return t;
}

__device__ __forceinline__ long long xfer_sum0(const Xfer& t) { return t.tie ? t.a + t.post : t.a; }

// L then R (order matters)
__device__ __forceinline__ Xfer xfer_compose(const Xfer& L, const Xfer& R)
{
    Xfer o;
    if (L.e == SS_E_ANY) return R;
    if (R.e == SS_E_ANY) return L;
    o.e = (L.e == R.e && L.e != SS_E_DIRTY) ? L.e : SS_E_DIRTY;
    const long long ls = xfer_sum0(L);
    o.minp = min(L.minp, ls + R.minp);
    o.maxp = max(L.maxp, ls + R.maxp);
    o.ntie = L.ntie + R.ntie;
    if (!L.tie) {
        o.a = L.a + R.a;
        o.post = R.post;
        o.tie = R.tie;
    } else {
        long long tot2 = R.a;
        if (R.tie) tot2 += ((L.post + R.a) & 1ll) + R.post; // parity entering R = parity(L.post)
        o.a = L.a;
        o.post = L.post + tot2;
        o.tie = 1;
    }
    return o;
}

__device__ __forceinline__ Xfer xfer_shfl_down(const Xfer& t, int off)
{
    Xfer o;
    o.a = __shfl_down_sync(0xffffffffu, t.a, off);
    o.post = __shfl_down_sync(0xffffffffu, t.post, off);
    o.minp = __shfl_down_sync(0xffffffffu, t.minp, off);
    o.maxp = __shfl_down_sync(0xffffffffu, t.maxp, off);
    o.ntie = __shfl_down_sync(0xffffffffu, t.ntie, off);
    o.e = __shfl_down_sync(0xffffffffu, t.e, off);
    o.tie = __shfl_down_sync(0xffffffffu, t.tie, off);
    return o;
}

// ordered reduction: lane 0 ends with t(0) o t(1) o ... o t(31)
__device__ __forceinline__ Xfer xfer_warp_reduce(Xfer t)
{
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        Xfer r = xfer_shfl_down(t, off);
        t = xfer_compose(t, r);
    }
    return t;
}

// 2^k as a double, k in [-1022, 1023]
__device__ __forceinline__ double pow2d(int k) { return __longlong_as_double((long long)(1023 + k) << 52); }

__device__ __forceinline__ int exponent_of(double v)
{
    return (int)((__double_as_longlong(v) >> 52) & 0x7ff) - 1023;
}

// one element under exponent e (scale = 2^(52-e)); returns false if it cannot be expressed safely
__device__ __forceinline__ bool xfer_element(double x, double scale, Xfer& t)
{
    t.a = t.post = t.minp = t.maxp = 0;
    t.ntie = 0;
    t.tie = 0;
    const double at = fabs(x) * scale;      // exact: power-of-two scaling
    if (!(at < 1099511627776.0)) return false; // |x/u| >= 2^40 (or inf/nan): x is not small against the binade
    const double qa = floor(at);
    const double fa = at - qa;              // exact
    const long long q = (long long)qa;
    const bool neg = x < 0.0;
    long long r;
    t.tie = 0;
    t.post = 0;
    t.ntie = 0;
    if (fa == 0.5) {
        r = neg ? -q - 1 : q;               // floor(x/u); the parity decides whether 1 is added
        t.tie = 1;
        t.ntie = 1;
    } else {
        const long long m = q + (fa > 0.5 ? 1 : 0);
        r = neg ? -m : m;
    }
    t.a = r;
    t.minp = min(0ll, r);
    t.maxp = max(0ll, r);
    return true;
}

// coordinates of point i, zeros if the point is skipped (any non-finite coordinate) or out of range
__device__ __forceinline__ void load_xyz(const char* __restrict__ pts, size_t n, size_t stride, size_t i, double& x,
                                         double& y, double& z, bool& ok)
{
    x = y = z = 0.0;
    ok = false;
    if (i < n) {
        const double* p = reinterpret_cast<const double*>(pts + i * stride);
        const double a = p[0], b = p[1], c = p[2];
        if (isfinite(a) && isfinite(b) && isfinite(c)) x = a, y = b, z = c, ok = true;
    }
}

// ---- pass 1: plain (tree-order) leaf sums, only used to PREDICT the binade of s at each leaf ----------------
__global__ void __launch_bounds__(128) ss_leaf_sum_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                          size_t nleaf, double* __restrict__ lsum /*[3][nleaf]*/)
{
    const int lane = threadIdx.x & 31;
    const size_t warp0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t leaf = warp0; leaf < nleaf; leaf += nwarp) {
        double sx = 0, sy = 0, sz = 0;
#pragma unroll
        for (int j = 0; j < SS_LEAF / 32; j++) {
            double x, y, z;
            bool ok;
            load_xyz(pts, n, stride, leaf * SS_LEAF + 32 * j + lane, x, y, z, ok);
            sx += x, sy += y, sz += z;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sx += __shfl_xor_sync(0xffffffffu, sx, o);
            sy += __shfl_xor_sync(0xffffffffu, sy, o);
            sz += __shfl_xor_sync(0xffffffffu, sz, o);
        }
        if (lane == 0) lsum[leaf] = sx, lsum[nleaf + leaf] = sy, lsum[2 * nleaf + leaf] = sz;
    }
}

// ---- pass 2: transfer record of every leaf under the predicted exponent ---------------------------------------
__device__ __forceinline__ int ss_pad(int i) { return i + (i >> 3); }

__global__ void __launch_bounds__(128) ss_leaf_xfer_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                           size_t nleaf, const double* __restrict__ lpre /*[3][nleaf]*/,
                                                           const double* __restrict__ carry3 /*or NULL*/,
                                                           Xfer* __restrict__ out /*[3][nleaf]*/)
{
    __shared__ double sm[4][3][SS_LEAF + SS_LEAF / 8];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const size_t warp0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t leaf = warp0; leaf < nleaf; leaf += nwarp) {
        __syncwarp();
#pragma unroll
        for (int j = 0; j < SS_LEAF / 32; j++) {
            double x, y, z;
            bool ok;
            load_xyz(pts, n, stride, leaf * SS_LEAF + 32 * j + lane, x, y, z, ok);
            const int s = ss_pad(32 * j + lane);
            sm[w][0][s] = x, sm[w][1][s] = y, sm[w][2][s] = z;
        }
        __syncwarp();
#pragma unroll
        for (int axis = 0; axis < 3; axis++) {
            const double pre = lpre[(size_t)axis * nleaf + leaf] + (carry3 ? carry3[axis] : 0.0);
            const int e = exponent_of(pre);
            const bool usable = (pre != 0.0) && e > -900 && e < 1000;
            const double scale = usable ? pow2d(52 - e) : 1.0;
            Xfer t = xfer_identity();
            bool good = usable;
#pragma unroll
            for (int j = 0; j < SS_LEAF / 32; j++) {
                const double v = sm[w][axis][ss_pad(8 * lane + j)];
                Xfer el;
                good = xfer_element(v, scale, el) && good;
                el.e = e;
                t = (j == 0) ? el : xfer_compose(t, el);
            }
            t.e = __all_sync(0xffffffffu, good) ? e : SS_E_DIRTY;
            t = xfer_warp_reduce(t);
            if (lane == 0) out[(size_t)axis * nleaf + leaf] = t;
        }
    }
}

// ---- pass 3: records of groups of SS_FAN children ---------------------------------------------------------------
__global__ void __launch_bounds__(128) ss_group_kernel(const Xfer* __restrict__ child, size_t nchild, size_t ngroup,
                                                       Xfer* __restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const size_t warp0 = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t nwarp = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t g = warp0; g < 3 * ngroup; g += nwarp) {
        const size_t axis = g / ngroup, gi = g % ngroup;
        const size_t c = gi * SS_FAN + lane;
        Xfer t = c < nchild ? child[axis * nchild + c] : xfer_identity();
        t = xfer_warp_reduce(t);
        if (lane == 0) out[axis * ngroup + gi] = t;
    }
}

// ---- pass 4: the walk -------------------------------------------------------------------------------------------
// Applies a record to the true running sum; false = the record does not hold for this s (replay the children).
__device__ __forceinline__ bool ss_apply(const Xfer& t, double& s)
{
    if (t.e == SS_E_ANY) return true;
    if (t.e == SS_E_DIRTY || s == 0.0 || exponent_of(s) != t.e) return false;
    const long long S = (long long)(s * pow2d(52 - t.e)); // exact: |S| in [2^52, 2^53)
    const long long lo = S + t.minp - t.ntie, hi = S + t.maxp + t.ntie;
    const long long B52 = 1ll << 52, B53 = 1ll << 53;
    const bool inside = (S > 0) ? (lo >= B52 && hi < B53) : (hi <= -B52 && lo > -B53);
    if (!inside) return false;
    long long tot = t.a;
    if (t.tie) tot += ((S + t.a) & 1ll) + t.post;
    s = (double)(S + tot) * pow2d(t.e - 52); // exact
    return true;
}

// element-by-element replay of points [first, first+count), every lane keeps the same s
__device__ __forceinline__ void ss_replay(const char* __restrict__ pts, size_t n, size_t stride, size_t first,
                                          size_t count, int axis, int lane, double& s)
{
    const size_t end = min(first + count, n);
    for (size_t base = first; base < end; base += 32) {
        double x, y, z;
        bool ok;
        load_xyz(pts, end, stride, base + lane, x, y, z, ok);
        const double v = axis == 0 ? x : (axis == 1 ? y : z);
        unsigned okmask = __ballot_sync(0xffffffffu, ok);
        if (okmask == 0xffffffffu) {
#pragma unroll
            for (int j = 0; j < 32; j++) s = __dadd_rn(s, __shfl_sync(0xffffffffu, v, j));
        } else {
            while (okmask) {
                const int j = __ffs(okmask) - 1;
                okmask &= okmask - 1;
                s = __dadd_rn(s, __shfl_sync(0xffffffffu, v, j));
            }
        }
    }
}

// stage up to 32 records [first, first+cnt) of `src` into this warp's shared buffer
__device__ __forceinline__ void ss_stage(const Xfer* __restrict__ src, size_t first, size_t total, int lane, Xfer* buf)
{
    __syncwarp();
    buf[lane] = (first + lane < total) ? src[first + lane] : xfer_identity();
    __syncwarp();
}

__global__ void __launch_bounds__(96) ss_walk_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                     const Xfer* __restrict__ leafx, size_t nleaf,
                                                     const Xfer* __restrict__ groupx, size_t ngroup,
                                                     const Xfer* __restrict__ superx, size_t nsuper,
                                                     double* __restrict__ out3, int carry,
                                                     unsigned long long* __restrict__ stats)
{
    __shared__ Xfer bufS[3][32], bufG[3][32], bufL[3][32];
    const int axis = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Xfer* lx = leafx + (size_t)axis * nleaf;
    const Xfer* gx = groupx + (size_t)axis * ngroup;
    const Xfer* sx = superx + (size_t)axis * nsuper;
    double s = carry ? out3[axis] : 0.0;
    unsigned long long replayed = 0;
    for (size_t s0 = 0; s0 < nsuper; s0 += 32) {
        ss_stage(sx, s0, nsuper, lane, bufS[axis]);
        for (int si = 0; si < 32 && s0 + si < nsuper; si++) {
            if (ss_apply(bufS[axis][si], s)) continue;
            const size_t g0 = (s0 + si) * SS_FAN;
            ss_stage(gx, g0, ngroup, lane, bufG[axis]);
            for (int gi = 0; gi < 32 && g0 + gi < ngroup; gi++) {
                if (ss_apply(bufG[axis][gi], s)) continue;
                const size_t l0 = (g0 + gi) * SS_FAN;
                ss_stage(lx, l0, nleaf, lane, bufL[axis]);
                for (int li = 0; li < 32 && l0 + li < nleaf; li++) {
                    if (ss_apply(bufL[axis][li], s)) continue;
                    ss_replay(pts, n, stride, (l0 + li) * SS_LEAF, SS_LEAF, axis, lane, s);
                    replayed++;
                }
            }
        }
    }
    if (lane == 0) {
        out3[axis] = s;
        if (stats) stats[axis] = replayed;
    }
}

// small clouds: one warp per axis, element by element
__global__ void __launch_bounds__(96) ss_small_kernel(const char* __restrict__ pts, size_t n, size_t stride,
                                                      double* __restrict__ out3, int carry)
{
    const int axis = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double s = carry ? out3[axis] : 0.0;
    ss_replay(pts, n, stride, 0, n, axis, lane, s);
    if (lane == 0) out3[axis] = s;
}

} // namespace

struct pcp_serial_plan {
    const char* p = nullptr;
    size_t n = 0, stride = 0, nleaf = 0, ngroup = 0, nsuper = 0;
    DevBuf<Xfer> leafx, groupx, superx;
};

int pcp_serial_sum_prepare(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, const double* d_pred3, pcp_serial_plan** out)
{
    pcp_serial_plan* P = new pcp_serial_plan();
    *out = P;
    P->p = static_cast<const char*>(dpts);
    P->n = n, P->stride = stride;
    if (n < 16384) return PCP_OK; // small clouds are walked element by element in `finish`
    const size_t nleaf = P->nleaf = (n + SS_LEAF - 1) / SS_LEAF;
    const size_t ngroup = P->ngroup = (nleaf + SS_FAN - 1) / SS_FAN;
    const size_t nsuper = P->nsuper = (ngroup + SS_FAN - 1) / SS_FAN;
    DevBuf<double> lsum, lpre;
    PCP_TRY(lsum.alloc(ctx, 3 * nleaf));
    PCP_TRY(lpre.alloc(ctx, 3 * nleaf));
    PCP_TRY(P->leafx.alloc(ctx, 3 * nleaf));
    PCP_TRY(P->groupx.alloc(ctx, 3 * ngroup));
    PCP_TRY(P->superx.alloc(ctx, 3 * nsuper));
    const int wide = ctx->sm_count * 16; // 4 warps per block, grid-stride over leaves
    const int blocks = (int)(((nleaf + 3) / 4 < (size_t)wide) ? (nleaf + 3) / 4 : (size_t)wide);
    ss_leaf_sum_kernel<<<blocks, 128, 0, ctx->stream>>>(P->p, n, stride, nleaf, lsum.p);
    PCP_LAUNCH_CHECK(ctx);
    for (int a = 0; a < 3; a++) PCP_TRY(pcp_exclusive_scan_f64(ctx, lsum.p + a * nleaf, lpre.p + a * nleaf, nleaf));
    ss_leaf_xfer_kernel<<<blocks, 128, 0, ctx->stream>>>(P->p, n, stride, nleaf, lpre.p, d_pred3, P->leafx.p);
    PCP_LAUNCH_CHECK(ctx);
    const int gb = (int)((3 * ngroup + 3) / 4 < (size_t)wide ? (3 * ngroup + 3) / 4 : (size_t)wide);
    ss_group_kernel<<<gb, 128, 0, ctx->stream>>>(P->leafx.p, nleaf, ngroup, P->groupx.p);
    PCP_LAUNCH_CHECK(ctx);
    const int sb = (int)((3 * nsuper + 3) / 4 < (size_t)wide ? (3 * nsuper + 3) / 4 : (size_t)wide);
    ss_group_kernel<<<sb, 128, 0, ctx->stream>>>(P->groupx.p, ngroup, nsuper, P->superx.p);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

int pcp_serial_sum_finish(pcp_ctx* ctx, pcp_serial_plan* P, double* d_carry3)
{
    if (P->n < 16384) {
        ss_small_kernel<<<1, 96, 0, ctx->stream>>>(P->p, P->n, P->stride, d_carry3, 1);
        PCP_LAUNCH_CHECK(ctx);
        return PCP_OK;
    }
    ss_walk_kernel<<<1, 96, 0, ctx->stream>>>(P->p, P->n, P->stride, P->leafx.p, P->nleaf, P->groupx.p, P->ngroup, P->superx.p,
                                            P->nsuper, d_carry3, 1, nullptr);
    PCP_LAUNCH_CHECK(ctx);
    return PCP_OK;
}

void pcp_serial_sum_release(pcp_ctx*, pcp_serial_plan* P) { delete P; }

// carry != 0: d_sum3 holds the running sums of the points that precede this chunk (chunked / sharded clouds)
int pcp_serial_sum_device(pcp_ctx* ctx, const void* dpts, size_t n, size_t stride, double* d_sum3, int carry)
{
    if (!carry) PCP_CUDA_TRY(ctx, cudaMemsetAsync(d_sum3, 0, 24, ctx->stream));
    pcp_serial_plan* P = nullptr;
    int st = pcp_serial_sum_prepare(ctx, dpts, n, stride, carry ? d_sum3 : nullptr, &P);
    if (st == PCP_OK) st = pcp_serial_sum_finish(ctx, P, d_sum3);
    pcp_serial_sum_release(ctx, P);
    return st;
}

// ---- C ABI of the two-step form (include/pcp_cuda.h) ----------------------------------------------------------------
struct pcp_serial_handle {
    pcp_ctx* ctx;
    pcp_serial_plan* plan;
};

extern "C" int pcp_serial_sum_begin(pcp_ctx* ctx, const void* dev_pts, size_t n, size_t stride, const double predicted3[3],
                                    void** handle)
{
    if (!ctx || !handle || stride < 24 || (n && !dev_pts)) return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_serial_sum_begin: bad argument");
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (n && pcp_classify(ctx, dev_pts) != PCP_PTR_DEVICE)
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_serial_sum_begin: the cloud must be in device memory (it is read again by pcp_serial_sum_end)");
    DevBuf<double> pred;
    if (predicted3) {
        PCP_TRY(pred.alloc(ctx, 3));
        PCP_TRY(pcp_writeto(ctx, pred.p, predicted3, 24));
    }
    pcp_serial_plan* P = nullptr;
    StageTimer t(ctx, "shard.serial_sum_begin", 24.0 * n);
    const int st = pcp_serial_sum_prepare(ctx, dev_pts, n, stride, predicted3 ? pred.p : nullptr, &P);
    if (st != PCP_OK) {
        pcp_serial_sum_release(ctx, P);
        return st;
    }
    *handle = new pcp_serial_handle{ctx, P};
    return PCP_OK;
}

extern "C" int pcp_serial_sum_end(void* handle, double* carry3_inout)
{
    pcp_serial_handle* H = static_cast<pcp_serial_handle*>(handle);
    if (!H) return PCP_ERR_INVALID;
    pcp_ctx* ctx = H->ctx;
    int st = PCP_OK;
    {
        std::lock_guard<std::recursive_mutex> lock(ctx->mu);
        cudaSetDevice(ctx->device);
        if (carry3_inout) {
            StageTimer t(ctx, "shard.serial_sum_end");
            // device array: walked in place, no synchronisation (the value can travel on to the next GPU);
            // host array: staged through the context
            if (pcp_classify(ctx, carry3_inout) == PCP_PTR_DEVICE) {
                st = pcp_serial_sum_finish(ctx, H->plan, carry3_inout);
            } else {
                DevBuf<double> c;
                st = c.alloc(ctx, 3);
                if (st == PCP_OK) st = pcp_writeto(ctx, c.p, carry3_inout, 24);
                if (st == PCP_OK) st = pcp_serial_sum_finish(ctx, H->plan, c.p);
                if (st == PCP_OK) st = pcp_readback(ctx, carry3_inout, c.p, 24);
            }
        }
        pcp_serial_sum_release(ctx, H->plan);
    }
    delete H;
    return st;
}
