// plane_fit.cuh -- CalculateFeature::calculate_plan_parameter_h_points (calculate_feature.cpp:119-206) for
// one neighbourhood, register-resident: mean, un-normalised scatter X*X^T, then the 3x3 symmetric
// eigen-decomposition cvEigenVV performs (OpenCV 2.4.x JacobiImpl_<double>: largest-off-diagonal pivot
// tracked through indR/indC, its own scaled hypot, eigenvalues sorted descending, eigenvectors as rows).
// Written for n = 3 with every matrix element in a named register -- no local-memory arrays.
// Built with -fmad=false: each product and sum is rounded separately, as in the reference's MSVC build.
#pragma once
#include <cfloat>
#include "../../include/pcp_cuda.h"

struct pcp_plane {
    float nx, ny, nz;
    float min_value, curvature, Distance;
    double l1, l2, l3;
};

// OpenCV's scaled hypot (a > b: a * sqrt(1 + (b/a)^2); else b > 0: b * sqrt(1 + (a/b)^2); else 0), with the larger
// and the smaller operand chosen by selects: the two sides of the branch are the same arithmetic, and a warp whose
// lanes disagree on a > b paid for the division and the square root twice.
__device__ __forceinline__ double pcp_cv_hypot(double a, double b)
{
    a = fabs(a);
    b = fabs(b);
    const bool ab = a > b;
    const double big = ab ? a : b, small = ab ? b : a;
    const double r = small / big;
    const double v = big * sqrt(1.0 + r * r);
    return big > 0 ? v : 0.0;
}

#define PCP_ROT(v0, v1)                                                                                      \
    {                                                                                                        \
        double a0_ = v0, b0_ = v1;                                                                           \
        v0 = a0_ * c - b0_ * s;                                                                              \
        v1 = a0_ * s + b0_ * c;                                                                              \
    }

// S = [[s00 s01 s02],[. s11 s12],[. . s22]].  Outputs W (descending) and rows of V.
// (INLINE: the plane-fit kernel takes it inline -- the eigenvalues and vectors stay in registers; the kernels whose
// code has to stay small call the copy below)
template <bool INLINE = true>
__device__ __forceinline__ void pcp_jacobi3_impl(double s00, double s01, double s02, double s11, double s12, double s22,
                                                 double W[3], double V[9])
{
    const double eps = DBL_EPSILON;
    double a01 = s01, a02 = s02, a12 = s12;
    double w0 = s00, w1 = s11, w2 = s22;
    double v00 = 1, v01 = 0, v02 = 0, v10 = 0, v11 = 1, v12 = 0, v20 = 0, v21 = 0, v22 = 1;
    // indR[0] in {1,2}; indR[1] == 2; indC[1] == 0; indC[2] in {0,1}
    int indR0 = (fabs(a01) < fabs(a02)) ? 2 : 1;
    int indC2 = (fabs(a02) < fabs(a12)) ? 1 : 0;
    for (int iters = 0; iters < 270; iters++) {
        int k = 0;
        double mv = fabs(indR0 == 1 ? a01 : a02);
        double val = fabs(a12);
        if (mv < val) mv = val, k = 1;
        int l = (k == 0) ? indR0 : 2;
        val = fabs(a01); // i = 1: A[indC[1]][1]
        if (mv < val) mv = val, k = 0, l = 1;
        val = fabs(indC2 == 0 ? a02 : a12); // i = 2: A[indC[2]][2]
        if (mv < val) mv = val, k = indC2, l = 2;
        const int pair = (k == 0) ? (l == 1 ? 0 : 1) : 2; // (0,1) (0,2) (1,2)
        double p = pair == 0 ? a01 : (pair == 1 ? a02 : a12);
        if (fabs(p) <= eps) break;
        double wk = (k == 0) ? w0 : w1;
        double wl = (l == 1) ? w1 : w2;
        double y = (wl - wk) * 0.5;
        double t = fabs(y) + pcp_cv_hypot(p, y);
        double s = pcp_cv_hypot(p, t);
        double c = t / s;
        s = p / s;
        t = (p / t) * p;
        if (y < 0) s = -s, t = -t;
        // the rotation of rows/columns (k, l) = (0,1) (0,2) (1,2), operands chosen by selects instead of three
        // copies of the same 26 operations under a three-way branch (the lanes of a warp rarely agree on the pivot)
        const bool p0 = pair == 0, p2 = pair == 2;
        const double nwk = wk - t, nwl = wl + t;
        w0 = p2 ? w0 : nwk;
        w1 = p0 ? nwl : (p2 ? nwk : w1);
        w2 = p0 ? w2 : nwl;
        double ra = p0 ? a02 : a01; // pair 0: (a02, a12), i = 2 > l; pair 1: (a01, a12), k < i = 1 < l;
        double rb = p2 ? a02 : a12; // pair 2: (a01, a02), i = 0 < k
        PCP_ROT(ra, rb);
        a01 = p0 ? 0.0 : ra;
        a02 = p0 ? ra : (p2 ? rb : 0.0);
        a12 = p2 ? 0.0 : rb;
        double k0 = p2 ? v10 : v00, k1 = p2 ? v11 : v01, k2 = p2 ? v12 : v02; // row k of V
        double l0 = p0 ? v10 : v20, l1 = p0 ? v11 : v21, l2 = p0 ? v12 : v22; // row l of V
        PCP_ROT(k0, l0);
        PCP_ROT(k1, l1);
        PCP_ROT(k2, l2);
        v00 = p2 ? v00 : k0, v01 = p2 ? v01 : k1, v02 = p2 ? v02 : k2;
        v10 = p0 ? l0 : (p2 ? k0 : v10), v11 = p0 ? l1 : (p2 ? k1 : v11), v12 = p0 ? l2 : (p2 ? k2 : v12);
        v20 = p0 ? v20 : l0, v21 = p0 ? v21 : l1, v22 = p0 ? v22 : l2;
        // OpenCV refreshes indR/indC only for idx in {k, l}
        if (k == 0) indR0 = (fabs(a01) < fabs(a02)) ? 2 : 1;
        if (l == 2) indC2 = (fabs(a02) < fabs(a12)) ? 1 : 0;
    }
    // selection sort, descending; rows of V follow
#define PCP_SWAPD(a, b)                                                                                      \
    {                                                                                                        \
        double t_ = a;                                                                                       \
        a = b;                                                                                               \
        b = t_;                                                                                              \
    }
    {
        int m = 0;
        if (w0 < w1) m = 1;
        if ((m == 0 ? w0 : w1) < w2) m = 2;
        if (m == 1) {
            PCP_SWAPD(w0, w1);
            PCP_SWAPD(v00, v10);
            PCP_SWAPD(v01, v11);
            PCP_SWAPD(v02, v12);
        } else if (m == 2) {
            PCP_SWAPD(w0, w2);
            PCP_SWAPD(v00, v20);
            PCP_SWAPD(v01, v21);
            PCP_SWAPD(v02, v22);
        }
        if (w1 < w2) {
            PCP_SWAPD(w1, w2);
            PCP_SWAPD(v10, v20);
            PCP_SWAPD(v11, v21);
            PCP_SWAPD(v12, v22);
        }
    }
#undef PCP_SWAPD
    W[0] = w0, W[1] = w1, W[2] = w2;
    V[0] = v00, V[1] = v01, V[2] = v02, V[3] = v10, V[4] = v11, V[5] = v12, V[6] = v20, V[7] = v21, V[8] = v22;
}
#undef PCP_ROT
static __device__ __noinline__ void pcp_jacobi3(double s00, double s01, double s02, double s11, double s12, double s22,
                                                double W[3], double V[9])
{
    pcp_jacobi3_impl(s00, s01, s02, s11, s12, s22, W, V);
}

// Everything after cvMatMul in calculate_plan_parameter_h_points (:165-199).
template <bool INLINE = false>
__device__ __forceinline__ void pcp_plane_from_scatter(double s00, double s01, double s02, double s11, double s12,
                                                       double s22, double XAve, double YAve, double ZAve,
                                                       pcp_plane& out)
{
    double W[3], V[9];
    if constexpr (INLINE) pcp_jacobi3_impl(s00, s01, s02, s11, s12, s22, W, V);
    else pcp_jacobi3(s00, s01, s02, s11, s12, s22, W, V);
    double valuemin = W[0], valuemax = W[0];
    int nummin = 0, nummax = 0;
#pragma unroll
    for (int i = 0; i < 3; i++) { // :166-179
        if (valuemin > W[i]) valuemin = W[i], nummin = i;
        if (valuemax < W[i]) valuemax = W[i], nummax = i;
    }
    double l1 = W[0], l2 = W[1], l3 = W[0]; // all-equal case: see oracle_feature.cpp
    if (nummin != nummax) {
        l3 = W[nummin];
        l1 = W[nummax];
        // lamada2 = the last index that is neither nummin nor nummax (:180-192)
        l2 = (nummin != 2 && nummax != 2) ? W[2] : ((nummin != 1 && nummax != 1) ? W[1] : W[0]);
    }
    double ex = nummin == 0 ? V[0] : (nummin == 1 ? V[3] : V[6]);
    double ey = nummin == 0 ? V[1] : (nummin == 1 ? V[4] : V[7]);
    double ez = nummin == 0 ? V[2] : (nummin == 1 ? V[5] : V[8]);
    out.nx = (float)ex; // PlanSegment members are float (data_struct.h:192-197)
    out.ny = (float)ey;
    out.nz = (float)ez;
    out.Distance = (float)(-((double)out.nx * XAve + (double)out.ny * YAve + (double)out.nz * ZAve));
    out.min_value = (float)l3;
    out.curvature = (float)(l3 / (l1 + l2 + l3));
    out.l1 = l1, out.l2 = l2, out.l3 = l3;
}

// LAS_POINT_PROPERTY record (48 bytes) written as three 16-byte stores.
__device__ __forceinline__ void pcp_store_props(void* props, size_t j, bool fitted, const pcp_plane& pl)
{
    float nx = 0.f, ny = 0.f, nz = 0.f;
    double dist = 0.0, curv = 1.0; // N <= 3 fallback, calculate_feature.cpp:356-363
    if (fitted) {
        nx = pl.nx, ny = pl.ny, nz = pl.nz;
        dist = (double)pl.Distance; // float member widened (:344-345)
        curv = (double)pl.curvature;
    }
    uint4* o = reinterpret_cast<uint4*>(static_cast<char*>(props) + j * 48);
    uint4 a, b, c;
    a.x = __float_as_uint(nx);
    a.y = __float_as_uint(ny);
    a.z = __float_as_uint(nz);
    a.w = 0u;
    unsigned long long db = (unsigned long long)__double_as_longlong(dist);
    unsigned long long cb = (unsigned long long)__double_as_longlong(curv);
    b.x = (unsigned)db;
    b.y = (unsigned)(db >> 32);
    b.z = (unsigned)cb;
    b.w = (unsigned)(cb >> 32);
    c.x = (unsigned)(int)j; // PointID
    c.y = 0u;               // SegmentID
    c.z = 0u;               // dis_from_point_plane
    c.w = 0u;
    o[0] = a;
    o[1] = b;
    o[2] = c;
}

__device__ __forceinline__ void pcp_store_extra(double* eig3, float* lps3, size_t j, bool fitted, const pcp_plane& pl)
{
    if (eig3) {
        eig3[3 * j + 0] = fitted ? pl.l1 : 0.0;
        eig3[3 * j + 1] = fitted ? pl.l2 : 0.0;
        eig3[3 * j + 2] = fitted ? pl.l3 : 0.0;
    }
    if (lps3) {
        lps3[3 * j + 0] = fitted ? (float)((pl.l1 - pl.l2) / pl.l1) : 0.f;
        lps3[3 * j + 1] = fitted ? (float)((pl.l2 - pl.l3) / pl.l1) : 0.f;
        lps3[3 * j + 2] = fitted ? (float)(pl.l3 / pl.l1) : 0.f;
    }
}
