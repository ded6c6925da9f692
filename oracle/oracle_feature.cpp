// oracle_feature.cpp -- TEST INFRASTRUCTURE (see oracle.h).  Plane fit = calculate_feature.cpp:119-206.
// The eigen-solver restates OpenCV 2.4.x lapack.cpp JacobiImpl_<double> (the routine behind cvEigenVV,
// calculate_feature.cpp:165); OpenCV is not vendored in /root/reference, so this is a restatement of
// its published algorithm, pinned by tests/golden/eigen_cv2.json (cv2.eigen outputs).
// Compile with -ffp-contract=off: every product and sum below is rounded separately.
#include "oracle.h"
#include <cmath>
#include <limits>
#include <utility>
#include <vector>

namespace {

// OpenCV's local hypot (lapack.cpp): scaled sqrt, only + * / sqrt.
inline double cv_hypot(double a, double b)
{
    a = std::fabs(a);
    b = std::fabs(b);
    if (a > b) {
        b /= a;
        return a * std::sqrt(1 + b * b);
    }
    if (b > 0) {
        a /= b;
        return b * std::sqrt(1 + a * a);
    }
    return 0;
}

} // namespace

extern "C" void orc_jacobi(double* A, int n, double* W, double* V)
{
    const double eps = std::numeric_limits<double>::epsilon();
    const int astep = n, vstep = n;
    int i, j, k, m;
    for (i = 0; i < n; i++) {
        for (j = 0; j < n; j++) V[i * vstep + j] = 0;
        V[i * vstep + i] = 1;
    }
    int iters, maxIters = n * n * 30;
    std::vector<int> indR(n), indC(n);
    double mv = 0;
    for (k = 0; k < n; k++) {
        W[k] = A[(astep + 1) * k];
        if (k < n - 1) {
            for (m = k + 1, mv = std::fabs(A[astep * k + m]), i = k + 2; i < n; i++) {
                double val = std::fabs(A[astep * k + i]);
                if (mv < val) mv = val, m = i;
            }
            indR[k] = m;
        }
        if (k > 0) {
            for (m = 0, mv = std::fabs(A[k]), i = 1; i < k; i++) {
                double val = std::fabs(A[astep * i + k]);
                if (mv < val) mv = val, m = i;
            }
            indC[k] = m;
        }
    }
    if (n > 1)
        for (iters = 0; iters < maxIters; iters++) {
            // pivot = largest |off-diagonal| of the upper triangle
            for (k = 0, mv = std::fabs(A[indR[0]]), i = 1; i < n - 1; i++) {
                double val = std::fabs(A[astep * i + indR[i]]);
                if (mv < val) mv = val, k = i;
            }
            int l = indR[k];
            for (i = 1; i < n; i++) {
                double val = std::fabs(A[astep * indC[i] + i]);
                if (mv < val) mv = val, k = indC[i], l = i;
            }
            double p = A[astep * k + l];
            if (std::fabs(p) <= eps) break;
            double y = (W[l] - W[k]) * 0.5;
            double t = std::fabs(y) + cv_hypot(p, y);
            double s = cv_hypot(p, t);
            double c = t / s;
            s = p / s;
            t = (p / t) * p;
            if (y < 0) s = -s, t = -t;
            A[astep * k + l] = 0;
            W[k] -= t;
            W[l] += t;
            double a0, b0;
#define ORC_ROTATE(v0, v1) a0 = v0, b0 = v1, v0 = a0 * c - b0 * s, v1 = a0 * s + b0 * c
            for (i = 0; i < k; i++) ORC_ROTATE(A[astep * i + k], A[astep * i + l]);
            for (i = k + 1; i < l; i++) ORC_ROTATE(A[astep * k + i], A[astep * i + l]);
            for (i = l + 1; i < n; i++) ORC_ROTATE(A[astep * k + i], A[astep * l + i]);
            for (i = 0; i < n; i++) ORC_ROTATE(V[vstep * k + i], V[vstep * l + i]);
#undef ORC_ROTATE
            for (j = 0; j < 2; j++) {
                int idx = j == 0 ? k : l;
                if (idx < n - 1) {
                    for (m = idx + 1, mv = std::fabs(A[astep * idx + m]), i = idx + 2; i < n; i++) {
                        double val = std::fabs(A[astep * idx + i]);
                        if (mv < val) mv = val, m = i;
                    }
                    indR[idx] = m;
                }
                if (idx > 0) {
                    for (m = 0, mv = std::fabs(A[idx]), i = 1; i < idx; i++) {
                        double val = std::fabs(A[astep * i + idx]);
                        if (mv < val) mv = val, m = i;
                    }
                    indC[idx] = m;
                }
            }
        }
    // descending eigenvalues, eigenvectors as rows
    for (k = 0; k < n - 1; k++) {
        m = k;
        for (i = k + 1; i < n; i++)
            if (W[m] < W[i]) m = i;
        if (k != m) {
            std::swap(W[m], W[k]);
            for (i = 0; i < n; i++) std::swap(V[vstep * m + i], V[vstep * k + i]);
        }
    }
}

// calculate_feature.cpp:119-206.  xyz = h packed triples, in the order the reference would hold them
// in cloud->points (for the kNN feature path: ascending (d2, index), self first).
extern "C" void orc_plane_fit(const double* xyz, int h, orc_plan_segment* out)
{
    double XAve = 0.0, YAve = 0.0, ZAve = 0.0;
    for (int i = 0; i < h; i++) { // :131-139
        XAve += xyz[3 * i + 0];
        YAve += xyz[3 * i + 1];
        ZAve += xyz[3 * i + 2];
    }
    XAve /= h; // :140-142
    YAve /= h;
    ZAve /= h;
    // X = A - mean (3 x h), XXT = X * X^T, un-normalised (:150-164); cvMatMul summed over the h columns
    // in order.
    std::vector<double> X(3 * (size_t)h);
    for (int n = 0; n < h; n++) {
        X[0 * h + n] = xyz[3 * n + 0] - XAve;
        X[1 * h + n] = xyz[3 * n + 1] - YAve;
        X[2 * h + n] = xyz[3 * n + 2] - ZAve;
    }
    double S[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int n = 0; n < h; n++) s += X[i * h + n] * X[j * h + n];
            S[i * 3 + j] = s;
        }
    double W[3], V[9];
    orc_jacobi(S, 3, W, V); // cvEigenVV(XXT, E, I) :165
    double valuemin = W[0], valuemax = W[0];
    int nummin = 0, nummax = 0;
    for (int i = 0; i < 3; i++) { // :166-179
        if (valuemin > W[i]) {
            valuemin = W[i];
            nummin = i;
        }
        if (valuemax < W[i]) {
            valuemax = W[i];
            nummax = i;
        }
    }
    double lamada1 = 0, lamada2 = 0, lamada3 = 0;
    for (int i = 0; i < 3; i++) { // :180-192 (when all three are equal nummin==nummax==0: lamada1 is
                                  // never assigned in the reference (UB); we take it as W[0])
        if (i == nummin) {
            lamada3 = W[nummin];
        } else if (i == nummax) {
            lamada1 = W[nummax];
        } else {
            lamada2 = W[i];
        }
    }
    if (nummin == nummax) lamada1 = W[0], lamada2 = W[1], lamada3 = W[0];
    out->normal_x = (float)V[nummin * 3 + 0]; // PlanSegment members are float, data_struct.h:192-197
    out->normal_y = (float)V[nummin * 3 + 1];
    out->normal_z = (float)V[nummin * 3 + 2];
    out->Distance = (float)(-(out->normal_x * XAve + out->normal_y * YAve + out->normal_z * ZAve)); // :197
    out->min_value = (float)lamada3;
    out->curvature = (float)(lamada3 / (lamada1 + lamada2 + lamada3));
    out->eig[0] = lamada1;
    out->eig[1] = lamada2;
    out->eig[2] = lamada3;
}
