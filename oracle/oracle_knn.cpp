// oracle_knn.cpp -- TEST INFRASTRUCTURE (see oracle.h).
// Exact kNN / radius search with KdTreeFLANN semantics (kd_tree.h:772-903, 928-956):
//   * non-finite points are dropped, results carry ORIGINAL cloud indices (index_mapping_, :837-842)
//   * k is clamped to the number of valid points (:820-821)
//   * squared L2 in double, accumulated x,y,z (FLANN L2_Simple: diff=a-b; result+=diff*diff), no FMA
//   * results ascending; ties ordered by original index (FLANN's own tie order is traversal-defined
//     and unpinned -- SURVEY 7.3-2; the harness reports rank-k boundary ties)
//   * radius: strict d2 < r*r (FLANN RadiusResultSet), sorted, capped at max_nn (0 = unlimited, :873-883)
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct Cand {
    double d2;
    int32_t idx;
};
inline bool cand_less(const Cand& a, const Cand& b) { return a.d2 < b.d2 || (a.d2 == b.d2 && a.idx < b.idx); }

inline const double* P(const void* base, size_t i, size_t stride)
{
    return reinterpret_cast<const double*>(static_cast<const char*>(base) + i * stride);
}
inline bool finite3(const double* p) { return std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]); }

inline double l2_simple(const double* a, const double* b)
{
    double result = 0;
    double diff = a[0] - b[0];
    result += diff * diff;
    diff = a[1] - b[1];
    result += diff * diff;
    diff = a[2] - b[2];
    result += diff * diff;
    return result;
}

// bounded max-heap on (d2, idx)
struct TopK {
    int k;
    std::vector<Cand> h;
    explicit TopK(int k_) : k(k_) { h.reserve(k_ > 0 ? k_ : 1); }
    inline bool full() const { return (int)h.size() >= k; }
    inline double worst() const { return full() ? h.front().d2 : std::numeric_limits<double>::infinity(); }
    inline void push(const Cand& c)
    {
        if (k <= 0) return;
        if (!full()) {
            h.push_back(c);
            std::push_heap(h.begin(), h.end(), cand_less);
        } else if (cand_less(c, h.front())) {
            std::pop_heap(h.begin(), h.end(), cand_less);
            h.back() = c;
            std::push_heap(h.begin(), h.end(), cand_less);
        }
    }
    void sorted(std::vector<Cand>& out)
    {
        out = h;
        std::sort(out.begin(), out.end(), cand_less);
    }
};

} // namespace

struct orc_kdtree {
    struct Node {
        int32_t left, right; // children, or -1
        int32_t begin, end;  // leaf range in perm
        int32_t dim;
        double split_lo, split_hi; // max of left child / min of right child along dim
    };
    std::vector<double> xyz;   // valid points, packed (kd_tree.h:928-956 cloud_)
    std::vector<int32_t> map;  // index_mapping_
    std::vector<int32_t> perm; // leaf ordering over valid ids
    std::vector<Node> nodes;
    double bb_lo[3], bb_hi[3];

    int32_t build(int32_t b, int32_t e)
    {
        Node nd;
        nd.left = nd.right = -1;
        nd.begin = b;
        nd.end = e;
        nd.dim = 0;
        nd.split_lo = nd.split_hi = 0;
        int32_t id = (int32_t)nodes.size();
        nodes.push_back(nd);
        if (e - b <= 15) return id; // KDTreeSingleIndexParams(15), kd_tree.h:795-797
        double lo[3], hi[3];
        for (int d = 0; d < 3; d++) lo[d] = std::numeric_limits<double>::max(), hi[d] = -lo[d];
        for (int32_t i = b; i < e; i++)
            for (int d = 0; d < 3; d++) {
                double v = xyz[3 * (size_t)perm[i] + d];
                lo[d] = std::min(lo[d], v);
                hi[d] = std::max(hi[d], v);
            }
        int dim = 0;
        for (int d = 1; d < 3; d++)
            if (hi[d] - lo[d] > hi[dim] - lo[dim]) dim = d;
        int32_t mid = b + (e - b) / 2;
        std::nth_element(perm.begin() + b, perm.begin() + mid, perm.begin() + e, [&](int32_t a, int32_t c) {
            double va = xyz[3 * (size_t)a + dim], vc = xyz[3 * (size_t)c + dim];
            return va < vc || (va == vc && a < c);
        });
        double slo = -std::numeric_limits<double>::max(), shi = std::numeric_limits<double>::max();
        for (int32_t i = b; i < mid; i++) slo = std::max(slo, xyz[3 * (size_t)perm[i] + dim]);
        for (int32_t i = mid; i < e; i++) shi = std::min(shi, xyz[3 * (size_t)perm[i] + dim]);
        int32_t l = build(b, mid);
        int32_t r = build(mid, e);
        nodes[id].left = l;
        nodes[id].right = r;
        nodes[id].dim = dim;
        nodes[id].split_lo = slo;
        nodes[id].split_hi = shi;
        return id;
    }

    // generic traversal; Visitor: bound() current pruning bound (d2), accept(d2, id)
    template <class V> void search(int32_t nid, const double* q, V& v) const
    {
        const Node& nd = nodes[nid];
        if (nd.left < 0) {
            for (int32_t i = nd.begin; i < nd.end; i++) {
                int32_t id = perm[i];
                v.accept(l2_simple(q, &xyz[3 * (size_t)id]), id);
            }
            return;
        }
        double qv = q[nd.dim];
        double dl = qv - nd.split_lo; // >0 => query right of every left point
        double dr = nd.split_hi - qv; // >0 => query left of every right point
        int32_t first, second;
        double gap;
        if (dl < dr) { // nearer to the left half
            first = nd.left;
            second = nd.right;
            gap = dr;
        } else {
            first = nd.right;
            second = nd.left;
            gap = dl;
        }
        search(first, q, v);
        // one-plane lower bound; fl(gap)^2 <= fl(q-p)^2 <= d2 by monotone rounding, so this never
        // prunes a point that could enter the result (ties use <=).
        if (gap <= 0 || gap * gap <= v.bound()) search(second, q, v);
    }
};

namespace {
struct KnnVisitor {
    TopK top;
    const int32_t* map;
    explicit KnnVisitor(int k, const int32_t* m) : top(k), map(m) {}
    inline double bound() const { return top.worst(); }
    inline void accept(double d2, int32_t id) { top.push(Cand{d2, map[id]}); }
};
struct RadiusVisitor {
    double r2;
    const int32_t* map;
    std::vector<Cand> got;
    inline double bound() const { return r2; }
    inline void accept(double d2, int32_t id)
    {
        if (d2 < r2) got.push_back(Cand{d2, map[id]});
    }
};
} // namespace

extern "C" {

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

size_t orc_valid_points(const void* pts, size_t n, size_t stride, int32_t* map)
{
    size_t m = 0;
    for (size_t i = 0; i < n; i++)
        if (finite3(P(pts, i, stride))) {
            if (map) map[m] = (int32_t)i;
            m++;
        }
    return m;
}

void orc_knn_brute(const void* pts, size_t n, size_t stride, const void* q, size_t nq, size_t qstride, int k,
                   int32_t* idx, double* d2, int32_t* found)
{
    std::vector<int32_t> map(n);
    size_t nv = orc_valid_points(pts, n, stride, map.data());
    int kk = (int)std::min<size_t>((size_t)std::max(k, 0), nv);
#pragma omp parallel for schedule(dynamic, 16)
    for (long long qi = 0; qi < (long long)nq; qi++) {
        const double* qp = P(q, qi, qstride);
        TopK top(kk);
        if (finite3(qp))
            for (size_t j = 0; j < nv; j++) top.push(Cand{l2_simple(qp, P(pts, map[j], stride)), map[j]});
        std::vector<Cand> s;
        top.sorted(s);
        for (int j = 0; j < k; j++) {
            bool ok = j < (int)s.size();
            idx[qi * k + j] = ok ? s[j].idx : -1;
            if (d2) d2[qi * k + j] = ok ? s[j].d2 : std::numeric_limits<double>::infinity();
        }
        if (found) found[qi] = (int32_t)s.size();
    }
}

orc_kdtree* orc_kdtree_build(const void* pts, size_t n, size_t stride)
{
    orc_kdtree* t = new orc_kdtree();
    t->map.reserve(n);
    t->xyz.reserve(3 * n);
    for (size_t i = 0; i < n; i++) {
        const double* p = P(pts, i, stride);
        if (!finite3(p)) continue;
        t->map.push_back((int32_t)i);
        t->xyz.push_back(p[0]);
        t->xyz.push_back(p[1]);
        t->xyz.push_back(p[2]);
    }
    t->perm.resize(t->map.size());
    std::iota(t->perm.begin(), t->perm.end(), 0);
    t->nodes.reserve(t->map.size() / 4 + 16);
    if (!t->perm.empty()) t->build(0, (int32_t)t->perm.size());
    return t;
}
void orc_kdtree_free(orc_kdtree* t) { delete t; }
size_t orc_kdtree_size(const orc_kdtree* t) { return t->map.size(); }

void orc_kdtree_knn(const orc_kdtree* t, const void* q, size_t nq, size_t qstride, int k, int32_t* idx,
                    double* d2, int32_t* found, int threads)
{
    int kk = (int)std::min<size_t>((size_t)std::max(k, 0), t->map.size());
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
    for (long long qi = 0; qi < (long long)nq; qi++) {
        const double* qp = P(q, qi, qstride);
        KnnVisitor v(kk, t->map.data());
        if (finite3(qp) && !t->nodes.empty() && kk > 0) t->search(0, qp, v);
        std::vector<Cand> s;
        v.top.sorted(s);
        for (int j = 0; j < k; j++) {
            bool ok = j < (int)s.size();
            idx[qi * k + j] = ok ? s[j].idx : -1;
            if (d2) d2[qi * k + j] = ok ? s[j].d2 : std::numeric_limits<double>::infinity();
        }
        if (found) found[qi] = (int32_t)s.size();
    }
}

static void radius_one(const orc_kdtree* t, const double* qp, double r, uint32_t max_nn, std::vector<Cand>& out)
{
    RadiusVisitor v;
    v.r2 = r * r; // kd_tree.h:888 passes radius*radius
    v.map = t->map.data();
    if (finite3(qp) && !t->nodes.empty()) t->search(0, qp, v);
    std::sort(v.got.begin(), v.got.end(), cand_less);
    size_t total = t->map.size();
    size_t cap = (max_nn == 0 || max_nn > total) ? total : max_nn; // :873-883
    if (v.got.size() > cap) v.got.resize(cap);
    out.swap(v.got);
}

void orc_kdtree_radius(const orc_kdtree* t, const void* q, size_t nq, size_t qstride, double r, uint32_t max_nn,
                       int64_t* offsets, int32_t* idx, double* d2, int threads)
{
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
    if (!idx) {
        offsets[0] = 0;
        std::vector<int64_t> cnt(nq);
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
        for (long long qi = 0; qi < (long long)nq; qi++) {
            std::vector<Cand> s;
            radius_one(t, P(q, qi, qstride), r, max_nn, s);
            cnt[qi] = (int64_t)s.size();
        }
        for (size_t i = 0; i < nq; i++) offsets[i + 1] = offsets[i] + cnt[i];
        return;
    }
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
    for (long long qi = 0; qi < (long long)nq; qi++) {
        std::vector<Cand> s;
        radius_one(t, P(q, qi, qstride), r, max_nn, s);
        int64_t o = offsets[qi];
        for (size_t j = 0; j < s.size() && o + (int64_t)j < offsets[qi + 1]; j++) {
            idx[o + j] = s[j].idx;
            if (d2) d2[o + j] = s[j].d2;
        }
    }
}

int64_t orc_count_boundary_ties(const orc_kdtree* t, const void* q, size_t nq, size_t qstride, int k, int threads)
{
    if ((size_t)k >= t->map.size() || k <= 0) return 0;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
    int64_t ties = 0;
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads) reduction(+ : ties)
    for (long long qi = 0; qi < (long long)nq; qi++) {
        const double* qp = P(q, qi, qstride);
        if (!finite3(qp)) continue;
        KnnVisitor v(k + 1, t->map.data());
        t->search(0, qp, v);
        std::vector<Cand> s;
        v.top.sorted(s);
        if ((int)s.size() == k + 1 && s[k].d2 == s[k - 1].d2) ties++;
    }
    return ties;
}

static void fill_props(orc_las_point_property* pr, int32_t j, const orc_plan_segment* pl)
{
    std::memset(pr, 0, sizeof(*pr));
    pr->PointID = j;
    if (pl) { // calculate_feature.cpp:340-346 (float members widened to double)
        pr->normal_x = pl->normal_x;
        pr->normal_y = pl->normal_y;
        pr->normal_z = pl->normal_z;
        pr->Distance = pl->Distance;
        pr->curvature = pl->curvature;
    } else { // :356-363
        pr->curvature = 1.0;
    }
}

static void fill_extra(double* eig3, float* lps3, size_t j, const orc_plan_segment* pl)
{
    if (eig3) {
        for (int c = 0; c < 3; c++) eig3[3 * j + c] = pl ? pl->eig[c] : 0.0;
    }
    if (lps3) { // build-defined: L=(l1-l2)/l1, P=(l2-l3)/l1, S=l3/l1
        if (pl) {
            double l1 = pl->eig[0], l2 = pl->eig[1], l3 = pl->eig[2];
            lps3[3 * j + 0] = (float)((l1 - l2) / l1);
            lps3[3 * j + 1] = (float)((l2 - l3) / l1);
            lps3[3 * j + 2] = (float)(l3 / l1);
        } else {
            lps3[3 * j + 0] = lps3[3 * j + 1] = lps3[3 * j + 2] = 0.f;
        }
    }
}

void orc_features_knn(const void* pts, size_t n, size_t stride, int k, orc_las_point_property* props, double* eig3,
                      float* lps3, int32_t* knn_idx, int threads)
{
    orc_kdtree* t = orc_kdtree_build(pts, n, stride); // calculate_feature.cpp:216-218
    int kk = (int)std::min<size_t>((size_t)std::max(k, 0), t->map.size());
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads) // calculate_feature.cpp:222
    for (long long j = 0; j < (long long)n; j++) {
        const double* qp = P(pts, j, stride);
        std::vector<Cand> s;
        if (finite3(qp) && kk > 0) {
            KnnVisitor v(kk, t->map.data());
            t->search(0, qp, v);
            v.top.sorted(s);
        }
        int N = (int)s.size();
        if (knn_idx)
            for (int c = 0; c < k; c++) knn_idx[j * k + c] = c < N ? s[c].idx : -1;
        if (N > 3) { // :238
            std::vector<double> xyz(3 * (size_t)N);
            for (int c = 0; c < N; c++) std::memcpy(&xyz[3 * c], P(pts, s[c].idx, stride), 24);
            orc_plan_segment pl;
            orc_plane_fit(xyz.data(), N, &pl);
            fill_props(&props[j], (int32_t)j, &pl);
            fill_extra(eig3, lps3, j, &pl);
        } else {
            fill_props(&props[j], (int32_t)j, nullptr);
            fill_extra(eig3, lps3, j, nullptr);
        }
    }
    orc_kdtree_free(t);
}

// ---- calculate_plan_parameter_rpca, calculate_feature.cpp:208-368, with a counter-based generator -----------------
// The reference draws its 3-point hypotheses with rand() under OpenMP (srand(time)): not reproducible run to run.
// Here draw n of hypothesis i of point j is splitmix64(seed, j, i, n) % N; everything else follows the reference:
// float-truncated neighbour coordinates in the h-subset (LAS_POINT_PROPERTY_sim is float, data_struct.h:181-187),
// float point-to-plane distances computed from float-truncated coordinates (extraction_tree.cpp:47-64), MAD test.
// Deviations (stated): the two std::sort calls are unstable in the reference -- ties here keep neighbour order;
// the best hypothesis is the first with the smallest lambda3; no valid hypothesis (every draw repeated an index,
// probability ~1e-28 for N = 20) gives the unfitted record instead of planes[0] on an empty vector.
static inline uint64_t rpca_mix(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
static inline int rpca_draw(uint64_t seed, uint64_t j, int i, int n, int N)
{
    return (int)((rpca_mix(seed * 0x9E3779B97F4A7C15ull + j * 0xD1B54A32D192ED03ull + (uint64_t)(i * 3 + n)) >> 33) % (uint64_t)N);
}
// TreeExtration::compute_distance_from_point_to_plane (extraction_tree.cpp:47-64): oracle_segment.cpp
static inline float rpca_plane_dist(const double* p, float a, float b, float c, float d)
{
    return orc_point_plane_distance(p, a, b, c, d);
}

extern "C" int orc_rpca_iterations(float Pr, float epi)
{
    // compute_iteration_number (calculate_feature.cpp:28-33): log10 of a float, pow(float, int) in double
    return (int)(std::log10(1 - Pr) / std::log10(1 - std::pow((double)(1 - epi), 3)));
}

extern "C" int orc_rpca_point(const double* nb, int N, int iterations, uint64_t seed, uint64_t j, orc_plan_segment* out)
{
    if (N <= 3) return 0; // :238
    const int h_free = (int)(2.0 / 3 * N);
    orc_plan_segment best;
    bool have = false;
    std::vector<float> dis(N);
    std::vector<int> order(N);
    std::vector<double> sub(3 * (size_t)std::max(h_free, 3));
    for (int i = 0; i < iterations; i++) {
        int num[3];
        for (int n = 0; n < 3; n++) num[n] = rpca_draw(seed, j, i, n, N);
        if (num[0] == num[1] || num[0] == num[2] || num[1] == num[2]) continue;
        double p3[9];
        for (int n = 0; n < 3; n++) std::memcpy(&p3[3 * n], &nb[3 * num[n]], 24);
        orc_plan_segment a;
        orc_plane_fit(p3, 3, &a); // calculate_plan_parameter_3points: the same arithmetic with h = 3
        for (int n = 0; n < N; n++) {
            dis[n] = rpca_plane_dist(&nb[3 * n], a.normal_x, a.normal_y, a.normal_z, a.Distance);
            order[n] = n;
        }
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return dis[x] < dis[y]; });
        for (int n = 0; n < h_free; n++)
            for (int c = 0; c < 3; c++) sub[3 * n + c] = (double)(float)nb[3 * order[n] + c]; // float members
        orc_plane_fit(sub.data(), h_free, &a);
        if (!have || a.min_value < best.min_value) best = a, have = true;
    }
    if (!have) return 0;
    std::vector<float> d(N), t(N);
    for (int n = 0; n < N; n++) d[n] = rpca_plane_dist(&nb[3 * n], best.normal_x, best.normal_y, best.normal_z, best.Distance);
    std::vector<float> srt(d);
    std::sort(srt.begin(), srt.end());
    const float median = srt[N / 2];
    for (int n = 0; n < N; n++) t[n] = std::fabs(d[n] - median);
    srt = t;
    std::sort(srt.begin(), srt.end());
    const float MAD = (float)(1.4826 * srt[N / 2]);
    std::vector<double> fin;
    for (int n = 0; n < N; n++) {
        bool keep = true;
        if (MAD != 0) {
            const float Rz = t[n] / MAD;
            keep = Rz < 2.5;
        }
        if (keep) fin.insert(fin.end(), &nb[3 * n], &nb[3 * n] + 3);
    }
    const int m = (int)(fin.size() / 3);
    if (m <= 3) return 0;
    orc_plane_fit(fin.data(), m, out);
    return 1;
}

void orc_features_rpca(const void* pts, size_t n, size_t stride, int k, float Pr, float epi, uint64_t seed,
                       orc_las_point_property* props, int threads)
{
    orc_kdtree* t = orc_kdtree_build(pts, n, stride);
    int kk = (int)std::min<size_t>((size_t)std::max(k, 0), t->map.size());
    const int iterations = orc_rpca_iterations(Pr, epi);
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
    for (long long j = 0; j < (long long)n; j++) {
        const double* qp = P(pts, j, stride);
        std::vector<Cand> s;
        if (finite3(qp) && kk > 0) {
            KnnVisitor v(kk, t->map.data());
            t->search(0, qp, v);
            v.top.sorted(s);
        }
        const int N = (int)s.size();
        std::vector<double> xyz(3 * (size_t)std::max(N, 1));
        for (int c = 0; c < N; c++) std::memcpy(&xyz[3 * c], P(pts, s[c].idx, stride), 24);
        orc_plan_segment pl;
        const int fitted = orc_rpca_point(xyz.data(), N, iterations, seed, (uint64_t)j, &pl);
        fill_props(&props[j], (int32_t)j, fitted ? &pl : nullptr);
    }
    orc_kdtree_free(t);
}

void orc_features_radius(const void* pts, size_t n, size_t stride, double r, uint32_t max_nn,
                         orc_las_point_property* props, double* eig3, float* lps3, int32_t* counts, int threads)
{
    orc_kdtree* t = orc_kdtree_build(pts, n, stride);
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
    for (long long j = 0; j < (long long)n; j++) {
        std::vector<Cand> s;
        radius_one(t, P(pts, j, stride), r, max_nn, s);
        int N = (int)s.size();
        if (counts) counts[j] = N;
        if (N > 3) {
            std::vector<double> xyz(3 * (size_t)N);
            for (int c = 0; c < N; c++) std::memcpy(&xyz[3 * c], P(pts, s[c].idx, stride), 24);
            orc_plan_segment pl;
            orc_plane_fit(xyz.data(), N, &pl);
            fill_props(&props[j], (int32_t)j, &pl);
            fill_extra(eig3, lps3, j, &pl);
        } else {
            fill_props(&props[j], (int32_t)j, nullptr);
            fill_extra(eig3, lps3, j, nullptr);
        }
    }
    orc_kdtree_free(t);
}

} // extern "C"
