// oracle_voxel.cpp -- TEST INFRASTRUCTURE (see oracle.h).
// VoxelGrid<PointXYZRGBA>::applyFilter (voxel_grid.h:811-1056, no field filter) and
// PointCloudHelper::remove_duplicate (point_cloud_helper.cpp:42-63) restated.
// Kept quirks: truncating static_cast<int> cell index (:835-840,:936-938), max seeded with DBL_MIN
// (point_cloud_helper.h:65), every field rounded through float before the double sum (concatenate.h:153),
// static_cast<T> write-back (concatenate.h:118), rgb re-pack with alpha 0 (voxel_grid.h:1045-1050),
// float leaf (point_cloud_helper.h:16).  Documented deviations: 64-bit cell key when the reference's
// int idx would overflow (undefined behaviour there); stable order inside a cell (std::sort is unstable).
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

namespace {
inline const char* R(const void* base, size_t i, size_t stride) { return static_cast<const char*>(base) + i * stride; }
inline char* Rw(void* base, size_t i, size_t stride) { return static_cast<char*>(base) + i * stride; }
inline bool finite3(const double* p) { return std::isfinite(p[0]) && std::isfinite(p[1]) && std::isfinite(p[2]); }
struct KeyIdx {
    uint64_t key;
    uint32_t pt;
};
} // namespace

extern "C" {

void orc_minmax(const void* pts, size_t n, size_t stride, int is_dense, double mn[3], double mx[3])
{
    for (int d = 0; d < 3; d++) {
        mn[d] = std::numeric_limits<double>::max();
        mx[d] = std::numeric_limits<double>::min(); // sic: smallest positive, point_cloud_helper.h:65
    }
    for (size_t i = 0; i < n; i++) {
        const double* p = reinterpret_cast<const double*>(R(pts, i, stride));
        if (!is_dense && !finite3(p)) continue;
        for (int d = 0; d < 3; d++) {
            mn[d] = std::min(mn[d], p[d]); // Eigen array min/max
            mx[d] = std::max(mx[d], p[d]);
        }
    }
}

void orc_centroid(const void* pts, size_t n, size_t stride, int is_dense, double c[3], uint32_t* count)
{
    c[0] = c[1] = c[2] = 0;
    uint32_t cp = 0;
    for (size_t i = 0; i < n; i++) { // serial-order sum, point_cloud_helper.h:203-226
        const double* p = reinterpret_cast<const double*>(R(pts, i, stride));
        if (!is_dense && !finite3(p)) continue;
        c[0] += p[0];
        c[1] += p[1];
        c[2] += p[2];
        ++cp;
    }
    if (cp) {
        double dn = static_cast<double>(cp);
        c[0] /= dn;
        c[1] /= dn;
        c[2] /= dn;
    }
    if (count) *count = cp;
}

int64_t orc_voxel_grid(const void* pts, size_t n, size_t stride, const double leaf[3], int downsample_all,
                       int is_dense, void* out_pts, int32_t min_b_out[3], int32_t div_b_out[3], int32_t* point_cell,
                       int32_t* cell_count, int* overflow32)
{
    if (overflow32) *overflow32 = 0;
    if (n == 0) {
        for (int d = 0; d < 3; d++) {
            if (min_b_out) min_b_out[d] = 0;
            if (div_b_out) div_b_out[d] = 0;
        }
        return 0;
    }
    double inv[3] = {1.0 / leaf[0], 1.0 / leaf[1], 1.0 / leaf[2]}; // setLeafSize :530-548
    double mn[3], mx[3];
    orc_minmax(pts, n, stride, is_dense, mn, mx);
    int min_b[3], max_b[3], div_b[3];
    for (int d = 0; d < 3; d++) { // :835-843
        min_b[d] = static_cast<int>(double(mn[d] * inv[d]));
        max_b[d] = static_cast<int>(double(mx[d] * inv[d]));
        div_b[d] = max_b[d] - min_b[d] + 1;
        if (min_b_out) min_b_out[d] = min_b[d];
        if (div_b_out) div_b_out[d] = div_b[d];
    }
    // the reference's idx is int: ijk0 + ijk1*div_x + ijk2*div_x*div_y (:941)
    long double cells = (long double)div_b[0] * div_b[1] * div_b[2];
    if (cells > 2147483647.0L && overflow32) *overflow32 = 1;

    std::vector<KeyIdx> iv;
    iv.reserve(n);
    for (size_t cp = 0; cp < n; ++cp) { // first pass :928-943
        const double* p = reinterpret_cast<const double*>(R(pts, cp, stride));
        if (!finite3(p)) {
            // reference skips only when !is_dense (:929-934); a NaN with is_dense=true is UB there.
            if (point_cell) point_cell[3 * cp] = point_cell[3 * cp + 1] = point_cell[3 * cp + 2] = -1;
            continue;
        }
        int ijk0 = static_cast<int>(double(p[0] * inv[0]) - min_b[0]);
        int ijk1 = static_cast<int>(double(p[1] * inv[1]) - min_b[1]);
        int ijk2 = static_cast<int>(double(p[2] * inv[2]) - min_b[2]);
        if (point_cell) {
            point_cell[3 * cp + 0] = ijk0;
            point_cell[3 * cp + 1] = ijk1;
            point_cell[3 * cp + 2] = ijk2;
        }
        uint64_t key = (uint64_t)ijk0 + (uint64_t)ijk1 * (uint64_t)div_b[0] +
                       (uint64_t)ijk2 * (uint64_t)div_b[0] * (uint64_t)div_b[1];
        iv.push_back(KeyIdx{key, (uint32_t)cp});
    }
    // second pass :948 (std::sort, unstable; stable here = ascending point index inside a cell)
    std::stable_sort(iv.begin(), iv.end(), [](const KeyIdx& a, const KeyIdx& b) { return a.key < b.key; });

    // fourth pass :986-1054.  FieldList of PointXYZRGBA = x,y,z,rgba,stamp_id (point_type.h:326-332):
    // centroid_size = 5 + 3 (rgb special case :862-874)
    int64_t index = 0;
    for (size_t cp = 0; cp < iv.size();) {
        size_t i = cp;
        double c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        bool first = true;
        while (i < iv.size() && iv[i].key == iv[cp].key) {
            const char* rec = R(pts, iv[i].pt, stride);
            double t[8];
            const double* p = reinterpret_cast<const double*>(rec);
            if (!downsample_all) { // :991-994, :1010-1013 full double xyz
                t[0] = p[0];
                t[1] = p[1];
                t[2] = p[2];
                t[3] = t[4] = t[5] = t[6] = t[7] = 0;
            } else {
                uint32_t rgba, stamp;
                std::memcpy(&rgba, rec + 32, 4);
                std::memcpy(&stamp, rec + 36, 4);
                t[5] = (double)((rgba >> 16) & 0xff); // r  (BGRA byte order, point_type.h:66-71)
                t[6] = (double)((rgba >> 8) & 0xff);  // g
                t[7] = (double)(rgba & 0xff);         // b
                t[0] = static_cast<float>(p[0]);      // concatenate.h:153
                t[1] = static_cast<float>(p[1]);
                t[2] = static_cast<float>(p[2]);
                t[3] = static_cast<float>(rgba);
                t[4] = static_cast<float>(stamp);
            }
            if (first) {
                for (int f = 0; f < 8; f++) c[f] = t[f];
                first = false;
            } else {
                for (int f = 0; f < 8; f++) c[f] += t[f];
            }
            ++i;
        }
        double cnt = static_cast<double>(i - cp);
        for (int f = 0; f < 8; f++) c[f] /= cnt; // :1034
        char* out = Rw(out_pts, (size_t)index, stride);
        if (!downsample_all) {
            // output.points.resize(total) default-constructs: x=y=z=0, data[3]=1, rgba=0 (point_type.h:86-91);
            // stamp_id is left uninitialised by the reference constructor, 0 here.
            std::memset(out, 0, stride < 48 ? stride : 48);
            double* o = reinterpret_cast<double*>(out);
            o[0] = c[0];
            o[1] = c[1];
            o[2] = c[2];
            if (stride >= 32) o[3] = 1.0;
        } else {
            std::memset(out, 0, stride < 48 ? stride : 48);
            double* o = reinterpret_cast<double*>(out);
            o[0] = static_cast<double>(c[0]); // concatenate.h:118
            o[1] = static_cast<double>(c[1]);
            o[2] = static_cast<double>(c[2]);
            if (stride >= 32) o[3] = 1.0;
            uint32_t stamp = static_cast<uint32_t>(c[4]);
            float r = (float)c[5], g = (float)c[6], b = (float)c[7]; // :1047-1049
            int rgb = (static_cast<int>(r) << 16) | (static_cast<int>(g) << 8) | static_cast<int>(b);
            if (stride >= 40) {
                std::memcpy(out + 32, &rgb, 4);
                std::memcpy(out + 36, &stamp, 4);
            }
        }
        if (cell_count) cell_count[index] = (int32_t)(i - cp);
        cp = i;
        ++index;
    }
    return index;
}

int64_t orc_remove_duplicate(void* pts, size_t n, size_t stride, float leaf, double* center3)
{
    if (n == 0) return 0;
    double c[3];
    // compute3DCentroid honours is_dense; the shim always treats clouds as possibly non-dense
    orc_centroid(pts, n, stride, 0, c, nullptr);
    if (center3) std::memcpy(center3, c, 24);
    std::vector<char> tmp(n * stride);
    std::memcpy(tmp.data(), pts, n * stride); // copyPointCloud :50
    for (size_t i = 0; i < n; i++) {          // transformPointCloud with R=I, t=-center :52
        double* p = reinterpret_cast<double*>(tmp.data() + i * stride);
        if (!finite3(p)) continue;
        p[0] = p[0] + (-c[0]);
        p[1] = p[1] + (-c[1]);
        p[2] = p[2] + (-c[2]);
    }
    double l = leaf; // const float -> setLeafSize(double...) :57
    double leaf3[3] = {l, l, l};
    int64_t m = orc_voxel_grid(tmp.data(), n, stride, leaf3, 1, 0, pts, nullptr, nullptr, nullptr, nullptr, nullptr);
    for (int64_t i = 0; i < m; i++) { // :60-61
        double* p = reinterpret_cast<double*>(Rw(pts, (size_t)i, stride));
        p[0] = p[0] + c[0];
        p[1] = p[1] + c[1];
        p[2] = p[2] + c[2];
    }
    return m;
}

} // extern "C"
