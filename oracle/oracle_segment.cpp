// oracle_segment.cpp -- TEST INFRASTRUCTURE ONLY (see oracle.h).
// CPU restatement of TreeExtration::region_growning, /root/reference/extraction_tree.cpp:170-272 (the live
// definition; :66-168 is the commented-out older one), with its two helpers
// compute_included_angle_between_vector (:25-33) and compute_distance_from_point_to_plane (:47-64).
//
// The reference runs one nearestKSearch(point, 50) per popped seed (:222).  The neighbour lists do not depend on the
// segmentation state, so this restatement computes all of them first (OpenMP over the oracle's leaf-15 kd-tree,
// ascending (d2, index) like everything else here) and then grows the regions strictly in the reference's order:
//   seeds     the lowest index still in UnSegment (std::set::begin, :191-192); a seed with curvature >= curv_t is
//             only erased (:268-269);
//   growth    FIFO over the region (:203-247): for each popped point p, its neighbours j in ascending distance that
//             are still UNSEGMENTATION join when |cos(normal of the FIRST seed, normal of j)| > cosfaT and the
//             distance of j to the plane of p is < distanceT (:233); joining erases j from UnSegment;
//   verdict   more than min_points members -> a PlanSegment with the first seed's normal and Distance, label++;
//             otherwise the members go back to UNSEGMENTATION (:257-260) but stay erased from UnSegment, so they
//             can be absorbed by a later region and never seed one.
#include "oracle.h"

#include <cmath>
#include <cstring>
#include <vector>

namespace {

inline const double* P(const void* base, size_t i, size_t stride)
{
    return reinterpret_cast<const double*>(static_cast<const char*>(base) + i * stride);
}

// extraction_tree.cpp:25-33: float throughout (MSVC's math.h gives sqrt(float) a float overload)
inline float included_angle(float vx1, float vy1, float vz1, float vx2, float vy2, float vz2)
{
    const float n_n1 = vx1 * vx2 + vy1 * vy2 + vz1 * vz2;
    const float n_n = std::sqrt(vx1 * vx1 + vy1 * vy1 + vz1 * vz1);
    const float n1_n1 = std::sqrt(vx2 * vx2 + vy2 * vy2 + vz2 * vz2);
    return std::fabs(n_n1 / (n_n * n1_n1));
}

} // namespace

// extraction_tree.cpp:47-64: coordinates truncated to float, float products, double sum, float result.  Shared
// with the RPCA restatement (oracle_knn.cpp).
extern "C" float orc_point_plane_distance(const double* p, float a, float b, float c, float d)
{
    const float x1 = (float)p[0], y1 = (float)p[1], z1 = (float)p[2];
    const double g = std::sqrt(a * a + b * b + c * c); // float argument -> float overload, widened (:57)
    const double f1 = a * x1, f2 = b * y1, f3 = c * z1, f4 = d;
    const double f = std::fabs(f1 + f2 + f3 + f4);
    return (float)(f / g);
}

extern "C" int64_t orc_region_growing(const void* pts, size_t n, size_t stride, orc_las_point_property* props, int k,
                                      double distanceT, double cosfaT, double curv_t, int min_points,
                                      int64_t* seg_offsets, int32_t* seg_points, float* seg_plane, int threads)
{
    std::vector<int32_t> nbr((size_t)n * (size_t)std::max(k, 1));
    {
        orc_kdtree* t = orc_kdtree_build(pts, n, stride);
        orc_kdtree_knn(t, pts, n, stride, k, nbr.data(), nullptr, nullptr, threads);
        orc_kdtree_free(t);
    }
    std::vector<unsigned char> unseg(n, 1); // the std::set UnSegment (:178-187)
    for (size_t i = 0; i < n; i++) props[i].SegmentID = -1;
    std::vector<int32_t> region;
    int64_t nseg = 0, total = 0;
    int labelOfSeg = 0;
    seg_offsets[0] = 0;
    for (size_t minNum = 0; minNum < n; minNum++) { // :190-192: begin() of an ordered set that only shrinks
        if (!unseg[minNum]) continue;
        unseg[minNum] = 0;
        if (!(props[minNum].curvature < curv_t)) continue; // :194, :268-269
        props[minNum].SegmentID = labelOfSeg;
        region.clear();
        region.push_back((int32_t)minNum);
        const double nx = props[minNum].normal_x, ny = props[minNum].normal_y, nz = props[minNum].normal_z;
        const double nd = props[minNum].Distance;
        for (size_t head = 0; head < region.size(); head++) { // the deque `seed`
            const int32_t point_id = region[head];
            const double nx_seed = props[point_id].normal_x, ny_seed = props[point_id].normal_y,
                         nz_seed = props[point_id].normal_z, d_seed = props[point_id].Distance;
            const int32_t row = props[point_id].PointID; // :220-221: the query is points[PointProperty[point_id].PointID]
            if (row < 0 || (size_t)row >= n) continue;
            for (int e = 0; e < k; e++) {
                const int32_t j = nbr[(size_t)row * k + e];
                if (j < 0) break;
                if (props[j].SegmentID != -1) continue; // :225
                const float cosn = included_angle((float)nx, (float)ny, (float)nz, props[j].normal_x,
                                                  props[j].normal_y, props[j].normal_z);
                const float dis = orc_point_plane_distance(P(pts, (size_t)j, stride), (float)nx_seed, (float)ny_seed,
                                                           (float)nz_seed, (float)d_seed);
                if (cosn > cosfaT && dis < distanceT) { // :233
                    props[j].SegmentID = labelOfSeg;
                    region.push_back(j);
                    unseg[j] = 0;
                }
            }
        }
        if ((int64_t)region.size() > min_points) { // :249-256
            std::memcpy(seg_points + total, region.data(), region.size() * sizeof(int32_t));
            total += (int64_t)region.size();
            seg_plane[4 * nseg + 0] = (float)nx;
            seg_plane[4 * nseg + 1] = (float)ny;
            seg_plane[4 * nseg + 2] = (float)nz;
            seg_plane[4 * nseg + 3] = (float)nd;
            nseg++;
            seg_offsets[nseg] = total;
            labelOfSeg++;
        } else {
            for (int32_t j : region) props[j].SegmentID = -1; // :257-260
        }
    }
    return nseg;
}
