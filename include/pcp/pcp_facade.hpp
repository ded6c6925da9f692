// pcp_facade.hpp -- source-compatible C++ facade over the C ABI (include/pcp_cuda.h) for the call shapes the
// reference's own callers use.  std-only: no Boost / Eigen / FLANN / OpenCV.
//
//   cloud_blend_double::PointXYZRGBA, PointCloud<PointT>      point_type.h:9-82, point_cloud.h:54-535
//   KdTreeFLANN<PointT>::setInputCloud / nearestKSearch / radiusSearch        kd_tree.h:772-903
//   VoxelGrid<PointT>::setInputCloud / setLeafSize / filter                   voxel_grid.h:482-549, 429-448
//   PointCloudHelper::remove_duplicate / getMinMax3D / compute3DCentroid      point_cloud_helper.h:16-90,192-230
//   CalculateFeature::calculate_plan_parameter[_rpca|_h_points]               calculate_feature.h:11-16
//   TreeExtration::region_growning, point_segment                             extraction_tree.cpp:170-272, static.cpp:8-21
//
// Every method computes on the GPU through libpcp_cuda.so; there is no CPU path behind this header.  Errors of
// the library surface as std::runtime_error (the reference's own error style -- cout + return code -- is kept
// for the documented cases: filter() without input prints and returns an empty cloud, voxel_grid.h:815-821).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../pcp_cuda.h"

namespace cloud_blend_double {

// point_type.h:9-82: 48 bytes, 16-aligned
struct alignas(16) PointXYZRGBA {
    union {
        double data[4];
        struct {
            double x, y, z;
        };
    };
    union {
        struct {
            uint8_t b, g, r, a;
        };
        uint32_t rgba;
    };
    uint32_t stamp_id;
    PointXYZRGBA() : rgba(0), stamp_id(0)
    {
        x = y = z = 0.0;
        data[3] = 1.0;
    }
};
static_assert(sizeof(PointXYZRGBA) == 48, "PointXYZRGBA must match point_type.h");

// data_struct.h:161-173 and :188-198
typedef pcp_point_property LAS_POINT_PROPERTY;
struct PlanSegment {
    int SegmentID = 0;
    float normal_x = 0, normal_y = 0, normal_z = 0;
    float min_value = 0, curvature = 0, Distance = 0;
    std::vector<int> PointID;
};

// point_cloud.h:295-312 without Eigen: the acquisition header and sensor pose travel with the cloud as plain members
// (Header: sensor_point_cloud2.h's seq / stamp / frame_id; sensor_origin_ x y z w; sensor_orientation_ w x y z)
struct Header {
    uint32_t seq = 0;
    uint64_t stamp = 0;
    std::string frame_id;
};
template <typename PointT> class PointCloud {
  public:
    typedef std::shared_ptr<PointCloud<PointT>> Ptr;
    typedef std::shared_ptr<const PointCloud<PointT>> ConstPtr;
    Header header;
    std::vector<PointT> points;
    uint32_t width = 0, height = 1;
    bool is_dense = true;
    double sensor_origin_[4] = {0, 0, 0, 0};      // Eigen::Vector4d::Zero (point_cloud.h:79)
    double sensor_orientation_[4] = {1, 0, 0, 0}; // Eigen::Quaterniond::Identity, w first
    size_t size() const { return points.size(); }
    void push_back(const PointT& p)
    {
        points.push_back(p);
        width = (uint32_t)points.size();
        height = 1;
    }
    void clear()
    {
        points.clear();
        width = 0;
        height = 1;
    }
    PointT& operator[](size_t i) { return points[i]; }
    const PointT& operator[](size_t i) const { return points[i]; }
};
// boost::shared_array<LAS_POINT_PROPERTY> of the reference (calculate_feature.h:13-16): indexed as PointProperty[j]
// (calculate_feature.cpp:229,341, extraction_tree.cpp:194); .get() gives the raw array
typedef std::shared_ptr<LAS_POINT_PROPERTY[]> PointPropertyArray;
typedef PointXYZRGBA CloudItem;              // cmm_types.h:11-14
typedef PointCloud<CloudItem> Cloud;
typedef Cloud::Ptr CloudPtr;

namespace detail {
inline void check(pcp_ctx* c, int st, const char* what)
{
    if (st != PCP_OK) throw std::runtime_error(std::string(what) + ": " + pcp_last_error(c));
}
// one context per process and device 0, created on first use
inline pcp_ctx* context()
{
    static pcp_ctx* ctx = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        int st = pcp_create(&ctx, 0);
        if (st != PCP_OK) throw std::runtime_error(std::string("pcp_create: ") + pcp_last_error(nullptr));
    });
    return ctx;
}
} // namespace detail

// kd_tree.h:636-1031.  The FLANN kd-tree is replaced by the library's Morton grid; results (indices ascending
// by squared distance, squared distances in double) follow kd_tree.h:814-845 / 863-903.
template <typename PointT> class KdTreeFLANN {
  public:
    typedef std::shared_ptr<const PointCloud<PointT>> PointCloudConstPtr;
    KdTreeFLANN(bool sorted = true) : sorted_(sorted) {}
    ~KdTreeFLANN() { pcp_index_free(index_); }
    KdTreeFLANN(const KdTreeFLANN&) = delete;
    KdTreeFLANN& operator=(const KdTreeFLANN&) = delete;
    void setEpsilon(double) {}                    // kd_tree.h:694-702: the search here is always exact
    void setSortedResults(bool s) { sorted_ = s; } // results are always delivered sorted
    void setInputCloud(const PointCloudConstPtr& cloud)
    {
        pcp_index_free(index_);
        index_ = nullptr;
        input_ = cloud;
        detail::check(detail::context(),
                      pcp_index_build(detail::context(), cloud->points.data(), cloud->points.size(), sizeof(PointT), 0.0,
                                      &index_),
                      "KdTreeFLANN::setInputCloud");
    }
    // const and safe to call from several host threads on one tree (calculate_feature.cpp:222-233)
    int nearestKSearch(const PointT& point, int k, std::vector<int>& k_indices, std::vector<double>& k_sqr_distances) const
    {
        const int64_t nvalid = pcp_index_size(index_);
        if (k > nvalid) k = (int)nvalid; // kd_tree.h:820-821
        k_indices.assign(k, -1);
        k_sqr_distances.assign(k, 0.0);
        if (k <= 0) return 0;
        int32_t found = 0;
        detail::check(detail::context(),
                      pcp_knn(index_, &point, 1, sizeof(PointT), k, k_indices.data(), k_sqr_distances.data(), &found),
                      "KdTreeFLANN::nearestKSearch");
        k_indices.resize(found);
        k_sqr_distances.resize(found);
        return found;
    }
    int radiusSearch(const PointT& point, double radius, std::vector<int>& k_indices, std::vector<double>& k_sqr_distances,
                     unsigned int max_nn = 0) const
    {
        int64_t off[2] = {0, 0};
        detail::check(detail::context(), pcp_radius_count(index_, &point, 1, sizeof(PointT), radius, max_nn, off),
                      "KdTreeFLANN::radiusSearch");
        k_indices.assign((size_t)off[1], -1);
        k_sqr_distances.assign((size_t)off[1], 0.0);
        if (off[1])
            detail::check(detail::context(),
                          pcp_radius_fill(index_, &point, 1, sizeof(PointT), radius, max_nn, off, k_indices.data(),
                                          k_sqr_distances.data()),
                          "KdTreeFLANN::radiusSearch");
        return (int)off[1];
    }
    // the (cloud, index) and (index) overloads of kd_tree.h:446-505 / 538-601
    int nearestKSearch(const PointCloud<PointT>& cloud, int index, int k, std::vector<int>& k_indices,
                       std::vector<double>& k_sqr_distances) const
    {
        return nearestKSearch(cloud.points.at((size_t)index), k, k_indices, k_sqr_distances);
    }
    int nearestKSearch(int index, int k, std::vector<int>& k_indices, std::vector<double>& k_sqr_distances) const
    {
        return nearestKSearch(input_->points.at((size_t)index), k, k_indices, k_sqr_distances);
    }
    int radiusSearch(const PointCloud<PointT>& cloud, int index, double radius, std::vector<int>& k_indices,
                     std::vector<double>& k_sqr_distances, unsigned int max_nn = 0) const
    {
        return radiusSearch(cloud.points.at((size_t)index), radius, k_indices, k_sqr_distances, max_nn);
    }
    int radiusSearch(int index, double radius, std::vector<int>& k_indices, std::vector<double>& k_sqr_distances,
                     unsigned int max_nn = 0) const
    {
        return radiusSearch(input_->points.at((size_t)index), radius, k_indices, k_sqr_distances, max_nn);
    }
    // One launch for a whole cloud of queries -- what the reference's per-point loops (main_blend.cpp:306-326,
    // extraction_tree.cpp:413-436: k = 1 closest points) become: row q of k_indices / k_sqr_distances holds the
    // found[q] <= k neighbours of queries[q] (-1 / +inf padded).
    void nearestKSearchBatch(const PointCloud<PointT>& queries, int k, std::vector<int>& k_indices,
                             std::vector<double>& k_sqr_distances, std::vector<int>& found) const
    {
        const size_t nq = queries.points.size();
        k_indices.assign(nq * (size_t)k, -1);
        k_sqr_distances.assign(nq * (size_t)k, 0.0);
        found.assign(nq, 0);
        if (nq == 0 || k <= 0) return;
        detail::check(detail::context(),
                      pcp_knn(index_, queries.points.data(), nq, sizeof(PointT), k, k_indices.data(), k_sqr_distances.data(),
                              found.data()),
                      "KdTreeFLANN::nearestKSearchBatch");
    }
    PointCloudConstPtr getInputCloud() const { return input_; }
    const pcp_index* handle() const { return index_; }

  private:
    bool sorted_;
    PointCloudConstPtr input_;
    pcp_index* index_ = nullptr;
};

// kd_tree_lod/kd_tree.h:8-22 (cloud_blend_double::KdTree, the float wrapper over trimesh2's KD-tree; dead code in the
// reference: not in the project file, every use commented out).  Same two calls; the answers here are the exact
// double-precision ones of KdTreeFLANN, not the float search + index recovery of kd_tree_lod/kd_tree.cpp:29-117.
class KdTree {
  public:
    void setInputCloud(PointCloud<PointXYZRGBA>::ConstPtr cloud) { tree_.setInputCloud(cloud); }
    int nearestKSearch(const PointXYZRGBA& point, int k, std::vector<int>& k_index, std::vector<double>& k_dis)
    {
        return tree_.nearestKSearch(point, k, k_index, k_dis);
    }

  private:
    KdTreeFLANN<PointXYZRGBA> tree_;
};

// voxel_grid.h:454-1056 (no field filter).
template <typename PointT> class VoxelGrid {
  public:
    typedef std::shared_ptr<const PointCloud<PointT>> PointCloudConstPtr;
    void setInputCloud(const PointCloudConstPtr& cloud) { input_ = cloud; }
    void setLeafSize(double lx, double ly, double lz) { leaf_[0] = lx, leaf_[1] = ly, leaf_[2] = lz; }
    void setDownsampleAllData(bool f) { downsample_all_data_ = f; }
    bool getDownsampleAllData() const { return downsample_all_data_; }
    void setSaveLeafLayout(bool f) { save_leaf_layout_ = f; }
    bool getSaveLeafLayout() const { return save_leaf_layout_; }
    // voxel_grid.h:592-615 (plain arrays where the reference returns Eigen::Vector3i)
    const int* getMinBoxCoordinates() const { return min_b_; }
    const int* getMaxBoxCoordinates() const { return max_b_; }
    const int* getNrDivisions() const { return div_b_; }
    const long long* getDivisionMultiplier() const { return divb_mul_; }
    // voxel_grid.h:672-678: truncating
    void getGridCoordinates(double x, double y, double z, int ijk[3]) const
    {
        ijk[0] = static_cast<int>(double(x * (1.0 / leaf_[0])));
        ijk[1] = static_cast<int>(double(y * (1.0 / leaf_[1])));
        ijk[2] = static_cast<int>(double(z * (1.0 / leaf_[2])));
    }
    // voxel_grid.h:684-696: -1 outside the grid, for an empty cell, or when the layout was not saved
    int getCentroidIndexAt(const int ijk[3]) const
    {
        if (layout_.empty()) return -1;
        long long idx = 0;
        for (int a = 0; a < 3; a++) idx += ((long long)ijk[a] - min_b_[a]) * divb_mul_[a];
        if (idx < 0 || idx >= (long long)div_b_[0] * div_b_[1] * div_b_[2]) return -1;
        auto it = layout_.find(idx);
        return it == layout_.end() ? -1 : it->second;
    }
    int getCentroidIndex(const PointT& p) const
    {
        int ijk[3];
        getGridCoordinates(p.x, p.y, p.z, ijk);
        return getCentroidIndexAt(ijk);
    }
    // voxel_grid.h:648-652: one entry per cell of the grid, -1 = empty (can be large: div_x * div_y * div_z)
    std::vector<int> getLeafLayout() const
    {
        std::vector<int> full;
        if (layout_.empty()) return full;
        full.assign((size_t)div_b_[0] * div_b_[1] * div_b_[2], -1);
        for (const auto& kv : layout_) full[(size_t)kv.first] = kv.second;
        return full;
    }
    void filter(PointCloud<PointT>& output)
    {
        output.clear();
        layout_.clear();
        if (!input_) { // voxel_grid.h:815-821
            std::cout << "[applyFilter] No input dataset given!" << std::endl;
            return;
        }
        const size_t n = input_->points.size();
        output.points.resize(n);
        std::vector<int32_t> cell(save_leaf_layout_ ? 3 * n : 0);
        size_t m = 0;
        detail::check(detail::context(),
                      pcp_voxel_grid(detail::context(), input_->points.data(), n, sizeof(PointT), leaf_,
                                     downsample_all_data_ ? 1 : 0, output.points.data(), &m, min_b_, div_b_,
                                     save_leaf_layout_ ? cell.data() : nullptr, nullptr),
                      "VoxelGrid::filter");
        output.points.resize(m);
        output.width = (uint32_t)m, output.height = 1, output.is_dense = true; // voxel_grid.h:824-825,1055
        output.header = input_->header;                                        // Filter::filter, voxel_grid.h:440-442
        for (int a = 0; a < 4; a++)
            output.sensor_origin_[a] = input_->sensor_origin_[a], output.sensor_orientation_[a] = input_->sensor_orientation_[a];
        for (int a = 0; a < 3; a++) max_b_[a] = min_b_[a] + div_b_[a] - 1;
        divb_mul_[0] = 1, divb_mul_[1] = div_b_[0], divb_mul_[2] = (long long)div_b_[0] * div_b_[1];
        if (save_leaf_layout_) { // outputs are in ascending cell index (voxel_grid.h:948, 1031)
            std::vector<long long> idx;
            idx.reserve(n);
            for (size_t i = 0; i < n; i++)
                if (cell[3 * i] >= 0) idx.push_back(cell[3 * i] + cell[3 * i + 1] * divb_mul_[1] + cell[3 * i + 2] * divb_mul_[2]);
            std::sort(idx.begin(), idx.end());
            idx.erase(std::unique(idx.begin(), idx.end()), idx.end());
            for (size_t i = 0; i < idx.size(); i++) layout_[idx[i]] = (int)i;
        }
    }

  private:
    PointCloudConstPtr input_;
    double leaf_[3] = {0, 0, 0};
    bool downsample_all_data_ = true, save_leaf_layout_ = false;
    int32_t min_b_[3] = {0, 0, 0}, div_b_[3] = {0, 0, 0}, max_b_[3] = {0, 0, 0};
    long long divb_mul_[3] = {0, 0, 0};
    std::unordered_map<long long, int> layout_; // cell index -> output index (sparse leaf_layout_)
};

class PointCloudHelper {
  public:
    // point_cloud_helper.cpp:42-63: in place, float leaf
    static void remove_duplicate(CloudPtr cloud, const float voxel_grid_size)
    {
        size_t n = cloud->points.size();
        detail::check(detail::context(),
                      pcp_remove_duplicate(detail::context(), cloud->points.data(), &n, sizeof(CloudItem), voxel_grid_size),
                      "PointCloudHelper::remove_duplicate");
        cloud->points.resize(n);
        cloud->width = (uint32_t)n, cloud->height = 1, cloud->is_dense = true;
    }
    // point_cloud_helper.cpp:65-73: keep the original if fewer than min_num points would remain
    static void remove_duplicate(CloudPtr cloud, const float voxel_grid_size, int min_num)
    {
        Cloud copy = *cloud;
        remove_duplicate(cloud, voxel_grid_size);
        if ((int)cloud->points.size() < min_num) *cloud = copy;
    }
    static void getMinMax3D(const Cloud& cloud, double mn[3], double mx[3])
    {
        detail::check(detail::context(),
                      pcp_minmax(detail::context(), cloud.points.data(), cloud.points.size(), sizeof(CloudItem), mn, mx),
                      "getMinMax3D");
    }
    static unsigned compute3DCentroid(const Cloud& cloud, double c[3])
    {
        uint64_t n = 0;
        detail::check(detail::context(),
                      pcp_centroid(detail::context(), cloud.points.data(), cloud.points.size(), sizeof(CloudItem), c, &n),
                      "compute3DCentroid");
        return (unsigned)n;
    }
};

class CalculateFeature {
  public:
    // calculate_feature.cpp:119-206
    static PlanSegment calculate_plan_parameter_h_points(CloudPtr cloud)
    {
        const int64_t off[2] = {0, (int64_t)cloud->points.size()};
        pcp_plan_segment s;
        detail::check(detail::context(),
                      pcp_plane_fit(detail::context(), cloud->points.data(), sizeof(CloudItem), off, 1, &s),
                      "calculate_plan_parameter_h_points");
        PlanSegment r;
        r.normal_x = s.normal_x, r.normal_y = s.normal_y, r.normal_z = s.normal_z;
        r.min_value = s.min_value, r.curvature = s.curvature, r.Distance = s.Distance;
        return r;
    }
    // calculate_feature.h:15 (declared, never defined in the reference): per-point plane fit on the k nearest
    // neighbours (radius <= 0) or on the neighbours within `radius`
    static PointPropertyArray calculate_plan_parameter(CloudPtr cloud, double radius, int k = 20)
    {
        PointPropertyArray out(new LAS_POINT_PROPERTY[cloud->points.size() ? cloud->points.size() : 1]);
        pcp_index* ix = nullptr;
        pcp_ctx* c = detail::context();
        detail::check(c, pcp_index_build(c, cloud->points.data(), cloud->points.size(), sizeof(CloudItem), 0.0, &ix),
                      "calculate_plan_parameter");
        int st = radius > 0 ? pcp_features_radius(ix, radius, 0, out.get(), nullptr, nullptr, nullptr)
                            : pcp_features_knn(ix, k, out.get(), nullptr, nullptr, nullptr);
        pcp_index_free(ix);
        detail::check(c, st, "calculate_plan_parameter");
        return out;
    }
    // calculate_feature.cpp:208-368: the robust fit on the 20 nearest neighbours (Radius is unused there too, :233).
    // The reference's rand() hypotheses are replaced by a counter-based generator keyed by `seed`
    // (pcp_features_rpca), so the result is reproducible.
    static PointPropertyArray calculate_plan_parameter_rpca(CloudPtr cloud, double /*Radius*/,
                                                                             float Pr = 0.99f, float epi = 0.5f,
                                                                             uint64_t seed = 1)
    {
        PointPropertyArray out(new LAS_POINT_PROPERTY[cloud->points.size() ? cloud->points.size() : 1]);
        pcp_index* ix = nullptr;
        pcp_ctx* c = detail::context();
        detail::check(c, pcp_index_build(c, cloud->points.data(), cloud->points.size(), sizeof(CloudItem), 0.0, &ix),
                      "calculate_plan_parameter_rpca");
        int st = pcp_features_rpca(ix, 20, Pr, epi, seed, out.get());
        pcp_index_free(ix);
        detail::check(c, st, "calculate_plan_parameter_rpca");
        return out;
    }
    // calculate_feature.cpp:28-33: h_free is not read there either (the exponent is the literal 3)
    static int compute_iteration_number(float Pr, float epi, int /*h_free*/ = 3) { return pcp_rpca_iterations(Pr, epi); }
};

// extraction_tree.h (the reference's spelling).  region_growning (extraction_tree.cpp:170-272): plane segments grown
// over the 50-neighbour graph from seeds in ascending point index; PointProperty[i].SegmentID is rewritten
// (-1 = UNSEGMENTATION) and the kept segments come back in the order found, members in the reference's FIFO order.
// `radius` is not read by the reference either.
class TreeExtration {
  public:
    std::vector<PlanSegment> region_growning(PointPropertyArray PointProperty, CloudPtr& cloud,
                                             double distanceT, double /*radius*/, double cosfaT)
    {
        const size_t n = cloud->points.size();
        std::vector<PlanSegment> result;
        if (n == 0) return result;
        const size_t cap = n / 6 + 1;
        std::vector<int64_t> off(cap + 1);
        std::vector<int32_t> members(n);
        std::vector<float> plane(4 * cap);
        size_t nseg = 0;
        pcp_index* ix = nullptr;
        pcp_ctx* c = detail::context();
        detail::check(c, pcp_index_build(c, cloud->points.data(), n, sizeof(CloudItem), 0.0, &ix), "region_growning");
        const int st = pcp_region_growing(ix, PointProperty.get(), 50, distanceT, cosfaT, 0.005, 5, off.data(), cap,
                                          members.data(), plane.data(), &nseg, nullptr);
        pcp_index_free(ix);
        detail::check(c, st, "region_growning");
        result.resize(nseg);
        for (size_t s = 0; s < nseg; s++) {
            PlanSegment& g = result[s];
            g.SegmentID = (int)s;
            g.normal_x = plane[4 * s], g.normal_y = plane[4 * s + 1], g.normal_z = plane[4 * s + 2], g.Distance = plane[4 * s + 3];
            g.PointID.assign(members.begin() + off[s], members.begin() + off[s + 1]);
        }
        return result;
    }
};

// static.cpp:8-21: robust per-point plane fit, then region growing, with the reference's constants
inline void point_segment(CloudPtr src_cloud, std::vector<PlanSegment>& Segment, uint64_t seed = 1)
{
    PointPropertyArray PointProperty = CalculateFeature::calculate_plan_parameter_rpca(src_cloud, 0.15, 0.99f, 0.5f, seed);
    TreeExtration Tree;
    Segment = Tree.region_growning(PointProperty, src_cloud, 0.5, 0.2, 0.940);
}

// static.cpp:24-36: the points of the segments with more than 50 members
inline void tree_filter(CloudPtr src_cloud, CloudPtr dst_cloud, uint64_t seed = 1)
{
    std::vector<PlanSegment> Segment;
    point_segment(src_cloud, Segment, seed);
    for (size_t i = 0; i < Segment.size(); i++)
        if (Segment[i].PointID.size() > 50)
            for (size_t j = 0; j < Segment[i].PointID.size(); j++) dst_cloud->push_back(src_cloud->points[Segment[i].PointID[j]]);
}

// io::loadPCDFile / savePCDFile* (pcd_helper.h:1343-1497): 0 = ok, -1 = error with the message on stdout, as
// PCDReader / PCDWriter report.  Host-side code of the library (csrc/pcd_io.cu); no device is needed.
namespace io {
namespace detail {
// PointXYZRGBA stores x y z rgba stamp_id; any other record type (PointC ...) stores its leading x y z doubles
template <typename PointT> struct point_kind {
    static const int value = PCP_PCD_POINT_XYZ;
};
template <> struct point_kind<PointXYZRGBA> {
    static const int value = PCP_PCD_POINT_XYZRGBA;
};
inline int report(int st)
{
    if (st == PCP_OK) return 0;
    std::cout << pcp_last_error(nullptr) << std::endl;
    return -1;
}
template <typename PointT> int save(const std::string& file_name, const PointCloud<PointT>& cloud, int mode, int precision)
{
    return report(pcp_pcd_save(file_name.c_str(), point_kind<PointT>::value, cloud.points.data(), cloud.points.size(),
                               sizeof(PointT), cloud.width, cloud.height, mode, precision));
}
} // namespace detail

template <typename PointT> int loadPCDFile(const std::string& file_name, PointCloud<PointT>& cloud)
{
    pcp_pcd_header h;
    if (pcp_pcd_read_header(file_name.c_str(), &h) != PCP_OK) return detail::report(-1);
    cloud.points.assign((size_t)h.points, PointT());
    size_t n = 0;
    int dense = 1;
    const int st = pcp_pcd_load(file_name.c_str(), detail::point_kind<PointT>::value, cloud.points.data(), cloud.points.size(),
                                sizeof(PointT), &n, &dense);
    if (st != PCP_OK) return detail::report(st);
    cloud.width = h.width, cloud.height = h.height, cloud.is_dense = dense != 0;
    return 0;
}
// several files into one cloud through the pipelined loader (pcp_pcd_load_many: host threads decode, the device
// assembles the records); offsets (optional) receives files.size() + 1 record offsets
template <typename PointT>
int loadPCDFiles(const std::vector<std::string>& files, PointCloud<PointT>& cloud, std::vector<size_t>* offsets = nullptr, int threads = 0)
{
    size_t total = 0;
    pcp_pcd_header h;
    for (size_t i = 0; i < files.size(); i++) {
        if (pcp_pcd_read_header(files[i].c_str(), &h) != PCP_OK) return detail::report(-1);
        total += (size_t)h.points;
    }
    std::vector<const char*> names(files.size());
    for (size_t i = 0; i < files.size(); i++) names[i] = files[i].c_str();
    std::vector<size_t> off(files.size() + 1);
    cloud.points.assign(total, PointT());
    int dense = 1;
    pcp_ctx* c = cloud_blend_double::detail::context();
    const int st = pcp_pcd_load_many(c, names.data(), names.size(), detail::point_kind<PointT>::value, cloud.points.data(), total,
                                     sizeof(PointT), threads, off.data(), &dense);
    if (st != PCP_OK) {
        std::cout << pcp_last_error(c) << std::endl;
        return -1;
    }
    cloud.width = (uint32_t)total, cloud.height = 1, cloud.is_dense = dense != 0;
    if (offsets) *offsets = off;
    return 0;
}
template <typename PointT> int savePCDFile(const std::string& file_name, const PointCloud<PointT>& cloud, bool binary_mode = true)
{
    return detail::save(file_name, cloud, binary_mode ? PCP_PCD_BINARY : PCP_PCD_ASCII, 8);
}
template <typename PointT> int savePCDFileASCII(const std::string& file_name, const PointCloud<PointT>& cloud)
{
    return detail::save(file_name, cloud, PCP_PCD_ASCII, 8);
}
template <typename PointT> int savePCDFileBinary(const std::string& file_name, const PointCloud<PointT>& cloud)
{
    return detail::save(file_name, cloud, PCP_PCD_BINARY, 8);
}
template <typename PointT> int savePCDFileBinaryCompressed(const std::string& file_name, const PointCloud<PointT>& cloud)
{
    return detail::save(file_name, cloud, PCP_PCD_BINARY_COMPRESSED, 8);
}
} // namespace io

} // namespace cloud_blend_double
