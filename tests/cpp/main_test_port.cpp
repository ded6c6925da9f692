// The reference's own smoke functions test_kd_tree and test_voxel_grid (main_test.cpp:126-188), written against
// include/pcp/pcp_facade.hpp exactly as they are written against the reference headers.  They print what the
// reference prints; the pytest wrapper compares the output with the derivable known answers.
#include <cmath>
#include <cstdio>
#include "../../include/pcp/pcp_facade.hpp"
using namespace cloud_blend_double;

static void test_voxel_grid()
{
    CloudPtr cloud(new Cloud);
    for (int i = 0; i < 10; i++) {
        CloudItem p;
        p.x = i, p.y = i, p.z = i;
        cloud->push_back(p);
    }
    for (int i = 0; i < 100; i++) {
        CloudItem p;
        p.x = 0.1 * i, p.y = 0.1 * i, p.z = 0.1 * i;
        cloud->push_back(p);
    }
    VoxelGrid<CloudItem> vg;
    vg.setInputCloud(cloud);
    vg.setLeafSize(1, 1, 1);
    vg.setSaveLeafLayout(true);
    Cloud out;
    vg.filter(out);
    printf("voxel %zu\n", out.size());
    for (size_t i = 0; i < out.size() && i < 10; i++) printf("%.17g %.17g %.17g\n", out[i].x, out[i].y, out[i].z);
    const int probe[3] = {4, 4, 4}, hole[3] = {4, 5, 4};
    printf("layout %d %d %d %zu\n", vg.getCentroidIndexAt(probe), vg.getCentroidIndexAt(hole), vg.getCentroidIndex(out[7]),
           vg.getLeafLayout().size());
}

static void test_kd_tree()
{
    CloudPtr cloud(new Cloud);
    for (int i = 0; i < 100; i++) {
        CloudItem p;
        p.x = 0.1 * i, p.y = 0.1 * i, p.z = 0.1 * i;
        cloud->push_back(p);
    }
    KdTreeFLANN<CloudItem> kdtree;
    kdtree.setInputCloud(cloud);
    CloudItem q;
    std::vector<int> index;
    std::vector<double> dis;
    int n = kdtree.nearestKSearch(q, 10, index, dis);
    printf("knn %d\n", n);
    for (int i = 0; i < n; i++) printf("%d %.17g\n", index[i], dis[i]);
    // the batched form of main_blend.cpp:306-326's loop (closest point of every query, one launch)
    std::vector<int> bi, bf;
    std::vector<double> bd;
    kdtree.nearestKSearchBatch(*cloud, 1, bi, bd, bf);
    int self = 0;
    for (size_t i = 0; i < cloud->size(); i++) self += bf[i] == 1 && bi[i] == (int)i && bd[i] == 0.0;
    int n5 = kdtree.nearestKSearch(5, 3, index, dis); // (index) overload: point 5 and its two neighbours
    printf("batch %d of %zu; by index %d: %d %d %d\n", self, cloud->size(), n5, index[0], index[1], index[2]);
    PointCloudHelper::remove_duplicate(cloud, 1.0f);
    printf("dedup %zu\n", cloud->size());
    PointPropertyArray props = CalculateFeature::calculate_plan_parameter_rpca(cloud, 0.15);
    printf("props %d %.9g\n", props[0].PointID, props[0].curvature); // indexed like boost::shared_array (calculate_feature.cpp:229)
}

// calculate_plan_parameter_rpca (calculate_feature.cpp:208-368) on a jittered plane; the Python side of the test
// builds the same cloud and compares the FNV-1a hash of the records with the oracle's
static void test_rpca()
{
    CloudPtr cloud(new Cloud);
    for (int i = 0; i < 2000; i++) {
        CloudItem p;
        p.x = 0.01 * (i % 50), p.y = 0.01 * (i / 50), p.z = 0.001 * ((i * 7919) % 13);
        cloud->push_back(p);
    }
    PointPropertyArray props = CalculateFeature::calculate_plan_parameter_rpca(cloud, 0.15, 0.99f, 0.5f, 4);
    unsigned long long h = 1469598103934665603ull;
    const unsigned char* b = reinterpret_cast<const unsigned char*>(props.get());
    for (size_t i = 0; i < cloud->size() * sizeof(LAS_POINT_PROPERTY); i++) h = (h ^ b[i]) * 1099511628211ull;
    printf("rpca %d %zu %llu\n", CalculateFeature::compute_iteration_number(0.99f, 0.5f, 3), sizeof(LAS_POINT_PROPERTY), h);
}

// The loop of calculate_feature.cpp:222-237 as the reference writes it: one tree shared by the OpenMP threads, one
// nearestKSearch per point into thread-local vectors, PointProperty[j] indexed like boost::shared_array; plus the
// kd_tree_lod call shape, and the header / sensor pose that VoxelGrid::filter copies (voxel_grid.h:440-442).
static void test_openmp_loop()
{
    CloudPtr cloud(new Cloud);
    for (int i = 0; i < 400; i++) {
        CloudItem p;
        p.x = 0.1 * (i % 20), p.y = 0.1 * (i / 20), p.z = 0.01 * ((i * 7919) % 13);
        cloud->push_back(p);
    }
    KdTreeFLANN<CloudItem> kdtree;
    kdtree.setInputCloud(cloud);
    const int point_num = (int)cloud->size();
    PointPropertyArray PointProperty(new LAS_POINT_PROPERTY[point_num]);
    long long idsum = 0;
#pragma omp parallel for schedule(dynamic) reduction(+ : idsum)
    for (int j = 0; j < point_num; j++) {
        std::vector<int> pointIdxNKNSearch(20);
        std::vector<double> pointNKNSquaredDistance(20);
        PointProperty[j].PointID = j;
        PointProperty[j].SegmentID = -1;
        const int N = kdtree.nearestKSearch(cloud->points[j], 20, pointIdxNKNSearch, pointNKNSquaredDistance);
        if (N > 3 && pointIdxNKNSearch[0] == j && pointNKNSquaredDistance[0] == 0.0) idsum += pointIdxNKNSearch[N - 1];
    }
    KdTree lod;
    lod.setInputCloud(cloud);
    std::vector<int> k_index;
    std::vector<double> k_dis;
    const int n1 = lod.nearestKSearch(cloud->points[7], 1, k_index, k_dis);
    cloud->header.seq = 42, cloud->header.frame_id = "velodyne";
    cloud->sensor_origin_[2] = 1.5, cloud->sensor_orientation_[0] = 0.5;
    VoxelGrid<CloudItem> vg;
    vg.setInputCloud(cloud);
    vg.setLeafSize(0.5, 0.5, 0.5);
    Cloud out;
    vg.filter(out);
    printf("omp %lld %d %d %.3g %u %s %.3g %.3g\n", idsum, n1, k_index[0], k_dis[0], out.header.seq, out.header.frame_id.c_str(),
           out.sensor_origin_[2], out.sensor_orientation_[0]);
}

// The sharded pipeline straight through the C ABI, as a caller shaped like main.cpp would use it: one cloud in host
// memory, pcp_mgpu_create / pcp_mgpu_features_knn over 3 shards (on one device here: the contexts share it), against
// remove_duplicate + calculate_plan_parameter on one context.
static void test_mgpu()
{
    CloudPtr cloud(new Cloud);
    for (int i = 0; i < 60000; i++) {
        CloudItem p;
        p.x = 0.004 * (i % 3000) + 1e-5 * ((i * 7919) % 17), p.y = 0.05 * (i / 3000) + 1e-5 * ((i * 31) % 13), p.z = 0.01 * ((i * 13) % 7);
        cloud->push_back(p);
    }
    const size_t n = cloud->size();
    std::vector<CloudItem> out(n);
    std::vector<LAS_POINT_PROPERTY> props(n);
    const int devices[3] = {0, 0, 0};
    pcp_mgpu* mg = nullptr;
    size_t m = 0;
    int st = pcp_mgpu_create(3, devices, &mg);
    if (st == PCP_OK) st = pcp_mgpu_features_knn(mg, cloud->points.data(), n, sizeof(CloudItem), 0.02f, 20, 0.5, out.data(), props.data(), &m);
    if (st != PCP_OK) throw std::runtime_error(std::string("pcp_mgpu: ") + pcp_last_error(nullptr));
    pcp_mgpu_destroy(mg);
    PointCloudHelper::remove_duplicate(cloud, 0.02f);
    PointPropertyArray ref = CalculateFeature::calculate_plan_parameter(cloud, 0.0, 20);
    // order-independent sums: the shards come slab after slab, the single device in voxel order
    double sx = 0, rx = 0, sc = 0, rc = 0;
    for (size_t i = 0; i < m; i++) sx += out[i].x + out[i].y + out[i].z, sc += props[i].curvature + props[i].normal_z * props[i].normal_z;
    for (size_t i = 0; i < cloud->size(); i++) rx += (*cloud)[i].x + (*cloud)[i].y + (*cloud)[i].z, rc += ref[i].curvature + ref[i].normal_z * ref[i].normal_z;
    printf("mgpu %zu %zu %d %d\n", m, cloud->size(), fabs(sx - rx) <= 1e-9 * fabs(rx), fabs(sc - rc) <= 1e-9 * fabs(rc));
}

// point_segment / tree_filter (static.cpp:8-36) on two sheets: the Python side runs the oracle on the same cloud
static void test_segment()
{
    CloudPtr cloud(new Cloud);
    for (int i = 0; i < 6000; i++) {
        CloudItem p;
        const int a = i % 3000;
        p.x = 0.02 * (a % 60) + 1e-4 * ((a * 7919) % 17), p.y = 0.02 * (a / 60), p.z = i < 3000 ? 1e-4 * ((a * 31) % 11) : 2.0 + 0.5 * p.x;
        cloud->push_back(p);
    }
    std::vector<PlanSegment> seg;
    point_segment(cloud, seg, 4);
    unsigned long long h = 1469598103934665603ull;
    size_t members = 0;
    for (size_t s = 0; s < seg.size(); s++) {
        members += seg[s].PointID.size();
        for (size_t j = 0; j < seg[s].PointID.size(); j++) h = (h ^ (unsigned long long)(unsigned)seg[s].PointID[j]) * 1099511628211ull;
    }
    CloudPtr kept(new Cloud);
    tree_filter(cloud, kept, 4);
    printf("segment %zu %zu %llu %zu\n", seg.size(), members, h, kept->size());
}

// main_test.cpp:188-208 (test_double_io): ten points at pi*i saved as a binary PCD and loaded back
static int test_double_io(const char* path)
{
    CloudPtr cloud(new Cloud);
    for (int i = 0; i < 10; i++) {
        CloudItem p;
        p.x = M_PI * i, p.y = M_PI * i, p.z = M_PI * i;
        cloud->push_back(p);
    }
    if (io::savePCDFileBinary(path, *cloud) != 0) return 1;
    cloud->clear();
    if (io::loadPCDFile(path, *cloud) != 0) return 1;
    printf("io %zu\n", cloud->size());
    for (size_t i = 0; i < cloud->size(); i++) printf("%.17g %.17g %.17g %u\n", (*cloud)[i].x, (*cloud)[i].y, (*cloud)[i].z, (*cloud)[i].rgba);
    Cloud missing;
    printf("missing %d\n", io::loadPCDFile(std::string(path) + ".not-there", missing));
    return 0;
}

int main(int argc, char** argv)
{
    if (argc == 3 && std::string(argv[1]) == "io") return test_double_io(argv[2]); // host only: no device needed
    try {
        test_voxel_grid();
        test_kd_tree();
        test_rpca();
        test_segment();
        test_openmp_loop();
        test_mgpu();
    } catch (const std::exception& e) {
        printf("error: %s\n", e.what());
        return 3;
    }
    return 0;
}
