// cloudclean_gpu.hpp — C++ host layer over the libccgpu C ABI with the reference's own call surface.
//
// This is the file a CloudClean maintainer adds to src/utilities/ (INTEGRATION.md): every function keeps the
// name, argument meaning, result layout and error behaviour of the reference function it replaces, and forwards
// the per-point loop to the GPU through include/ccgpu.h.  No Qt, no PCL needed to compile it:
//   * with -DCCGPU_WITH_PCL the real pcl::PointXYZI / pcl::Normal / Eigen::Vector3f are used;
//   * otherwise layout-compatible stand-ins (same size, same field offsets — SURVEY.md §8 "Host layouts") are
//     defined in namespace ccgpu so the adapters and their tests build in an image without PCL.
//
// Reference functions mirrored (file:line):
//   getPCA<PointT>(cloud*, radius, max_nn)                    src/utilities/cv.h:76-116
//   getHist(cloud, radius, max_nn)                            src/utilities/cv.cpp:249-308
//   NormalEstimator::estimateNormals(cloud)                   src/plugins/normalestimation/normalestimation.cpp:268-470
//   OutlierFilter::filter()  (index list handed to Select)    src/plugins/outlierfilter/outlierfilter.cpp:120-179
//   octreeDownsample<PointT>(cloud*, resolution, sub_idxs)    src/utilities/utils.h:16-62
//   normal_stdev<PointT,NormalT> / stdev_dist                 src/utilities/cv.h:123-206, cv.cpp:166-230
//   principalCurvatures (pcl::PrincipalCurvaturesEstimation)  markov.cpp:389-402, featureeval.cpp:947-956
//   PointCloud::load_ptx(filename, decimation_factor)         src/model/pointcloud.cpp:127-330
//   fpfh (pcl::FPFHEstimation, radius search)                 src/plugins/featureeval/featureeval.cpp:925-932
//   MinCut::buildGraph neighbour + weight loop               src/plugins/graphcut/mincut.cpp:458-475,574-585
//   Markov::mkFeatureVector eigen block                       src/plugins/markov/markov.cpp:449-475
//   FeatureEval::pca_eigen_value_ratio / plane similarity     src/plugins/featureeval/featureeval.cpp:1104-1137,1201-1220
// plus the two north-star operators that have no reference function (PCL definitions):
//   knnNormals (kNN(k)+PCA normals/curvature), statisticalOutliers (mean + alpha*sigma).
//
// Error behaviour: the reference logs with qDebug and returns early / returns an empty result; nothing throws
// across the plugin boundary (App::notify swallows exceptions, app.cpp:345-355).  Here every adapter returns an
// empty result on failure and leaves the message in Session::lastError().
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../include/ccgpu.h"

#ifdef CCGPU_WITH_PCL
#include <pcl/point_types.h>
#include <Eigen/Core>
#endif

namespace ccgpu {

#ifdef CCGPU_WITH_PCL
using PointXYZI = pcl::PointXYZI;
using PointXYZINormal = pcl::PointXYZINormal;
using Normal = pcl::Normal;
using PrincipalCurvatures = pcl::PrincipalCurvatures;
using Vector3f = Eigen::Vector3f;
#else
// pcl::PointXYZI: 32 bytes, x y z 1.0f | intensity pad pad pad (the reference also keeps cloud/grid indices in
// data_c[2..3], pointcloud.cpp:294-295)
struct alignas(16) PointXYZI {
    float x, y, z, w;
    float intensity, c1, c2, c3;
};
// pcl::PrincipalCurvatures: 20 bytes (featureeval.cpp:963-964 reads it as 5 floats, pc1 at offset 3)
struct PrincipalCurvatures {
    float principal_curvature_x, principal_curvature_y, principal_curvature_z;
    float pc1, pc2;
};
// pcl::PointXYZINormal: 48 bytes
struct alignas(16) PointXYZINormal {
    float x, y, z, w;
    float normal_x, normal_y, normal_z, nw;
    float intensity, curvature, c2, c3;
};
// pcl::Normal: 32 bytes, normal[3] data_n[3] | curvature pad pad pad
struct alignas(16) Normal {
    float normal_x, normal_y, normal_z, data_n3;
    float curvature, c1, c2, c3;
};
// Eigen::Vector3f: 12 bytes, not aligned (featureeval.cpp:1176 relies on the 12-byte stride)
struct Vector3f {
    float v[3];
    float &operator[](int i) { return v[i]; }
    const float &operator[](int i) const { return v[i]; }
};
#endif
static_assert(sizeof(PointXYZI) == 32, "pcl::PointXYZI layout");
static_assert(sizeof(PointXYZINormal) == 48, "pcl::PointXYZINormal layout");
static_assert(sizeof(Normal) == 32, "pcl::Normal layout");
static_assert(sizeof(Vector3f) == 12, "Eigen::Vector3f layout");

// One GPU context per thread of use (the estimateNormals worker thread of the reference owns its own).
class Session {
 public:
    explicit Session(int device = 0) {
        const int rc = ccg_create(&ctx_, device);
        if (rc != CCG_OK) {
            err_ = rc == CCG_ERR_CUDA ? "ccg_create: no usable CUDA device (libccgpu has no CPU path)" : "ccg_create: invalid device id";
            ctx_ = nullptr;
        }
    }
    ~Session() {
        if (ctx_) ccg_destroy(ctx_);
    }
    Session(const Session &) = delete;
    Session &operator=(const Session &) = delete;
    bool ok() const { return ctx_ != nullptr; }
    ccg_ctx *ctx() const { return ctx_; }
    const std::string &lastError() const { return err_; }
    // PointCloud::deleting_ (normalestimation.cpp:305-306): polled between stages
    void setCancelFlag(const volatile int *flag) {
        if (ctx_) ccg_set_cancel_flag(ctx_, flag);
    }
    // KdTreeFLANN::setInputCloud / octree build equivalent; any point type whose first three floats are x, y, z
    template <class PointT>
    bool setInputCloud(const PointT *points, size_t n) {
        return check(ccg_set_cloud(ctx_, points, n, sizeof(PointT)));
    }
    bool check(int rc) {
        if (rc == CCG_OK) return true;
        err_ = ctx_ ? ccg_last_error(ctx_) : "no GPU context";
        return false;
    }

 private:
    ccg_ctx *ctx_ = nullptr;
    std::string err_;
};

// ---------------------------------------------------------------------------------------------------------
// getPCA — src/utilities/cv.h:76-116.  Same template, same arguments, same result: one Vector3f of eigenvalues
// (descending) per point, (0,0,1) for neighbourhoods below 3 points.  `cloud` is anything with data()/size()
// of PointT (pcl::PointCloud<PointT>::points is a std::vector<PointT>), so the reference call sites
// `getPCA(cloud.get(), radius, max_nn)` keep compiling with a one-line overload.
// The reference's `continue` on "too many" (cv.h:100-103) can never trigger with a capped search and is dropped.
// ---------------------------------------------------------------------------------------------------------
template <class PointT>
std::shared_ptr<std::vector<Vector3f>> getPCA(Session &s, const PointT *points, size_t n, double radius, unsigned max_nn = 0,
                                               int cov_norm = CCG_COV_NM1) {
    auto eigen_vals = std::make_shared<std::vector<Vector3f>>(n);
    if (!s.ok() || !s.setInputCloud(points, n)) return std::make_shared<std::vector<Vector3f>>();
    if (n == 0) return eigen_vals;
    if (!s.check(ccg_pca_radius(s.ctx(), radius, (int)max_nn, cov_norm, /*normalise_hist=*/0, nullptr,
                                reinterpret_cast<float *>(eigen_vals->data()), nullptr, nullptr)))
        return std::make_shared<std::vector<Vector3f>>();
    return eigen_vals;
}
template <class Cloud>
auto getPCA(Session &s, const Cloud *cloud, double radius, unsigned max_nn = 0)
    -> decltype(getPCA(s, cloud->data(), cloud->size(), radius, max_nn)) {
    return getPCA(s, cloud->data(), cloud->size(), radius, max_nn);
}

// getHist — src/utilities/cv.cpp:249-308: as getPCA, eigenvalue vector L2-normalised (cv.cpp:293).
template <class PointT>
std::shared_ptr<std::vector<Vector3f>> getHist(Session &s, const PointT *points, size_t n, double radius, unsigned max_nn = 0,
                                                int cov_norm = CCG_COV_NM1) {
    auto eigen_vals = std::make_shared<std::vector<Vector3f>>(n);
    if (!s.ok() || !s.setInputCloud(points, n)) return std::make_shared<std::vector<Vector3f>>();
    if (n == 0) return eigen_vals;
    if (!s.check(ccg_pca_radius(s.ctx(), radius, (int)max_nn, cov_norm, /*normalise_hist=*/1, nullptr,
                                reinterpret_cast<float *>(eigen_vals->data()), nullptr, nullptr)))
        return std::make_shared<std::vector<Vector3f>>();
    return eigen_vals;
}

// ---------------------------------------------------------------------------------------------------------
// NormalEstimator::estimateNormals — normalestimation.cpp:268-470.  Inputs are what the reference reads from
// PointCloud: points, cloudToGridMap() (pointcloud.cpp:389-391), scan_width()/scan_height().  Result: one
// pcl::Normal per point, curvature 0, data_n[3] = 1 for fan normals, NaN where nothing could be estimated.
// Returns an empty vector when cancelled (the reference returns nullptr when cloud->deleting_, :305-306).
// ---------------------------------------------------------------------------------------------------------
template <class PointT>
std::vector<Normal> estimateNormals(Session &s, const PointT *points, size_t n, const std::vector<int> &cloud_to_grid_map,
                                    int scan_width, int scan_height, bool knn_fallback = true) {
    std::vector<Normal> normals;
    if (!s.ok() || cloud_to_grid_map.size() != n || !s.setInputCloud(points, n)) return normals;
    std::vector<float> n4(n * 4);
    int32_t missing = 0;
    static_assert(sizeof(int) == sizeof(int32_t), "int is 32-bit");
    if (!s.check(ccg_grid_normals(s.ctx(), reinterpret_cast<const int32_t *>(cloud_to_grid_map.data()), scan_width, scan_height,
                                  knn_fallback ? 1 : 0, n4.data(), &missing)))
        return normals;
    normals.resize(n);
    for (size_t i = 0; i < n; ++i) {
        Normal &o = normals[i];
        std::memset(&o, 0, sizeof(Normal));
        o.normal_x = n4[4 * i]; o.normal_y = n4[4 * i + 1]; o.normal_z = n4[4 * i + 2];
#ifdef CCGPU_WITH_PCL
        o.data_n[3] = n4[4 * i + 3];
#else
        o.data_n3 = n4[4 * i + 3];
#endif
    }
    return normals;
}

// North-star operator a12 (PCL NormalEstimation definition): kNN(k) + PCA normal and curvature, optionally
// flipped towards the sensor origin (PointCloud::sensor_origin_).
template <class PointT>
std::vector<Normal> knnNormals(Session &s, const PointT *points, size_t n, int k, const float *sensor_origin3 = nullptr) {
    std::vector<Normal> normals;
    if (!s.ok() || !s.setInputCloud(points, n)) return normals;
    std::vector<float> n4(n * 4);
    if (!s.check(ccg_pca_knn(s.ctx(), k, CCG_COV_N, sensor_origin3, nullptr, n4.data()))) return normals;
    normals.resize(n);
    for (size_t i = 0; i < n; ++i) {
        Normal &o = normals[i];
        std::memset(&o, 0, sizeof(Normal));
        o.normal_x = n4[4 * i]; o.normal_y = n4[4 * i + 1]; o.normal_z = n4[4 * i + 2];
        o.curvature = n4[4 * i + 3];
    }
    return normals;
}

// ---------------------------------------------------------------------------------------------------------
// OutlierFilter::filter — outlierfilter.cpp:120-179.  radius_cm and min_neighbours are the plugin's spin-box
// values (defaults 100 and 10, :46-47); the search radius is radius_cm/100 metres (:169).  Returns the index
// list the plugin pushes as `new Select(cloud, indices, ...)` (:177), ascending.
// ---------------------------------------------------------------------------------------------------------
inline std::shared_ptr<std::vector<int>> mask_to_indices(const std::vector<uint8_t> &mask) {
    auto indices = std::make_shared<std::vector<int>>();
    for (size_t i = 0; i < mask.size(); ++i)
        if (mask[i]) indices->push_back((int)i);
    return indices;
}
template <class PointT>
std::shared_ptr<std::vector<int>> outlierFilter(Session &s, const PointT *points, size_t n, double radius_cm = 100.0,
                                                int min_neighbours = 10, bool octree_prefilter = true) {
    if (!s.ok() || !s.setInputCloud(points, n)) return std::make_shared<std::vector<int>>();
    std::vector<uint8_t> mask(n);
    if (!s.check(ccg_radius_count_mask(s.ctx(), radius_cm / 100.0, min_neighbours, octree_prefilter ? 1 : 0, mask.data())))
        return std::make_shared<std::vector<int>>();
    return mask_to_indices(mask);
}
// Statistical outlier filter (north-star row a11, PCL StatisticalOutlierRemoval definition).
template <class PointT>
std::shared_ptr<std::vector<int>> statisticalOutliers(Session &s, const PointT *points, size_t n, int mean_k, double stddev_mul) {
    if (!s.ok() || !s.setInputCloud(points, n)) return std::make_shared<std::vector<int>>();
    std::vector<uint8_t> mask(n);
    if (!s.check(ccg_sor_mask(s.ctx(), mean_k, stddev_mul, mask.data(), nullptr))) return std::make_shared<std::vector<int>>();
    return mask_to_indices(mask);
}

// ---------------------------------------------------------------------------------------------------------
// Eigen features — Markov::mkFeatureVector's block markov.cpp:449-475 (7 values per point, as written there) and
// FeatureEval::pca_eigen_value_ratio (featureeval.cpp:1104-1137) / eigen_plane_consine_similarity (:1201-1220).
// ---------------------------------------------------------------------------------------------------------
inline std::vector<float> eigenFeatures(Session &s, const std::vector<Vector3f> &pca, bool as_written = true) {
    std::vector<float> out(pca.size() * 7);
    if (!s.ok() || !s.check(ccg_features(s.ctx(), reinterpret_cast<const float *>(pca.data()), pca.size(), as_written ? 0 : 1, out.data())))
        out.clear();
    return out;
}
inline std::vector<float> pcaEigenValueRatio(Session &s, const std::vector<Vector3f> &pca) {
    std::vector<float> out(pca.size());
    if (!s.ok() || !s.check(ccg_eigen_scalar(s.ctx(), reinterpret_cast<const float *>(pca.data()), pca.size(), 0, out.data()))) out.clear();
    return out;
}
inline std::vector<float> eigenPlaneCosineSimilarity(Session &s, const std::vector<Vector3f> &pca) {
    std::vector<float> out(pca.size());
    if (!s.ok() || !s.check(ccg_eigen_scalar(s.ctx(), reinterpret_cast<const float *>(pca.data()), pca.size(), 1, out.data()))) out.clear();
    return out;
}

// ---------------------------------------------------------------------------------------------------------
// octreeDownsample — src/utilities/utils.h:16-62: one averaged point per voxel of edge `resolution` (ALL float
// fields of PointT, accumulator = default-constructed PCL point as in the reference) and the big -> small index map
// in sub_idxs.  map_small_to_big (utils.h:74-86) is the reference's own host loop and stays as it is.
// Voxel numbering is ascending packed key, not PCL's octree DFS order (include/ccgpu.h): consumers index the small
// cloud only through sub_idxs, which is consistent with it.
// ---------------------------------------------------------------------------------------------------------
template <class PointT>
std::vector<PointT> octreeDownsample(Session &s, const PointT *input, size_t n, float resolution, std::vector<int> &sub_idxs) {
    static_assert(sizeof(PointT) % sizeof(float) == 0 && sizeof(PointT) / sizeof(float) <= 16, "point type is a short array of floats");
    std::vector<PointT> output;
    sub_idxs.assign(n, 0);
    if (!s.ok() || !s.setInputCloud(input, n) || n == 0) return output;
    int64_t n_out = 0;
    if (!s.check(ccg_voxel_downsample(s.ctx(), input, sizeof(PointT), (int)(sizeof(PointT) / sizeof(float)), resolution, /*as_written=*/1,
                                      reinterpret_cast<int32_t *>(sub_idxs.data()), &n_out)))
        return output;
    output.resize((size_t)n_out);
    if (!s.check(ccg_voxel_downsample_fetch(s.ctx(), reinterpret_cast<float *>(output.data())))) output.clear();
    return output;
}

// ---------------------------------------------------------------------------------------------------------
// kNN graph for MinCut::buildGraph (graphcut/mincut.cpp:458-475) and Flood::flood (flood.cpp:313-356): the neighbour
// table the reference obtains with one nearestKSearch per node, plus calculateBinaryPotential (mincut.cpp:574-585)
// per edge.  neighbours / weights are n*k row-major, entry 0 of a row is the node itself (the reference starts its
// edge loop at 1, :467).  Returns false on failure.
// ---------------------------------------------------------------------------------------------------------
template <class PointT>
bool knnGraph(Session &s, const PointT *points, size_t n, int k, double inverse_sigma, std::vector<int> &neighbours,
              std::vector<double> &weights) {
    neighbours.assign(n * (size_t)k, -1);
    weights.assign(n * (size_t)k, 0.0);
    if (!s.ok() || !s.setInputCloud(points, n)) return false;
    if (n == 0) return true;
    return s.check(ccg_knn_graph(s.ctx(), k, inverse_sigma, reinterpret_cast<int32_t *>(neighbours.data()), weights.data()));
}

// ---------------------------------------------------------------------------------------------------------
// KdTreeFLANN::radiusSearch(i, radius, k_indices, k_sqr_distances, max_nn) for EVERY point of the cloud at once
// (cv.h:98, cv.h:162, cv.cpp:273, featureeval/utils.h:50,145, markov.cpp:360-367): row i of the result is
// indices[offsets[i] .. offsets[i+1]) / sqr_distances[...], ascending distance, the point itself first; max_nn = 0
// means every point inside the radius, any positive cap keeps the max_nn nearest.  Returns false on failure.
// ---------------------------------------------------------------------------------------------------------
template <class PointT>
bool radiusSearchAll(Session &s, const PointT *points, size_t n, double radius, unsigned max_nn, std::vector<int64_t> &offsets,
                     std::vector<int> &indices, std::vector<float> &sqr_distances) {
    offsets.assign(n + 1, 0);
    indices.clear(); sqr_distances.clear();
    if (!s.ok() || !s.setInputCloud(points, n)) return false;
    if (n == 0) return true;
    int64_t total = 0;
    if (!s.check(ccg_radius_csr(s.ctx(), radius, (int)max_nn, 1, offsets.data(), nullptr, nullptr, 0, &total))) return false;
    indices.resize((size_t)total); sqr_distances.resize((size_t)total);
    if (total == 0) return true;
    return s.check(ccg_radius_csr(s.ctx(), radius, (int)max_nn, 1, offsets.data(), reinterpret_cast<int32_t *>(indices.data()),
                                  sqr_distances.data(), total, &total));
}

// ---------------------------------------------------------------------------------------------------------
// normal_stdev<PointT,NormalT> (cv.h:123-206) and stdev_dist (cv.cpp:166-230): one float per point, arithmetic and
// quirks as written in the reference (include/ccgpu.h).  Any max_nn the reference accepts (0 = every point inside r).
// ---------------------------------------------------------------------------------------------------------
template <class PointT, class NormalT>
std::shared_ptr<std::vector<float>> normal_stdev(Session &s, const PointT *cloud, const NormalT *normals, size_t n, double radius, int max_nn) {
    auto std_devs = std::make_shared<std::vector<float>>(n, 0.f);
    if (!s.ok() || !s.setInputCloud(cloud, n) || n == 0) return std_devs;
    if (!s.check(ccg_normal_stdev(s.ctx(), radius, max_nn, normals, sizeof(NormalT), std_devs->data()))) std_devs->assign(n, 0.f);
    return std_devs;
}
template <class PointT>
std::shared_ptr<std::vector<float>> stdev_dist(Session &s, const PointT *cloud, size_t n, double radius, int max_nn, bool use_depth = false,
                                               const float *sensor_origin3 = nullptr) {
    auto stdevs = std::make_shared<std::vector<float>>(n, 0.f);
    if (!s.ok() || !s.setInputCloud(cloud, n) || n == 0) return stdevs;
    if (!s.check(ccg_stdev_dist(s.ctx(), radius, max_nn, use_depth ? 1 : 0, sensor_origin3, stdevs->data()))) stdevs->assign(n, 0.f);
    return stdevs;
}

// calcDistHist / calcIntensityHist (src/plugins/featureeval/utils.h:18-110, 113-195): one `bins`-bin histogram per point.
template <class PointT>
std::shared_ptr<std::vector<std::vector<float>>> calcDistHist(Session &s, const PointT *cloud, size_t n, int bins, double radius, int max_nn) {
    auto histograms = std::make_shared<std::vector<std::vector<float>>>(n, std::vector<float>((size_t)bins, 0.0f));
    if (!s.ok() || !s.setInputCloud(cloud, n) || n == 0) return histograms;
    std::vector<float> flat(n * (size_t)bins);
    if (!s.check(ccg_neighbour_hist(s.ctx(), radius, max_nn, bins, 0, nullptr, 0, flat.data()))) return histograms;
    for (size_t i = 0; i < n; ++i) (*histograms)[i].assign(flat.begin() + i * bins, flat.begin() + (i + 1) * bins);
    return histograms;
}
// PointT must have an `intensity` member (pcl::PointXYZI / PointXYZINormal)
template <class PointT>
std::shared_ptr<std::vector<std::vector<float>>> calcIntensityHist(Session &s, const PointT *cloud, size_t n, int bins, double radius,
                                                                   int max_nn) {
    auto histograms = std::make_shared<std::vector<std::vector<float>>>(n, std::vector<float>((size_t)bins, 0.0f));
    if (!s.ok() || !s.setInputCloud(cloud, n) || n == 0) return histograms;
    std::vector<float> flat(n * (size_t)bins);
    if (!s.check(ccg_neighbour_hist(s.ctx(), radius, max_nn, bins, 1, &cloud[0].intensity, sizeof(PointT), flat.data()))) return histograms;
    for (size_t i = 0; i < n; ++i) (*histograms)[i].assign(flat.begin() + i * bins, flat.begin() + (i + 1) * bins);
    return histograms;
}

// don<PointT,NormalT>(cloud, normals, radius1, radius2) — featureeval.cpp:694-739, the difference of normals.
template <class PointT, class NormalT>
std::vector<Normal> don(Session &s, const PointT *cloud, const NormalT *normals, size_t n, float radius1 = 0.2f, float radius2 = 0.5f) {
    std::vector<Normal> donormals;
    if (!s.ok() || !s.setInputCloud(cloud, n) || n == 0) return donormals;
    std::vector<float> d4(n * 4);
    if (!s.check(ccg_don(s.ctx(), normals, sizeof(NormalT), radius1, radius2, d4.data()))) return donormals;
    donormals.resize(n);
    for (size_t i = 0; i < n; ++i) {
        std::memset(&donormals[i], 0, sizeof(Normal));
        donormals[i].normal_x = d4[4 * i]; donormals[i].normal_y = d4[4 * i + 1]; donormals[i].normal_z = d4[4 * i + 2];
    }
    return donormals;
}

// pcl::PrincipalCurvaturesEstimation<PointT, NormalT, pcl::PrincipalCurvatures> with setRadiusSearch(radius) —
// the block at markov.cpp:389-402 and featureeval.cpp:947-956 (setInputCloud, setInputNormals, compute) as one call.
template <class PointT, class NormalT>
std::vector<PrincipalCurvatures> principalCurvatures(Session &s, const PointT *cloud, const NormalT *normals, size_t n, double radius) {
    std::vector<PrincipalCurvatures> out;
    static_assert(sizeof(PrincipalCurvatures) == 20, "five packed floats");
    if (!s.ok() || !s.setInputCloud(cloud, n) || n == 0) return out;
    out.resize(n);
    const void *nrm = &normals[0].normal_x;          // pcl::Normal (offset 0) or pcl::PointXYZINormal (offset 16), as the reference uses
    if (!s.check(ccg_principal_curvatures(s.ctx(), radius, nrm, sizeof(NormalT), reinterpret_cast<float *>(out.data())))) out.clear();
    return out;
}

// pcl::FPFHEstimation<PointT, NormalT, pcl::FPFHSignature33> with setRadiusSearch(radius) — the block at
// featureeval.cpp:925-932 as one call; rows are pcl::FPFHSignature33 (float histogram[33]).
struct FPFHSignature33 {
    float histogram[33];
};
template <class PointT, class NormalT>
std::vector<FPFHSignature33> fpfh(Session &s, const PointT *cloud, const NormalT *normals, size_t n, double radius) {
    std::vector<FPFHSignature33> out;
    static_assert(sizeof(FPFHSignature33) == 132, "33 packed floats");
    if (!s.ok() || !s.setInputCloud(cloud, n) || n == 0) return out;
    out.resize(n);
    const void *nrm = &normals[0].normal_x;
    if (!s.check(ccg_fpfh(s.ctx(), radius, nrm, sizeof(NormalT), reinterpret_cast<float *>(out.data()), nullptr))) out.clear();
    return out;
}

// PointCloud::load_ptx(filename, decimation_factor) — src/model/pointcloud.cpp:127-330.  Everything the method leaves in
// the PointCloud object: points (with data_c[2..3] = cloud / grid index, :284-285), cloud_to_grid_map_, scan_width_ /
// scan_height_, sensor_origin_ and the 3x3 axes (the caller stores Quaternionf(orientation^T), :167-172).  ok == false
// mirrors `return false` (file cannot be opened, read or parsed; the message is in Session::lastError()).
struct PtxCloud {
    bool ok = false;
    std::vector<PointXYZI> points;
    std::vector<int> cloud_to_grid_map;
    int scan_width = 0, scan_height = 0;
    float sensor_origin[4] = {0, 0, 0, 0};
    float orientation[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
};
inline PtxCloud load_ptx(Session &s, const char *filename, int decimation_factor = 1, bool keep_as_session_cloud = false) {
    PtxCloud c;
    static_assert(sizeof(PointXYZI) == 32, "pcl::PointXYZI is 32 bytes");
    if (!s.ok()) return c;
    ccg_ptx_info info;
    if (!s.check(ccg_ptx_load(s.ctx(), filename, decimation_factor, &info))) return c;
    c.points.resize((size_t)info.points);
    c.cloud_to_grid_map.resize((size_t)info.points);
    static_assert(sizeof(int) == sizeof(int32_t), "cloud_to_grid_map_ is std::vector<int>");
    if (info.points && !s.check(ccg_ptx_fetch(s.ctx(), c.points.data(), reinterpret_cast<int32_t *>(c.cloud_to_grid_map.data())))) return c;
    // the parsed rows are already on the device: raw-ABI feature calls that follow a load can skip their upload
    if (keep_as_session_cloud && info.points && !s.check(ccg_ptx_set_cloud(s.ctx()))) return c;
    c.scan_width = info.scan_width; c.scan_height = info.scan_height;
    for (int i = 0; i < 3; ++i) c.sensor_origin[i] = info.sensor_origin[i];
    for (int i = 0; i < 9; ++i) c.orientation[i] = info.orientation[i];
    c.ok = true;
    return c;
}

}  // namespace ccgpu
