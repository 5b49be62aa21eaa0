// ndtb_pclomp.hpp -- header-only C++ facade over the C ABI of libndtb.so (include/ndtb.h).
//
// It re-creates the class and method names the reference's callers use, so that
// src/registercallback.cpp and run/pipeline_*.cpp of khalisfadil/slam-sam compile against it with a
// two-line change:
//   pclomp::NormalDistributionsTransform<PointSource, PointTarget>      extern/ndt_omp/include/pclomp/ndt_omp.h:78-553
//   pclomp::NdtResult / NdtParams / NeighborSearchMethod                extern/ndt_omp/include/pclomp/ndt_struct.hpp:65-105
//   svn_ndt::SvnNormalDistributionsTransform<PointSource, PointTarget>  extern/svn_ndt/include/svn_ndt.h:65-316
//   svn_ndt::SvnNdtResult                                               extern/svn_ndt/include/svn_ndt.h:40-51
// Same names, argument meaning and error behaviour (no exceptions on the hot path: an empty or
// overflowing target leaves the grid empty and align() returns without converging, like the
// PCL_WARN / early-return paths of the reference).  There is no CPU path behind this header.
//
// Matrix types: with Eigen available (the reference's build) the facade uses Eigen::Matrix4f /
// Eigen::Matrix<double,6,6>; define NDTB_FACADE_NO_EIGEN (or build where <Eigen/Core> is absent) and it
// falls back to the small row-major structs below, which is what the tests in this repository compile.
// Clouds: any type with a `points` std::vector-like member of PointT whose first three floats are
// x, y, z (pcl::PointXYZ: 16-byte stride, pcl::PointXYZI: 32-byte stride) held by shared_ptr.
#ifndef NDTB_PCLOMP_HPP_
#define NDTB_PCLOMP_HPP_

#include "ndtb.h"

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#if !defined(NDTB_FACADE_NO_EIGEN) && defined(__has_include)
#if __has_include(<Eigen/Core>)
#include <Eigen/Core>
#define NDTB_FACADE_EIGEN 1
#endif
#endif

namespace ndtb_facade {

#ifdef NDTB_FACADE_EIGEN
using Matrix4f = Eigen::Matrix4f;
using Matrix4d = Eigen::Matrix4d;
using Matrix6d = Eigen::Matrix<double, 6, 6>;
template <class M> inline M identity() { return M::Identity(); }
#else
template <class T, int N>
struct SmallMatrix {  // row-major N x N
  T m[N * N] = {};
  T& operator()(int r, int c) { return m[N * r + c]; }
  const T& operator()(int r, int c) const { return m[N * r + c]; }
  static SmallMatrix Identity() {
    SmallMatrix a;
    for (int i = 0; i < N; ++i) a.m[N * i + i] = T(1);
    return a;
  }
};
using Matrix4f = SmallMatrix<float, 4>;
using Matrix4d = SmallMatrix<double, 4>;
using Matrix6d = SmallMatrix<double, 6>;
template <class M> inline M identity() { return M::Identity(); }
#endif

// row-major C arrays <-> matrix types (works for Eigen's column-major storage too)
template <class M, class T> inline void to_rows(const M& a, int n, T* out) {
  for (int r = 0; r < n; ++r) for (int c = 0; c < n; ++c) out[n * r + c] = static_cast<T>(a(r, c));
}
template <class M, class T> inline void from_rows(const T* in, int n, M& a) {
  for (int r = 0; r < n; ++r) for (int c = 0; c < n; ++c) a(r, c) = in[n * r + c];
}

// Minimal stand-in for pcl::PointCloud<PointT> (only what the facade touches) for builds without PCL.
template <class PointT>
struct PointCloud {
  std::vector<PointT> points;
  bool is_dense = true;
  using Ptr = std::shared_ptr<PointCloud<PointT>>;
  using ConstPtr = std::shared_ptr<const PointCloud<PointT>>;
  std::size_t size() const { return points.size(); }
};

struct Handle {  // RAII over ndtb_handle
  ndtb_handle h = nullptr;
  explicit Handle(int device) {
    if (ndtb_create(device, &h) != NDTB_OK) throw std::runtime_error("ndtb_create failed: no CUDA device (there is no CPU fallback)");
  }
  ~Handle() { if (h) ndtb_destroy(h); }
  Handle(const Handle&) = delete;
  Handle& operator=(const Handle&) = delete;
};

template <class Cloud>
inline int upload(ndtb_handle h, const Cloud& cloud, bool target) {
  using PointT = typename std::remove_reference<decltype(cloud.points[0])>::type;
  static_assert(sizeof(PointT) >= 12, "point type must start with float x, y, z");
  const float* xyz = cloud.points.empty() ? nullptr : reinterpret_cast<const float*>(cloud.points.data());
  return target ? ndtb_set_target(h, xyz, cloud.points.size(), sizeof(PointT)) : ndtb_set_source(h, xyz, cloud.points.size(), sizeof(PointT));
}

}  // namespace ndtb_facade

// include/ndtb_pcl_adapter.hpp (the pcl::Registration-derived drop-in, for builds WITH PCL) re-homes the classes below into
// namespace ndtb_core and defines pclomp:: / svn_ndt:: itself; without it they ARE pclomp:: / svn_ndt::.
#ifndef NDTB_PCLOMP_NS
#define NDTB_PCLOMP_NS pclomp
#endif
#ifndef NDTB_SVN_NS
#define NDTB_SVN_NS svn_ndt
#endif

namespace NDTB_PCLOMP_NS {

enum NeighborSearchMethod { KDTREE, DIRECT26, DIRECT7, DIRECT1 };  // ndt_struct.hpp:65

struct NdtResult {  // ndt_struct.hpp:67-88
  ndtb_facade::Matrix4f pose = ndtb_facade::identity<ndtb_facade::Matrix4f>();
  float transform_probability = 0.f;
  float nearest_voxel_transformation_likelihood = 0.f;
  int iteration_num = 0;
  std::vector<ndtb_facade::Matrix4f> transformation_array;
  ndtb_facade::Matrix6d hessian;
  std::vector<float> transform_probability_array;
  std::vector<float> nearest_voxel_transformation_likelihood_array;
};

struct NdtParams {  // ndt_struct.hpp:90-105
  double trans_epsilon{};
  double step_size{};
  double resolution{};
  int max_iterations{};
  NeighborSearchMethod search_method{};
  int num_threads{};
  float regularization_scale_factor{};
  bool use_line_search = false;
};

// One exported leaf of getTargetCells().getLeaves() (voxel_grid_covariance_omp.h:101-179), as the map
// export of include/pipeline.hpp:163-222 consumes it.
struct Leaf {
  int nr_points = 0;
  double mean[3] = {0, 0, 0};
  double cov[9] = {}, icov[9] = {}, evecs[9] = {}, evals[3] = {};
};

template <typename PointSource, typename PointTarget, typename CloudSource = ndtb_facade::PointCloud<PointSource>,
          typename CloudTarget = ndtb_facade::PointCloud<PointTarget>>
class NormalDistributionsTransform {
public:
  using PointCloudSource = CloudSource;
  using PointCloudSourceConstPtr = std::shared_ptr<const CloudSource>;
  using PointCloudTarget = CloudTarget;
  using PointCloudTargetConstPtr = std::shared_ptr<const CloudTarget>;
  using Ptr = std::shared_ptr<NormalDistributionsTransform>;
  using Matrix4f = ndtb_facade::Matrix4f;

  explicit NormalDistributionsTransform(int cuda_device = 0) : h_(cuda_device) {  // ndt_omp_impl.hpp:51-81
    ndtb_default_params(NDTB_FLAVOUR_PCLOMP, &p_);
    final_transformation_ = ndtb_facade::identity<Matrix4f>();
  }

  void setNumThreads(int n) { num_threads_ = n; }  // ndt_omp.h:119 (kept for API compatibility: the GPU needs no thread count)
  int getNumThreads() const { return num_threads_; }

  void setInputTarget(const PointCloudTargetConstPtr& cloud) {  // ndt_omp.h:126-130 -> init() :324-330
    target_ = cloud;
    window_target_ = false;
    push_params();
    target_rc_ = cloud ? ndtb_facade::upload(h_.h, *cloud, true) : NDTB_ERR_EMPTY;
  }
  void setInputSource(const PointCloudSourceConstPtr& cloud) {  // pcl::Registration
    source_ = cloud;
    source_rc_ = cloud ? ndtb_facade::upload(h_.h, *cloud, false) : NDTB_ERR_EMPTY;
  }

  // --- extension: sliding-window target kept on the device (replaces the transformPointCloud + `+=` loop of
  // run/pipeline_ligo_tc.cpp:519-526 and the setInputTarget that follows it) ---------------------------------
  int putKeyframe(std::int64_t id, const CloudTarget& cloud) {  // pointsArchive[id].points, uploaded once
    using PointT = typename std::remove_reference<decltype(cloud.points[0])>::type;
    return ndtb_keyframe_put(h_.h, id, cloud.points.empty() ? nullptr : reinterpret_cast<const float*>(cloud.points.data()), cloud.points.size(), sizeof(PointT));
  }
  int dropKeyframe(std::int64_t id) { return ndtb_keyframe_drop(h_.h, id); }
  template <class Matrix4dLike>
  void setInputTargetWindow(const std::vector<std::int64_t>& ids, const std::vector<Matrix4dLike>& poses_b2m) {
    if (ids.size() != poses_b2m.size()) throw std::invalid_argument("ndtb: one pose per keyframe id");
    std::vector<double> P(16 * ids.size());
    for (std::size_t k = 0; k < ids.size(); ++k) ndtb_facade::to_rows(poses_b2m[k], 4, P.data() + 16 * k);
    push_params();
    target_.reset();
    window_target_ = true;
    target_rc_ = ndtb_set_target_window(h_.h, static_cast<int>(ids.size()), ids.data(), P.data());
  }

  void setResolution(float resolution) {  // ndt_omp.h:135-142: rebuilds only if the value changed and a target is set
    if (p_.resolution != resolution) { p_.resolution = resolution; if (target_ || window_target_) target_rc_ = push_params(); }
  }
  float getResolution() const { return p_.resolution; }
  double getStepSize() const { return p_.step_size; }
  void setStepSize(double s) { p_.step_size = s; }
  double getOutlierRatio() const { return p_.outlier_ratio; }
  void setOutlierRatio(double r) { p_.outlier_ratio = r; }
  void setNeighborhoodSearchMethod(NeighborSearchMethod m) {
    p_.search_method = (m == DIRECT7) ? NDTB_DIRECT7 : (m == DIRECT1 ? NDTB_DIRECT1 : (m == DIRECT26 ? NDTB_DIRECT26 : NDTB_KDTREE));
  }
  NeighborSearchMethod getNeighborhoodSearchMethod() const {
    return p_.search_method == NDTB_DIRECT7 ? DIRECT7 : (p_.search_method == NDTB_DIRECT1 ? DIRECT1 : (p_.search_method == NDTB_DIRECT26 ? DIRECT26 : KDTREE));
  }
  void setTransformationEpsilon(double e) { p_.trans_epsilon = e; }
  double getTransformationEpsilon() { return p_.trans_epsilon; }
  void setMaximumIterations(int n) { p_.max_iterations = n; }
  int getMaxIterations() const { return p_.max_iterations; }
  void setRegularizationScaleFactor(float f) { p_.regularization_scale_factor = f; }
  void setRegularizationPose(const Matrix4f& pose) {  // ndt_omp.h:254
    float T[16];
    ndtb_facade::to_rows(pose, 4, T);
    ndtb_set_regularization_pose(h_.h, T);
  }
  void unsetRegularizationPose() { ndtb_set_regularization_pose(h_.h, nullptr); }
  void setMinPointPerVoxel(int n) { p_.min_points_per_voxel = n; if (target_) target_rc_ = push_params(); }
  void setParams(const NdtParams& q) {
    p_.trans_epsilon = q.trans_epsilon; p_.step_size = q.step_size; p_.max_iterations = q.max_iterations;
    p_.regularization_scale_factor = q.regularization_scale_factor; p_.use_line_search = q.use_line_search ? 1 : 0;
    num_threads_ = q.num_threads;
    setNeighborhoodSearchMethod(q.search_method);
    setResolution(static_cast<float>(q.resolution));
  }
  NdtParams getParams() const {
    NdtParams q;
    q.trans_epsilon = p_.trans_epsilon; q.step_size = p_.step_size; q.resolution = p_.resolution; q.max_iterations = p_.max_iterations;
    q.search_method = getNeighborhoodSearchMethod(); q.num_threads = num_threads_;
    q.regularization_scale_factor = p_.regularization_scale_factor; q.use_line_search = p_.use_line_search != 0;
    return q;
  }

  // pcl::Registration::align(output, guess) -> computeTransformation (ndt_omp_impl.hpp:85-202)
  void align(CloudSource& output, const Matrix4f& guess) {
    converged_ = false;
    nr_iterations_ = 0;
    if (push_params() != NDTB_OK || target_rc_ != NDTB_OK || source_rc_ != NDTB_OK || !source_) return;  // reference: PCL_WARN + early out
    float g[16];
    ndtb_facade::to_rows(guess, 4, g);
    ndtb_result r;
    std::vector<float> arr(16 * kMaxArray);
    const std::size_t n = source_->points.size();
    std::vector<float> out(3 * n);
    if (ndtb_align(h_.h, g, &r, arr.data(), kMaxArray, out.data()) != NDTB_OK) return;
    ndtb_facade::from_rows(r.pose, 4, final_transformation_);
    ndtb_facade::from_rows(r.hessian, 6, hessian_);
    trans_probability_ = r.transform_probability;
    nvtl_ = r.nearest_voxel_transformation_likelihood;
    nr_iterations_ = r.iteration_num;
    converged_ = r.converged != 0;
    transformation_array_.resize(r.transformation_array_len);
    for (int i = 0; i < r.transformation_array_len; ++i) ndtb_facade::from_rows(arr.data() + 16 * i, 4, transformation_array_[i]);
    output = *source_;  // keeps intensity etc.; xyz replaced by the transformed coordinates
    for (std::size_t i = 0; i < n; ++i) std::memcpy(reinterpret_cast<float*>(&output.points[i]), out.data() + 3 * i, 12);
  }
  void align(CloudSource& output) { align(output, ndtb_facade::identity<Matrix4f>()); }

  Matrix4f getFinalTransformation() const { return final_transformation_; }
  bool hasConverged() const { return converged_; }
  double getTransformationProbability() const { return trans_probability_; }
  double getNearestVoxelTransformationLikelihood() const { return nvtl_; }
  int getFinalNumIteration() const { return nr_iterations_; }
  ndtb_facade::Matrix6d getHessian() const { return hessian_; }
  const std::vector<Matrix4f>& getFinalTransformationArray() const { return transformation_array_; }
  NdtResult getResult() {  // ndt_omp.h:529-538
    NdtResult r;
    r.pose = final_transformation_;
    r.transform_probability = static_cast<float>(trans_probability_);
    r.nearest_voxel_transformation_likelihood = static_cast<float>(nvtl_);
    r.iteration_num = nr_iterations_;
    r.transformation_array = transformation_array_;
    r.hessian = hessian_;
    return r;
  }
  // calculateTransformationProbability / calculateNearestVoxelTransformationLikelihood (ndt_omp_impl.hpp:1169-1281)
  double calculateTransformationProbability(const Matrix4f& trans) { double tp = 0, nv = 0; scores(trans, &tp, &nv); return tp; }
  double calculateNearestVoxelTransformationLikelihood(const Matrix4f& trans) { double tp = 0, nv = 0; scores(trans, &tp, &nv); return nv; }

  // getTargetCells().getLeaves() / getMinPointPerVoxel() (include/pipeline.hpp:178-206): linear index -> leaf, ascending
  std::vector<std::pair<std::size_t, Leaf>> getLeaves() {
    std::vector<std::pair<std::size_t, Leaf>> out;
    std::size_t V = 0;
    if (ndtb_get_voxel_count(h_.h, &V, nullptr, nullptr, nullptr) != NDTB_OK || V == 0) return out;
    std::vector<int64_t> idx(V); std::vector<int32_t> n(V);
    std::vector<double> mean(3 * V), cov(9 * V), icov(9 * V), evecs(9 * V), evals(3 * V);
    if (ndtb_get_voxels(h_.h, idx.data(), n.data(), mean.data(), cov.data(), icov.data(), evecs.data(), evals.data()) != NDTB_OK) return out;
    out.resize(V);
    for (std::size_t v = 0; v < V; ++v) {
      out[v].first = static_cast<std::size_t>(idx[v]);
      Leaf& l = out[v].second;
      l.nr_points = n[v];
      std::memcpy(l.mean, &mean[3 * v], 24); std::memcpy(l.cov, &cov[9 * v], 72); std::memcpy(l.icov, &icov[9 * v], 72);
      std::memcpy(l.evecs, &evecs[9 * v], 72); std::memcpy(l.evals, &evals[3 * v], 24);
    }
    return out;
  }
  int getMinPointPerVoxel() const { return p_.min_points_per_voxel; }
  std::string lastError() const { return ndtb_last_error(h_.h); }
  ndtb_handle native() { return h_.h; }

private:
  static constexpr int kMaxArray = 64;
  int push_params() { return ndtb_set_params(h_.h, &p_); }
  void scores(const Matrix4f& trans, double* tp, double* nv) {
    float T[16];
    ndtb_facade::to_rows(trans, 4, T);
    ndtb_calculate_scores(h_.h, T, tp, nv);
  }
  ndtb_facade::Handle h_;
  ndtb_params p_{};
  int num_threads_ = 1;
  PointCloudTargetConstPtr target_;
  bool window_target_ = false;  // the target was assembled on the device from keyframes
  PointCloudSourceConstPtr source_;
  int target_rc_ = NDTB_ERR_NOT_READY, source_rc_ = NDTB_ERR_NOT_READY;
  Matrix4f final_transformation_;
  ndtb_facade::Matrix6d hessian_;
  std::vector<Matrix4f> transformation_array_;
  double trans_probability_ = 0, nvtl_ = 0;
  int nr_iterations_ = 0;
  bool converged_ = false;
};

// pclomp::MultiGridNormalDistributionsTransform (extern/ndt_omp/include/multigrid_pclomp/multigrid_ndt_omp.h:96-567):
// the dynamic-map variant -- targets come and go by id, the neighbour search is the radius search over leaf
// centroids, derivatives are FP64.  Same method names and argument meaning as the reference class.
template <typename PointSource, typename PointTarget, typename CloudSource = ndtb_facade::PointCloud<PointSource>,
          typename CloudTarget = ndtb_facade::PointCloud<PointTarget>>
class MultiGridNormalDistributionsTransform {
public:
  using PointCloudSource = CloudSource;
  using PointCloudSourceConstPtr = std::shared_ptr<const CloudSource>;
  using PointCloudTarget = CloudTarget;
  using PointCloudTargetConstPtr = std::shared_ptr<const CloudTarget>;
  using Ptr = std::shared_ptr<MultiGridNormalDistributionsTransform>;
  using Matrix4f = ndtb_facade::Matrix4f;

  explicit MultiGridNormalDistributionsTransform(int cuda_device = 0) : h_(cuda_device) {  // multigrid_ndt_omp_impl.hpp:180-207
    ndtb_default_params(NDTB_FLAVOUR_PCLOMP, &p_);
    final_transformation_ = ndtb_facade::identity<Matrix4f>();
  }
  void setNumThreads(int n) { num_threads_ = n; }  // multigrid_ndt_omp.h:119-123
  int getNumThreads() const { return num_threads_; }
  void setInputSource(const PointCloudSourceConstPtr& cloud) {  // :127-135 (a null source is fatal in the reference)
    if (!cloud) throw std::invalid_argument("Error: Null input source cloud is not allowed");
    source_ = cloud;
    source_rc_ = ndtb_facade::upload(h_.h, *cloud, false);
  }
  void setInputTarget(const PointCloudTargetConstPtr& cloud) {  // :143-148
    addTarget(cloud, "default");
    createVoxelKdtree();
  }
  void addTarget(const PointCloudTargetConstPtr& cloud, const std::string& target_id) {  // :150-155
    using PointT = typename std::remove_reference<decltype(cloud->points[0])>::type;
    ndtb_set_params(h_.h, &p_);
    target_rc_ = cloud ? ndtb_mg_add_target(h_.h, target_id.c_str(), cloud->points.empty() ? nullptr : reinterpret_cast<const float*>(cloud->points.data()),
                                            cloud->points.size(), sizeof(PointT))
                       : NDTB_ERR_EMPTY;
  }
  void removeTarget(const std::string target_id) { ndtb_mg_remove_target(h_.h, target_id.c_str()); }  // :157
  void createVoxelKdtree() { tree_rc_ = ndtb_mg_create_kdtree(h_.h); }                                 // :159-163
  void setResolution(float resolution) { p_.resolution = resolution; }  // :165-171 (the reference re-inits only the default target)
  float getResolution() const { return p_.resolution; }
  double getStepSize() const { return p_.step_size; }
  void setStepSize(double s) { p_.step_size = s; }
  double getOutlierRatio() const { return p_.outlier_ratio; }
  void setOutlierRatio(double r) { p_.outlier_ratio = r; }
  void setTransformationEpsilon(double e) { p_.trans_epsilon = e; }
  double getTransformationEpsilon() { return p_.trans_epsilon; }
  void setMaximumIterations(int n) { p_.max_iterations = n; }
  int getMaxIterations() const { return p_.max_iterations; }
  void setRegularizationScaleFactor(float f) { p_.regularization_scale_factor = f; }
  void setRegularizationPose(const Matrix4f& pose) {
    float T[16];
    ndtb_facade::to_rows(pose, 4, T);
    ndtb_set_regularization_pose(h_.h, T);
  }
  void unsetRegularizationPose() { ndtb_set_regularization_pose(h_.h, nullptr); }
  void setParams(const NdtParams& q) {  // :302-307
    p_.trans_epsilon = q.trans_epsilon; p_.step_size = q.step_size; p_.max_iterations = q.max_iterations;
    p_.regularization_scale_factor = q.regularization_scale_factor; p_.use_line_search = q.use_line_search ? 1 : 0;
    p_.resolution = static_cast<float>(q.resolution);
    num_threads_ = q.num_threads;
  }
  NdtParams getParams() const {
    NdtParams q;
    q.trans_epsilon = p_.trans_epsilon; q.step_size = p_.step_size; q.resolution = p_.resolution; q.max_iterations = p_.max_iterations;
    q.search_method = KDTREE; q.num_threads = num_threads_;
    q.regularization_scale_factor = p_.regularization_scale_factor; q.use_line_search = p_.use_line_search != 0;
    return q;
  }
  std::vector<std::string> getCurrentMapIDs() const {  // :325
    std::size_t need = 0;
    ndtb_mg_get_grid_ids(h_.h, nullptr, 0, &need);
    std::string buf(need, '\0');
    ndtb_mg_get_grid_ids(h_.h, &buf[0], need, nullptr);
    std::vector<std::string> out;
    std::size_t a = 0;
    const std::string s(buf.c_str());
    while (a < s.size()) { std::size_t b = s.find('\n', a); if (b == std::string::npos) b = s.size(); out.push_back(s.substr(a, b - a)); a = b + 1; }
    return out;
  }
  CloudTarget getVoxelPCD() const {  // :323: the leaf centroids the kd-tree is built over
    CloudTarget out;
    std::size_t k = 0;
    if (ndtb_mg_get_leaves(h_.h, 0, nullptr, nullptr, nullptr, nullptr, &k) != NDTB_OK || k == 0) return out;
    std::vector<float> c(3 * k);
    ndtb_mg_get_leaves(h_.h, k, nullptr, nullptr, nullptr, c.data(), nullptr);
    out.points.resize(k);
    for (std::size_t i = 0; i < k; ++i) std::memcpy(reinterpret_cast<float*>(&out.points[i]), c.data() + 3 * i, 12);
    return out;
  }

  void align(CloudSource& output, const Matrix4f& guess) {  // pcl::Registration::align -> computeTransformation :210-323
    converged_ = false;
    nr_iterations_ = 0;
    if (ndtb_set_params(h_.h, &p_) != NDTB_OK || target_rc_ != NDTB_OK || tree_rc_ != NDTB_OK || source_rc_ != NDTB_OK || !source_) return;
    float g[16];
    ndtb_facade::to_rows(guess, 4, g);
    ndtb_result r;
    std::vector<float> arr(16 * kMaxArray);
    const std::size_t n = source_->points.size();
    std::vector<float> out(3 * n);
    if (ndtb_mg_align(h_.h, g, &r, arr.data(), kMaxArray, out.data()) != NDTB_OK) return;
    ndtb_facade::from_rows(r.pose, 4, final_transformation_);
    ndtb_facade::from_rows(r.hessian, 6, hessian_);
    trans_probability_ = r.transform_probability;
    nvtl_ = r.nearest_voxel_transformation_likelihood;
    nr_iterations_ = r.iteration_num;
    converged_ = r.converged != 0;
    transformation_array_.resize(r.transformation_array_len);
    for (int i = 0; i < r.transformation_array_len; ++i) ndtb_facade::from_rows(arr.data() + 16 * i, 4, transformation_array_[i]);
    output = *source_;
    for (std::size_t i = 0; i < n; ++i) std::memcpy(reinterpret_cast<float*>(&output.points[i]), out.data() + 3 * i, 12);
  }
  void align(CloudSource& output) { align(output, ndtb_facade::identity<Matrix4f>()); }

  Matrix4f getFinalTransformation() const { return final_transformation_; }
  bool hasConverged() const { return converged_; }
  double getTransformationProbability() const { return trans_probability_; }
  double getNearestVoxelTransformationLikelihood() const { return nvtl_; }
  int getFinalNumIteration() const { return nr_iterations_; }
  ndtb_facade::Matrix6d getHessian() const { return hessian_; }
  const std::vector<Matrix4f>& getFinalTransformationArray() const { return transformation_array_; }
  NdtResult getResult() {  // :283-296
    NdtResult r;
    r.pose = final_transformation_;
    r.transform_probability = static_cast<float>(trans_probability_);
    r.nearest_voxel_transformation_likelihood = static_cast<float>(nvtl_);
    r.iteration_num = nr_iterations_;
    r.transformation_array = transformation_array_;
    r.hessian = hessian_;
    return r;
  }
  // the reference takes an already transformed cloud; here the transform is passed and applied on the device
  double calculateTransformationProbability(const Matrix4f& trans) { double tp = 0, nv = 0; scores(trans, &tp, &nv); return tp; }
  double calculateNearestVoxelTransformationLikelihood(const Matrix4f& trans) { double tp = 0, nv = 0; scores(trans, &tp, &nv); return nv; }
  std::string lastError() const { return ndtb_last_error(h_.h); }
  ndtb_handle native() { return h_.h; }

private:
  static constexpr int kMaxArray = 64;
  void scores(const Matrix4f& trans, double* tp, double* nv) {
    float T[16];
    ndtb_facade::to_rows(trans, 4, T);
    ndtb_mg_calculate_scores(h_.h, T, tp, nv);
  }
  ndtb_facade::Handle h_;
  ndtb_params p_{};
  int num_threads_ = 1;
  PointCloudSourceConstPtr source_;
  int target_rc_ = NDTB_ERR_NOT_READY, tree_rc_ = NDTB_ERR_NOT_READY, source_rc_ = NDTB_ERR_NOT_READY;
  Matrix4f final_transformation_;
  ndtb_facade::Matrix6d hessian_;
  std::vector<Matrix4f> transformation_array_;
  double trans_probability_ = 0, nvtl_ = 0;
  int nr_iterations_ = 0;
  bool converged_ = false;
};

}  // namespace NDTB_PCLOMP_NS (pclomp)

namespace NDTB_SVN_NS {

enum class NeighborSearchMethod { KDTREE, DIRECT7, DIRECT1 };  // svn_ndt.h:31-36

struct SvnNdtResult {  // svn_ndt.h:40-51
  ndtb_facade::Matrix4d final_pose = ndtb_facade::identity<ndtb_facade::Matrix4d>();            // gtsam::Pose3 as 4x4
  ndtb_facade::Matrix6d final_covariance = ndtb_facade::identity<ndtb_facade::Matrix6d>();      // GTSAM tangent order (rx,ry,rz,x,y,z)
  bool converged = false;
  int iterations = 0;
};

// svn_ndt::SvnNormalDistributionsTransform (svn_ndt.h:65-316).  The reference takes / returns gtsam::Pose3;
// this facade takes the pose as a 4x4 (Pose3::matrix()) so that it builds without GTSAM -- with GTSAM the
// caller passes `prior.matrix()` and builds `gtsam::Pose3(result.final_pose)`.
template <typename PointSource, typename PointTarget, typename CloudSource = ndtb_facade::PointCloud<PointSource>,
          typename CloudTarget = ndtb_facade::PointCloud<PointTarget>>
class SvnNormalDistributionsTransform {
public:
  using PointCloudSource = CloudSource;
  using PointCloudTargetConstPtr = std::shared_ptr<const CloudTarget>;
  using Matrix4d = ndtb_facade::Matrix4d;

  explicit SvnNormalDistributionsTransform(int cuda_device = 0) : h_(cuda_device) {  // svn_ndt_impl.hpp:61-77
    ndtb_default_params(NDTB_FLAVOUR_SVN, &p_);
    ndtb_svn_default_params(&s_);
    ndtb_set_params(h_.h, &p_);
  }
  void setInputTarget(const PointCloudTargetConstPtr& cloud) {  // svn_ndt_impl.hpp:137-159
    target_ = cloud;
    ndtb_set_params(h_.h, &p_);
    target_rc_ = cloud ? ndtb_facade::upload(h_.h, *cloud, true) : NDTB_ERR_EMPTY;
  }
  void setResolution(float r) {  // svn_ndt_impl.hpp:162-176
    if (p_.resolution != r) { p_.resolution = r; if (target_) target_rc_ = ndtb_set_params(h_.h, &p_); }
  }
  float getResolution() const { return p_.resolution; }
  void setMinPointPerVoxel(int n) { p_.min_points_per_voxel = n; if (target_) target_rc_ = ndtb_set_params(h_.h, &p_); }
  void setOutlierRatio(double r) { p_.outlier_ratio = r; }
  void setNeighborhoodSearchMethod(NeighborSearchMethod m) {
    p_.search_method = (m == NeighborSearchMethod::DIRECT7) ? NDTB_DIRECT7 : (m == NeighborSearchMethod::KDTREE ? NDTB_KDTREE : NDTB_DIRECT1);
  }
  void setNumThreads(int n) { num_threads_ = n; }
  void setParticleCount(int k) { s_.particle_count = k; }
  void setMaxIterations(int n) { s_.max_iterations = n; }
  void setKernelBandwidth(double h) { s_.kernel_bandwidth = h; }
  void setStepSize(double e) { s_.step_size = e; }
  void setEarlyStopThreshold(double t) { s_.stop_threshold = t; }
  void setUseGaussNewtonHessian(bool gn) { s_.use_gauss_newton = gn ? 1 : 0; }

  // align(source, prior) (svn_ndt_impl.hpp:675-964).  The reference seeds its particles from the clock (:712);
  // pass `init_particles` (K x 4x4, row-major doubles) or a seed for a reproducible run.
  SvnNdtResult align(const CloudSource& source, const Matrix4d& prior, const double* init_particles = nullptr, uint64_t seed = 0) {
    SvnNdtResult out;
    out.final_pose = prior;  // invalid inputs: converged = false, pose = prior, covariance = I (:682-702)
    if (ndtb_set_params(h_.h, &p_) != NDTB_OK || ndtb_svn_set_params(h_.h, &s_) != NDTB_OK || target_rc_ != NDTB_OK) return out;
    if (ndtb_facade::upload(h_.h, source, false) != NDTB_OK) return out;
    double P[16];
    ndtb_facade::to_rows(prior, 4, P);
    ndtb_svn_result r;
    if (ndtb_svn_align(h_.h, P, init_particles, seed, &r, nullptr) != NDTB_OK) return out;
    ndtb_facade::from_rows(r.final_pose, 4, out.final_pose);
    ndtb_facade::from_rows(r.final_covariance, 6, out.final_covariance);
    out.converged = r.converged != 0;
    out.iterations = r.iterations;
    return out;
  }
  std::string lastError() const { return ndtb_last_error(h_.h); }
  ndtb_handle native() { return h_.h; }

private:
  ndtb_facade::Handle h_;
  ndtb_params p_{};
  ndtb_svn_params s_{};
  int num_threads_ = 1;
  PointCloudTargetConstPtr target_;
  int target_rc_ = NDTB_ERR_NOT_READY;
};

}  // namespace NDTB_SVN_NS (svn_ndt)

#endif  // NDTB_PCLOMP_HPP_
