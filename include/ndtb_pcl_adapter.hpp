// ndtb_pcl_adapter.hpp -- the pcl::Registration-derived drop-in, for builds WITH PCL (the reference's own build).
//
// north_star: "registercallback.cpp and the run/pipeline_*.cpp drivers link against it unchanged".  The reference parks its
// registrar in a BASE-CLASS pointer, `pcl::Registration<pcl::PointXYZI, pcl::PointXYZI>::Ptr registration`
// (include/registercallback.hpp:35), assigns a pclomp::NormalDistributionsTransform to it (run/pipeline.cpp:481) and drives it
// through the base (`registration->setInputTarget / setInputSource / align / getFinalTransformation`, run/pipeline.cpp:557-566).
// So the replacement has to BE a pcl::Registration.  This header defines
//     pclomp::NormalDistributionsTransform<PointSource, PointTarget> : public pcl::Registration<PointSource, PointTarget>
// (extern/ndt_omp/include/pclomp/ndt_omp.h:78-553) with the reference's public method set, forwarding to the PCL-free facade of
// include/ndtb_pclomp.hpp (re-homed into namespace ndtb_core) and through it to the C ABI of libndtb.so.  With
//     #include <ndtb_pcl_adapter.hpp>      instead of      #include <pclomp/ndt_omp.h>
// include/registercallback.hpp, src/registercallback.cpp and run/pipeline*.cpp compile UNCHANGED -- the member keeps its type.
//
// Only PCL's PUBLIC base API is used: the virtual setInputSource / setInputTarget, the protected virtual
// computeTransformation(output, guess) that pcl::Registration::align calls, and the protected result members align reads back
// (final_transformation_, transformation_, previous_transformation_, converged_, nr_iterations_, max_iterations_,
// transformation_epsilon_, reg_name_).  Matrices are only touched through operator()(row, col), never by Eigen name.
// This image has no PCL, so the header is compile-checked here against a minimal stand-in of those PCL declarations
// (tests/cpp/mock_pcl/, test infrastructure) and exercised on the GPU through the base pointer (tests/cpp/adapter_example.cpp).
//
// One behaviour of the PCL base is worth knowing: pcl::Registration::align -> initCompute() builds a FLANN kd-tree over the
// TARGET cloud whenever the target changed (it serves getFitnessScore and correspondence estimation; the reference pays for
// it too).  NDT never queries that tree; call setSkipTargetKdTree(true) to hand PCL the existing tree with
// force_no_recompute (pcl::Registration::setSearchMethodTarget) and skip that host-side build.
#ifndef NDTB_PCL_ADAPTER_HPP_
#define NDTB_PCL_ADAPTER_HPP_

#if !defined(__has_include)
#error "ndtb_pcl_adapter.hpp needs __has_include (C++17)"
#endif
#if !__has_include(<pcl/registration/registration.h>)
#error "ndtb_pcl_adapter.hpp needs PCL (<pcl/registration/registration.h>); without PCL use include/ndtb_pclomp.hpp directly"
#endif

#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#include <pcl/registration/registration.h>

#define NDTB_PCLOMP_NS ndtb_core
#define NDTB_SVN_NS ndtb_core_svn
#include "ndtb_pclomp.hpp"

namespace pclomp {

using ndtb_core::NeighborSearchMethod;
using ndtb_core::KDTREE;
using ndtb_core::DIRECT26;
using ndtb_core::DIRECT7;
using ndtb_core::DIRECT1;
using ndtb_core::NdtParams;
using ndtb_core::Leaf;

template <typename PointSource, typename PointTarget>
class NormalDistributionsTransform : public pcl::Registration<PointSource, PointTarget> {
  using Base = pcl::Registration<PointSource, PointTarget>;
  using Core = ndtb_core::NormalDistributionsTransform<PointSource, PointTarget, pcl::PointCloud<PointSource>, pcl::PointCloud<PointTarget>>;

protected:
  using PointCloudSource = typename Base::PointCloudSource;
  using PointCloudSourceConstPtr = typename PointCloudSource::ConstPtr;
  using PointCloudTarget = typename Base::PointCloudTarget;
  using PointCloudTargetConstPtr = typename PointCloudTarget::ConstPtr;
  using Matrix4 = typename Base::Matrix4;   // Eigen::Matrix4f in PCL

  using Base::converged_;
  using Base::final_transformation_;
  using Base::input_;
  using Base::max_iterations_;
  using Base::nr_iterations_;
  using Base::previous_transformation_;
  using Base::reg_name_;
  using Base::target_;
  using Base::transformation_;
  using Base::transformation_epsilon_;

public:
  using Ptr = std::shared_ptr<NormalDistributionsTransform<PointSource, PointTarget>>;
  using ConstPtr = std::shared_ptr<const NormalDistributionsTransform<PointSource, PointTarget>>;

  // pclomp::NdtResult (ndt_struct.hpp:67-88) with PCL's matrix types
  struct NdtResult {
    Matrix4 pose;
    float transform_probability = 0.f;
    float nearest_voxel_transformation_likelihood = 0.f;
    int iteration_num = 0;
    std::vector<Matrix4> transformation_array;
    ndtb_facade::Matrix6d hessian;
    std::vector<float> transform_probability_array;
    std::vector<float> nearest_voxel_transformation_likelihood_array;
  };

  explicit NormalDistributionsTransform(int cuda_device = 0) : core_(cuda_device) {  // ndt_omp_impl.hpp:51-81
    reg_name_ = "NormalDistributionsTransform";
    transformation_epsilon_ = 0.1;
    max_iterations_ = 35;
  }
  ~NormalDistributionsTransform() override {}

  void setNumThreads(int n) { core_.setNumThreads(n); }
  int getNumThreads() const { return core_.getNumThreads(); }

  // ndt_omp.h:126-130: Registration::setInputTarget + init() (the voxel-map build)
  void setInputTarget(const PointCloudTargetConstPtr& cloud) override {
    Base::setInputTarget(cloud);
    core_.setInputTarget(cloud);
  }
  void setInputSource(const PointCloudSourceConstPtr& cloud) override {
    Base::setInputSource(cloud);
    core_.setInputSource(cloud);
  }

  void setResolution(float resolution) { core_.setResolution(resolution); }   // ndt_omp.h:135-142
  float getResolution() const { return core_.getResolution(); }
  double getStepSize() const { return core_.getStepSize(); }
  void setStepSize(double s) { core_.setStepSize(s); }
  double getOutlierRatio() const { return core_.getOutlierRatio(); }
  void setOutlierRatio(double r) { core_.setOutlierRatio(r); }
  void setNeighborhoodSearchMethod(NeighborSearchMethod m) { core_.setNeighborhoodSearchMethod(m); }
  NeighborSearchMethod getNeighborhoodSearchMethod() const { return core_.getNeighborhoodSearchMethod(); }
  void setRegularizationScaleFactor(float f) { core_.setRegularizationScaleFactor(f); }
  void setRegularizationPose(const Matrix4& pose) {   // ndt_omp.h:277-283
    ndtb_facade::Matrix4f T;
    copy4(pose, T);
    core_.setRegularizationPose(T);
  }
  void unsetRegularizationPose() { core_.unsetRegularizationPose(); }
  void setMinPointPerVoxel(int n) { core_.setMinPointPerVoxel(n); }
  void setParams(const NdtParams& q) {
    core_.setParams(q);
    transformation_epsilon_ = q.trans_epsilon;
    max_iterations_ = q.max_iterations;
  }
  NdtParams getParams() const {
    NdtParams q = core_.getParams();
    q.trans_epsilon = transformation_epsilon_;
    q.max_iterations = max_iterations_;
    return q;
  }

  double getTransformationProbability() const { return core_.getTransformationProbability(); }
  double getNearestVoxelTransformationLikelihood() const { return core_.getNearestVoxelTransformationLikelihood(); }
  int getFinalNumIteration() const { return nr_iterations_; }
  ndtb_facade::Matrix6d getHessian() const { return core_.getHessian(); }
  std::vector<Matrix4> getFinalTransformationArray() const {
    std::vector<Matrix4> out(core_.getFinalTransformationArray().size());
    for (std::size_t i = 0; i < out.size(); ++i) copy4(core_.getFinalTransformationArray()[i], out[i]);
    return out;
  }
  NdtResult getResult() {   // ndt_omp.h:529-538
    NdtResult r;
    r.pose = final_transformation_;
    r.transform_probability = static_cast<float>(core_.getTransformationProbability());
    r.nearest_voxel_transformation_likelihood = static_cast<float>(core_.getNearestVoxelTransformationLikelihood());
    r.iteration_num = nr_iterations_;
    r.transformation_array = getFinalTransformationArray();
    r.hessian = core_.getHessian();
    return r;
  }
  double calculateTransformationProbability(const Matrix4& trans) { ndtb_facade::Matrix4f T; copy4(trans, T); return core_.calculateTransformationProbability(T); }
  double calculateNearestVoxelTransformationLikelihood(const Matrix4& trans) {
    ndtb_facade::Matrix4f T; copy4(trans, T);
    return core_.calculateNearestVoxelTransformationLikelihood(T);
  }
  // export path of include/pipeline.hpp:163-222 (getTargetCells().getLeaves(), getMinPointPerVoxel())
  std::vector<std::pair<std::size_t, Leaf>> getLeaves() { return core_.getLeaves(); }
  int getMinPointPerVoxel() const { return core_.getMinPointPerVoxel(); }

  // see the header comment: skip PCL's own kd-tree over the target in initCompute()
  void setSkipTargetKdTree(bool skip) { this->setSearchMethodTarget(this->getSearchMethodTarget(), skip); }

  std::string lastError() const { return core_.lastError(); }
  ndtb_handle native() { return core_.native(); }

  // the reference's gtest calls the protected member directly (extern/svn_ndt/test/test_svn_ndt.cpp:171)
  void computeTransformation(PointCloudSource& output) { computeTransformation(output, identity4()); }

protected:
  // pcl::Registration::align(output, guess) has copied the source into `output`, reset final_/transformation_ to identity and
  // now asks for the transformation (ndt_omp_impl.hpp:85-202).  The Newton loop runs on the GPU; PCL's members get the outcome.
  void computeTransformation(PointCloudSource& output, const Matrix4& guess) override {
    nr_iterations_ = 0;
    converged_ = false;
    core_.setTransformationEpsilon(transformation_epsilon_);   // pcl::Registration::setTransformationEpsilon / setMaximumIterations
    core_.setMaximumIterations(max_iterations_);
    ndtb_facade::Matrix4f g;
    copy4(guess, g);
    core_.align(output, g);
    copy4(core_.getFinalTransformation(), final_transformation_);
    nr_iterations_ = core_.getFinalNumIteration();
    converged_ = core_.hasConverged();
    // transformation_ = the last Newton step, previous_transformation_ = the one before (ndt_omp_impl.hpp:165-178)
    const auto& arr = core_.getFinalTransformationArray();
    if (arr.size() >= 2) {
      copy4(arr[arr.size() - 2], previous_transformation_);
      copy4(arr[arr.size() - 1], transformation_);
    }
  }

private:
  template <class A, class B>
  static void copy4(const A& a, B& b) {
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) b(r, c) = a(r, c);
  }
  static Matrix4 identity4() {
    Matrix4 m;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m(r, c) = (r == c) ? 1.0f : 0.0f;
    return m;
  }
  Core core_;
};

}  // namespace pclomp

// svn_ndt::SvnNormalDistributionsTransform has no base class in the reference (extern/svn_ndt/include/svn_ndt.h:65): the facade
// class is the drop-in; with PCL its clouds are pcl::PointCloud, and with GTSAM present align() takes / returns gtsam::Pose3 like
// the reference (svn_ndt.h:178-182, run/pipeline_lo_svn.cpp:388).
#if __has_include(<gtsam/geometry/Pose3.h>)
#include <gtsam/geometry/Pose3.h>
#define NDTB_HAVE_GTSAM 1
#endif

namespace svn_ndt {

using NeighborSearchMethod = ndtb_core_svn::NeighborSearchMethod;

#ifdef NDTB_HAVE_GTSAM
struct SvnNdtResult {   // svn_ndt.h:40-51
  gtsam::Pose3 final_pose;
  Eigen::Matrix<double, 6, 6> final_covariance = Eigen::Matrix<double, 6, 6>::Identity();
  bool converged = false;
  int iterations = 0;
};
#else
using SvnNdtResult = ndtb_core_svn::SvnNdtResult;
#endif

template <typename PointSource, typename PointTarget>
class SvnNormalDistributionsTransform
    : public ndtb_core_svn::SvnNormalDistributionsTransform<PointSource, PointTarget, pcl::PointCloud<PointSource>, pcl::PointCloud<PointTarget>> {
  using Core = ndtb_core_svn::SvnNormalDistributionsTransform<PointSource, PointTarget, pcl::PointCloud<PointSource>, pcl::PointCloud<PointTarget>>;

public:
  using Core::Core;
  using Core::align;
#ifdef NDTB_HAVE_GTSAM
  SvnNdtResult align(const pcl::PointCloud<PointSource>& source, const gtsam::Pose3& prior) {
    ndtb_facade::Matrix4d P;
    const auto M = prior.matrix();
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) P(r, c) = M(r, c);
    const ndtb_core_svn::SvnNdtResult r = Core::align(source, P);
    SvnNdtResult out;
    Eigen::Matrix4d F;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) F(i, j) = r.final_pose(i, j);
    out.final_pose = gtsam::Pose3(F);
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) out.final_covariance(i, j) = r.final_covariance(i, j);
    out.converged = r.converged;
    out.iterations = r.iterations;
    return out;
  }
#endif
};

}  // namespace svn_ndt

#endif  // NDTB_PCL_ADAPTER_HPP_
