// TEST INFRASTRUCTURE -- stand-in for pcl::Registration (PCL >= 1.14, registration/registration.h + impl/registration.hpp): the
// public methods the reference's drivers call through the base pointer, the protected virtual computeTransformation that align()
// dispatches to, and the protected members a derived registrar reports through.  align() follows PCL's: the source is copied to
// `output`, final_/transformation_/previous_ are reset to identity, converged_ cleared, computeTransformation(output, guess) runs.
#pragma once
#include <pcl/point_cloud.h>
#include <string>
namespace pcl {
namespace mock {
struct Matrix4f {   // what the adapter needs of Eigen::Matrix4f: element access
  float m[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  float& operator()(int r, int c) { return m[4 * r + c]; }
  const float& operator()(int r, int c) const { return m[4 * r + c]; }
  static Matrix4f Identity() { return Matrix4f(); }
};
struct KdTree {};
}  // namespace mock

template <typename PointSource, typename PointTarget, typename Scalar = float>
class Registration {
public:
  using Matrix4 = mock::Matrix4f;
  using Ptr = shared_ptr<Registration<PointSource, PointTarget, Scalar>>;
  using PointCloudSource = pcl::PointCloud<PointSource>;
  using PointCloudSourcePtr = typename PointCloudSource::Ptr;
  using PointCloudSourceConstPtr = typename PointCloudSource::ConstPtr;
  using PointCloudTarget = pcl::PointCloud<PointTarget>;
  using PointCloudTargetConstPtr = typename PointCloudTarget::ConstPtr;
  using KdTreePtr = shared_ptr<mock::KdTree>;

  Registration() : tree_(new mock::KdTree) {}
  virtual ~Registration() = default;
  virtual void setInputSource(const PointCloudSourceConstPtr& cloud) { input_ = cloud; }
  virtual void setInputTarget(const PointCloudTargetConstPtr& cloud) { target_ = cloud; target_cloud_updated_ = true; }
  PointCloudTargetConstPtr const getInputTarget() { return target_; }
  void setSearchMethodTarget(const KdTreePtr& tree, bool force_no_recompute = false) { tree_ = tree; force_no_recompute_ = force_no_recompute; }
  KdTreePtr getSearchMethodTarget() const { return tree_; }
  Matrix4 getFinalTransformation() { return final_transformation_; }
  Matrix4 getLastIncrementalTransformation() { return transformation_; }
  void setMaximumIterations(int n) { max_iterations_ = n; }
  int getMaximumIterations() { return max_iterations_; }
  void setTransformationEpsilon(double e) { transformation_epsilon_ = e; }
  double getTransformationEpsilon() { return transformation_epsilon_; }
  bool hasConverged() const { return converged_; }
  void align(PointCloudSource& output) { align(output, Matrix4::Identity()); }
  void align(PointCloudSource& output, const Matrix4& guess) {
    if (!input_ || !target_) return;
    output = *input_;
    converged_ = false;
    final_transformation_ = transformation_ = previous_transformation_ = Matrix4::Identity();
    computeTransformation(output, guess);
  }
  const std::string& getClassName() const { return reg_name_; }

protected:
  std::string reg_name_;
  KdTreePtr tree_;
  int nr_iterations_ = 0, max_iterations_ = 10;
  PointCloudSourceConstPtr input_;
  PointCloudTargetConstPtr target_;
  Matrix4 final_transformation_, transformation_, previous_transformation_;
  double transformation_epsilon_ = 0.0;
  bool converged_ = false, target_cloud_updated_ = true, force_no_recompute_ = false;
  virtual void computeTransformation(PointCloudSource& output, const Matrix4& guess) = 0;
};
}  // namespace pcl
