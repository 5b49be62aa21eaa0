// TEST INFRASTRUCTURE -- stand-in for pcl::PointCloud (see point_types.h).
#pragma once
#include <cstddef>
#include <memory>
#include <vector>
namespace pcl {
template <class T> using shared_ptr = std::shared_ptr<T>;
template <class PointT>
class PointCloud {
public:
  using Ptr = shared_ptr<PointCloud<PointT>>;
  using ConstPtr = shared_ptr<const PointCloud<PointT>>;
  std::vector<PointT> points;
  unsigned width = 0, height = 1;
  bool is_dense = true;
  std::size_t size() const { return points.size(); }
  void push_back(const PointT& p) { points.push_back(p); width = static_cast<unsigned>(points.size()); }
};
}  // namespace pcl
