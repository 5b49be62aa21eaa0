// TEST INFRASTRUCTURE -- a minimal stand-in for the PCL declarations include/ndtb_pcl_adapter.hpp touches, so that the adapter
// can be compiled and run in an image without PCL.  Layouts follow PCL's: PointXYZ 16 bytes, PointXYZI 32 bytes.
#pragma once
namespace pcl {
struct alignas(16) PointXYZ { float x = 0, y = 0, z = 0, pad_ = 1.0f; };
struct alignas(16) PointXYZI { float x = 0, y = 0, z = 0, pad_ = 1.0f; float intensity = 0, pad2_[3] = {0, 0, 0}; };
}  // namespace pcl
