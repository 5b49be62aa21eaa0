// The registrar held and driven through the pcl::Registration BASE pointer, exactly as the reference does it
// (include/registercallback.hpp:35 `pcl::Registration<PointXYZI,PointXYZI>::Ptr registration`; run/pipeline.cpp:460-481 builds a
// pclomp::NormalDistributionsTransform and assigns it; run/pipeline.cpp:557-566 calls setInputTarget / setInputSource / align /
// getFinalTransformation through the base and getResult() through the derived pointer).  Compiled against include/ndtb_pcl_adapter.hpp
// with tests/cpp/mock_pcl/ standing in for PCL's headers (this image has no PCL).  Usage: adapter_example target.bin source.bin g00..g33
#include <ndtb_pcl_adapter.hpp>

#include <cstdio>
#include <cstdlib>
#include <fstream>

using Cloud = pcl::PointCloud<pcl::PointXYZI>;

struct RegisterCallback {  // the member of the reference's class, type unchanged
  pcl::Registration<pcl::PointXYZI, pcl::PointXYZI>::Ptr registration;
};

static Cloud::Ptr load(const char* path) {  // raw float32 xyzw records -> PointXYZI (32-byte stride)
  Cloud::Ptr c(new Cloud());
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  if (!f) return c;
  const std::size_t n = static_cast<std::size_t>(f.tellg()) / 16;
  std::vector<float> raw(4 * n);
  f.seekg(0);
  f.read(reinterpret_cast<char*>(raw.data()), static_cast<std::streamsize>(16 * n));
  c->points.resize(n);
  for (std::size_t i = 0; i < n; ++i) { c->points[i].x = raw[4 * i]; c->points[i].y = raw[4 * i + 1]; c->points[i].z = raw[4 * i + 2]; c->points[i].intensity = 42.0f; }
  return c;
}

int main(int argc, char** argv) {
  if (argc < 19) { std::fprintf(stderr, "usage: %s target.bin source.bin g00 .. g33\n", argv[0]); return 2; }
  RegisterCallback registerCallback;
  // run/pipeline.cpp:460-481
  pclomp::NormalDistributionsTransform<pcl::PointXYZI, pcl::PointXYZI>::Ptr ndt_omp(new pclomp::NormalDistributionsTransform<pcl::PointXYZI, pcl::PointXYZI>());
  ndt_omp->setNumThreads(8);
  ndt_omp->setResolution(1.0f);
  ndt_omp->setTransformationEpsilon(1e-4);
  ndt_omp->setMaximumIterations(50);
  ndt_omp->setStepSize(0.1);
  ndt_omp->setNeighborhoodSearchMethod(pclomp::DIRECT7);
  registerCallback.registration = ndt_omp;
  // run/pipeline.cpp:557-566
  Cloud::Ptr lidarFactorPointsTarget = load(argv[1]), pointsBody = load(argv[2]);
  Cloud::Ptr lidarFactorPointsSource(new Cloud());
  pcl::Registration<pcl::PointXYZI, pcl::PointXYZI>::Matrix4 guess;
  for (int i = 0; i < 16; ++i) guess(i / 4, i % 4) = std::strtof(argv[3 + i], nullptr);
  registerCallback.registration->setInputTarget(lidarFactorPointsTarget);
  registerCallback.registration->setInputSource(pointsBody);
  registerCallback.registration->align(*lidarFactorPointsSource, guess);
  const auto T = registerCallback.registration->getFinalTransformation();
  auto ndt_result = ndt_omp->getResult();
  std::printf("%d %d", registerCallback.registration->hasConverged() ? 1 : 0, ndt_result.iteration_num);
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) std::printf(" %.9g", static_cast<double>(T(i, j)));
  const auto& p0 = lidarFactorPointsSource->points[0];
  std::printf(" %zu %.9g %.9g %.9g %.9g %zu\n", lidarFactorPointsSource->points.size(), static_cast<double>(p0.x), static_cast<double>(p0.y),
              static_cast<double>(p0.z), static_cast<double>(p0.intensity), ndt_omp->getLeaves().size());
  return 0;
}
