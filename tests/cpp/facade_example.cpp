// Reads like TEST(ConvergenceComparison, PclOmp) of extern/svn_ndt/test/test_svn_ndt.cpp:138-199, written against
// include/ndtb_pclomp.hpp.  Usage: facade_example target.bin source.bin  (raw float32 xyzw records);
// prints "converged iterations t00..t33".  Built (and, on a GPU box, run) by tests/test_facade_cpp.py.
#include "ndtb_pclomp.hpp"
#include <cstdio>
#include <cstdlib>
#include <fstream>

struct PointXYZ { float x, y, z, pad; };  // layout of pcl::PointXYZ
using Cloud = ndtb_facade::PointCloud<PointXYZ>;

static Cloud::Ptr load(const char* path) {
  auto c = std::make_shared<Cloud>();
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  if (!f) return c;
  const std::size_t bytes = static_cast<std::size_t>(f.tellg());
  c->points.resize(bytes / sizeof(PointXYZ));
  f.seekg(0);
  f.read(reinterpret_cast<char*>(c->points.data()), static_cast<std::streamsize>(c->points.size() * sizeof(PointXYZ)));
  return c;
}

int main(int argc, char** argv) {
  if (argc < 3) { std::fprintf(stderr, "usage: %s target.bin source.bin\n", argv[0]); return 2; }
  pclomp::NormalDistributionsTransform<PointXYZ, PointXYZ> ndt;
  ndt.setResolution(1.0f);
  ndt.setNeighborhoodSearchMethod(pclomp::DIRECT7);
  ndt.setMaximumIterations(50);
  ndt.setTransformationEpsilon(1e-4);
  ndt.setStepSize(0.1);
  ndt.setNumThreads(20);
  ndt.setInputTarget(load(argv[1]));
  ndt.setInputSource(load(argv[2]));
  Cloud aligned;
  ndtb_facade::Matrix4f guess = ndtb_facade::identity<ndtb_facade::Matrix4f>();
  if (argc >= 19) for (int i = 0; i < 16; ++i) guess(i / 4, i % 4) = std::strtof(argv[3 + i], nullptr);
  ndt.align(aligned, guess);
  const auto T = ndt.getFinalTransformation();
  const pclomp::NdtResult r = ndt.getResult();
  std::printf("%d %d", ndt.hasConverged() ? 1 : 0, ndt.getFinalNumIteration());
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) std::printf(" %.9g", static_cast<double>(T(i, j)));
  std::printf(" %zu %zu %.9g\n", aligned.points.size(), ndt.getLeaves().size(), static_cast<double>(r.transform_probability));
  if (argc >= 20) {  // the svn facade, same clouds (svn_ndt.h:100-182); K = 4 keeps it short
    svn_ndt::SvnNormalDistributionsTransform<PointXYZ, PointXYZ> svn;
    svn.setResolution(1.0f);
    svn.setMinPointPerVoxel(3);
    svn.setNeighborhoodSearchMethod(svn_ndt::NeighborSearchMethod::DIRECT7);
    svn.setParticleCount(4);
    svn.setMaxIterations(5);
    svn.setInputTarget(load(argv[1]));
    ndtb_facade::Matrix4d prior = ndtb_facade::identity<ndtb_facade::Matrix4d>();
    for (int i = 0; i < 16; ++i) prior(i / 4, i % 4) = guess(i / 4, i % 4);
    const svn_ndt::SvnNdtResult s = svn.align(*load(argv[2]), prior, nullptr, 7);
    std::printf("svn %d %d %.9g\n", s.converged ? 1 : 0, s.iterations, s.final_pose(0, 3));
    // the same pclomp alignment with the target assembled on the device from an archived keyframe (identity pose)
    pclomp::NormalDistributionsTransform<PointXYZ, PointXYZ> win;
    win.setResolution(1.0f); win.setMaximumIterations(50); win.setTransformationEpsilon(1e-4); win.setStepSize(0.1);
    win.putKeyframe(7, *load(argv[1]));
    win.setInputTargetWindow(std::vector<std::int64_t>{7}, std::vector<ndtb_facade::Matrix4d>{ndtb_facade::identity<ndtb_facade::Matrix4d>()});
    win.setInputSource(load(argv[2]));
    Cloud aligned2;
    win.align(aligned2, guess);
    const auto Tw = win.getFinalTransformation();
    std::printf("win %d %d", win.hasConverged() ? 1 : 0, win.getFinalNumIteration());
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) std::printf(" %.9g", static_cast<double>(Tw(i, j)));
    std::printf("\n");
    // multigrid NDT: the target handed over as two tiles (even / odd points), as a dynamic map loader would
    pclomp::MultiGridNormalDistributionsTransform<PointXYZ, PointXYZ> mg;
    mg.setResolution(1.0f); mg.setMaximumIterations(50); mg.setTransformationEpsilon(1e-4); mg.setStepSize(0.1);
    auto whole = load(argv[1]);
    auto t0 = std::make_shared<Cloud>(), t1 = std::make_shared<Cloud>();
    for (std::size_t i = 0; i < whole->points.size(); ++i) (whole->points[i].x < 0.f ? t0 : t1)->points.push_back(whole->points[i]);
    mg.addTarget(t0, "west"); mg.addTarget(t1, "east"); mg.addTarget(t1, "stale"); mg.removeTarget("stale");
    mg.createVoxelKdtree();
    mg.setInputSource(load(argv[2]));
    Cloud aligned3;
    mg.align(aligned3, guess);
    const auto Tm = mg.getFinalTransformation();
    std::printf("mg %d %d", mg.hasConverged() ? 1 : 0, mg.getFinalNumIteration());
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) std::printf(" %.9g", static_cast<double>(Tm(i, j)));
    std::printf(" %zu %zu %.9g\n", mg.getCurrentMapIDs().size(), mg.getVoxelPCD().points.size(), mg.calculateNearestVoxelTransformationLikelihood(Tm));
  }
  return 0;
}
