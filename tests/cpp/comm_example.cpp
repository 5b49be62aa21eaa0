// Multi-GPU SVN from ONE process, as run/pipeline_lo_svn.cpp:299-319, 387-388 would drive it: one
// svn_ndt::SvnNormalDistributionsTransform per GPU, ndtb_comm_init over their handles, then the same align() on every registrar
// from one thread each (the call is collective).  Prints one line per rank: "rank converged iterations t03 t13 t23 checksum".
// Usage: comm_example n_gpus target.bin source.bin   (raw float32 xyzw records)
#include "ndtb_pclomp.hpp"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <thread>
#include <vector>

struct PointXYZ { float x, y, z, pad; };
using Cloud = ndtb_facade::PointCloud<PointXYZ>;
using Svn = svn_ndt::SvnNormalDistributionsTransform<PointXYZ, PointXYZ>;

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
  if (argc < 4) { std::fprintf(stderr, "usage: %s n_gpus target.bin source.bin\n", argv[0]); return 2; }
  const int n = std::atoi(argv[1]);
  auto target = load(argv[2]), source = load(argv[3]);
  std::vector<std::unique_ptr<Svn>> regs;
  std::vector<ndtb_handle> hs;
  for (int d = 0; d < n; ++d) {
    regs.emplace_back(new Svn(d));
    Svn& s = *regs.back();
    s.setResolution(1.0f); s.setMinPointPerVoxel(3); s.setNeighborhoodSearchMethod(svn_ndt::NeighborSearchMethod::DIRECT7);
    s.setParticleCount(16); s.setMaxIterations(30); s.setKernelBandwidth(1.0); s.setStepSize(1.0); s.setEarlyStopThreshold(1e-4);
    s.setInputTarget(target);
    hs.push_back(s.native());
  }
  if (n > 1 && ndtb_comm_init(hs.data(), n) != NDTB_OK) { std::fprintf(stderr, "ndtb_comm_init: %s\n", ndtb_last_error(hs[0])); return 1; }
  const ndtb_facade::Matrix4d prior = ndtb_facade::identity<ndtb_facade::Matrix4d>();
  std::vector<svn_ndt::SvnNdtResult> out(n);
  std::vector<std::thread> th;
  for (int r = 0; r < n; ++r) th.emplace_back([&, r] { out[r] = regs[r]->align(*source, prior, nullptr, /*seed=*/7); });
  for (auto& t : th) t.join();
  for (int r = 0; r < n; ++r) {
    double sum = 0;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) sum += out[r].final_pose(i, j) * (1 + 4 * i + j);
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) sum += out[r].final_covariance(i, j) * (17 + 6 * i + j);
    std::printf("%d %d %d %.17g %.17g %.17g %.17g\n", r, out[r].converged ? 1 : 0, out[r].iterations, out[r].final_pose(0, 3), out[r].final_pose(1, 3),
                out[r].final_pose(2, 3), sum);
  }
  return 0;
}
