// Host-side cost of one Newton iteration between "results seen" and "next evaluation sent":
// svd6_solve + pose_to_T + angle tables.  nvcc -O3 -fmad=false -o /tmp/host_iter tools/micro/host_iter_cost.cu
#include "../../slam_sam_b200/csrc/ndtb_internal.cuh"
#include "../../slam_sam_b200/csrc/newton_math.cuh"
#include "../../slam_sam_b200/csrc/se3_math.cuh"
#include <chrono>
#include <cstdio>
using namespace ndtb;
int main() {
  double H[36], g[6], x[6];
  unsigned s = 12345;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (s >> 8) / double(1 << 24) - 0.5; };
  double acc = 0;
  const int R = 20000;
  for (int i = 0; i < 36; ++i) H[i] = 0;
  for (int i = 0; i < 6; ++i) for (int j = 0; j <= i; ++j) { double v = rnd() * 1e4; H[6 * i + j] = v; H[6 * j + i] = v; }
  for (int i = 0; i < 6; ++i) { H[7 * i] -= 5e4; g[i] = rnd() * 100; }
  auto t0 = std::chrono::steady_clock::now();
  for (int r = 0; r < R; ++r) { H[0] += 1e-3; svd6_solve(H, g, x); acc += x[0]; }
  auto t1 = std::chrono::steady_clock::now();
  {
    double xs[6], xg[6], worst = 0;
    auto t3 = std::chrono::steady_clock::now();
    for (int r = 0; r < R; ++r) { H[0] += 1e-3; bool ok = gecp6_solve(H, g, xg, 1e-6); acc += xg[0] + ok; }
    auto t4 = std::chrono::steady_clock::now();
    printf("gecp6_solve %.3f us\n", std::chrono::duration<double, std::micro>(t4 - t3).count() / R);
    for (int trial = 0; trial < 2000; ++trial) {   // NDT-like Hessians: translation block ~1e4, rotation block ~1e6..1e7, negative definite + noise
      double B[36], Hh[36];
      for (int i = 0; i < 36; ++i) B[i] = rnd();
      const double sc[6] = {1e2, 1e2, 1e2, 2e3 * (1 + rnd()), 2e3, 3e3};
      for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { double v = 0; for (int k = 0; k < 6; ++k) v += B[6 * i + k] * B[6 * j + k]; Hh[6 * i + j] = -(v + (i == j ? 0.3 : 0)) * sc[i] * sc[j] * (1 + 1e-7 * rnd()); }
      double gg[6]; for (int i = 0; i < 6; ++i) gg[i] = rnd() * 1e3 * sc[i];
      svd6_solve(Hh, gg, xs);
      if (!gecp6_solve(Hh, gg, xg, 1e-6)) { continue; }
      double dn = 0, nn = 0; for (int i = 0; i < 6; ++i) { dn += (xs[i] - xg[i]) * (xs[i] - xg[i]); nn += xs[i] * xs[i]; }
      worst = fmax(worst, sqrt(dn / nn));
    }
    printf("worst relative |x_svd - x_gecp| over 2000 NDT-like systems: %.3e\n", worst);
  }
  float T[16]; float ja[24], ha[45];
  double p[6] = {1, 2, 3, 0.01, 0.02, 0.5};
  for (int r = 0; r < R; ++r) { p[5] += 1e-6; pose_to_T(p, T); angle_tables(p, 10e-5, ja, ha); acc += T[0] + ja[3] + ha[7]; }
  auto t2 = std::chrono::steady_clock::now();
  printf("svd6_solve %.2f us, pose_to_T+angle_tables %.2f us (acc %g)\n", std::chrono::duration<double, std::micro>(t1 - t0).count() / R,
         std::chrono::duration<double, std::micro>(t2 - t1).count() / R, acc);
}
