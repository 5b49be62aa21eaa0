// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points (ctypes) for the SVN-NDT restatement.
#include "svn_ndt.hpp"

using namespace oracle;

extern "C" {

void* orc_svn_create() { return new SvnNdt(); }
void orc_svn_destroy(void* h) { delete static_cast<SvnNdt*>(h); }
void* orc_svn_grid(void* h) { return &static_cast<SvnNdt*>(h)->grid; }
void orc_svn_set_params(void* h, float resolution, double outlier_ratio, int search_method, int num_threads, int K, int max_iter,
                        double kernel_h, double step_size, double stop_thresh, int use_gn, int min_pts, double eig_mult) {
  SvnNdt* s = static_cast<SvnNdt*>(h);
  s->resolution = resolution; s->outlier_ratio = outlier_ratio; s->search_method = search_method; s->num_threads = num_threads;
  s->K = K; s->max_iter = max_iter; s->kernel_h = kernel_h; s->step_size = step_size; s->stop_thresh = stop_thresh; s->use_gn = use_gn != 0;
  s->grid.min_points_per_voxel = min_pts < 3 ? 3 : min_pts;  // voxel_grid_covariance.h:176-184
  s->grid.min_covar_eigvalue_mult = eig_mult;
  s->update_constants();
}
void orc_svn_gauss(void* h, double* d) { SvnNdt* s = static_cast<SvnNdt*>(h); d[0] = s->gauss_d1; d[1] = s->gauss_d2; d[2] = s->gauss_d3; }
int orc_svn_set_target(void* h, const float* xyz, size_t n, size_t sf) { return static_cast<SvnNdt*>(h)->set_target(xyz, n, sf); }
void orc_svn_set_source(void* h, const float* xyz, size_t n, size_t sf) { static_cast<SvnNdt*>(h)->set_source(xyz, n, sf); }

// prior: 4x4 row-major double.  particles_out: K x 16 doubles.
void orc_svn_sample_particles(const double* prior16, int K, unsigned long long seed, double* particles_out) {
  Pose3 prior; pose_from_matrix16(prior16, prior);
  std::vector<Pose3> P; SvnNdt::sample_particles(prior, K, seed, P);
  for (int k = 0; k < K; ++k) pose_to_matrix16(P[k], particles_out + 16 * k);
}
// Stage 1 for one particle pose (4x4 row-major double): score, g[6], H[36] in NDT order.
double orc_svn_particle_derivatives(void* h, const double* pose16, double* g, double* H) {
  Pose3 P; pose_from_matrix16(pose16, P);
  return static_cast<SvnNdt*>(h)->stage1(P, g, H);
}
// One SVN iteration on K particles (in/out K x 16), returning scores[K], grads[6K], hess[36K], updates[6K].
void orc_svn_step(void* h, int K, double* particles, double* scores, double* grads, double* hess, double* updates) {
  std::vector<Pose3> P(K);
  for (int k = 0; k < K; ++k) pose_from_matrix16(particles + 16 * k, P[k]);
  std::vector<double> s, g, H, u;
  static_cast<SvnNdt*>(h)->step(P, s, g, H, u);
  for (int k = 0; k < K; ++k) pose_to_matrix16(P[k], particles + 16 * k);
  if (scores) std::memcpy(scores, s.data(), 8 * K);
  if (grads) std::memcpy(grads, g.data(), 48 * K);
  if (hess) std::memcpy(hess, H.data(), 288 * K);
  if (updates) std::memcpy(updates, u.data(), 48 * K);
}
// Stage 2 only (parity hook for the K x K Stein kernel).
void orc_svn_stage2(void* h, int K, const double* particles, const double* grads, const double* hess, double* updates) {
  std::vector<Pose3> P(K);
  for (int k = 0; k < K; ++k) pose_from_matrix16(particles + 16 * k, P[k]);
  std::vector<double> g(grads, grads + 6 * K), H(hess, hess + 36 * K), u;
  static_cast<SvnNdt*>(h)->stage2(P, g, H, u);
  std::memcpy(updates, u.data(), 48 * K);
}
// Full align.  particles: K x 16 in/out.  Returns iterations; fills final pose (16), covariance (36, GTSAM order).
int orc_svn_align(void* h, const double* prior16, int K, double* particles, double* final_pose16, double* cov36, int* converged) {
  Pose3 prior; pose_from_matrix16(prior16, prior);
  std::vector<Pose3> P(K);
  for (int k = 0; k < K; ++k) pose_from_matrix16(particles + 16 * k, P[k]);
  SvnResult r = static_cast<SvnNdt*>(h)->align(prior, P);
  for (int k = 0; k < K; ++k) pose_to_matrix16(P[k], particles + 16 * k);
  pose_to_matrix16(r.final_pose, final_pose16);
  std::memcpy(cov36, r.final_covariance, 288);
  *converged = r.converged ? 1 : 0;
  return r.iterations;
}
// SE(3) probes for tests
void orc_pose_expmap(const double* xi, double* M16) { pose_to_matrix16(pose_expmap(xi), M16); }
void orc_pose_logmap(const double* M16, double* xi) { Pose3 p; pose_from_matrix16(M16, p); pose_logmap(p, xi); }
void orc_rot_rpy(const double* M16, double* rpy) { Pose3 p; pose_from_matrix16(M16, p); rot_rpy(p.R, rpy); }

}  // extern "C"
