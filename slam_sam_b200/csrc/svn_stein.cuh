// SVN-NDT on the device (svn_stein.cu): particle states, the K x K RBF-kernel / Stein variational Newton update, mean pose and
// stop test -- everything of svn_ndt::align (svn_ndt_impl.hpp:675-964) that is not the per-point derivative kernel.
#pragma once
#include "handle.cuh"

namespace ndtb {

// Window layout (ndtb_handle_s::SvnDev::d_window): flags [2][CAPACITY] u64 | records [2][CAPACITY][43] f64.
constexpr size_t SVN_WINDOW_FLAG_BYTES = 2 * static_cast<size_t>(ndtb_handle_s::SvnDev::CAPACITY) * sizeof(unsigned long long);
constexpr size_t SVN_WINDOW_BYTES = SVN_WINDOW_FLAG_BYTES + 2 * static_cast<size_t>(ndtb_handle_s::SvnDev::CAPACITY) * SVN_REC * sizeof(double);
inline unsigned long long* svn_window_flags(void* base) { return static_cast<unsigned long long*>(base); }
inline double* svn_window_recs(void* base) { return reinterpret_cast<double*>(static_cast<unsigned char*>(base) + SVN_WINDOW_FLAG_BYTES); }

int svn_dev_ensure(ndtb_handle_s* h);                                  // allocates the device-resident SVN state once
int svn_dev_upload_poses(ndtb_handle_s* h, int K, const double* poses);  // host K x 16 -> current particles

// Stage 1 on the handle's stream for the particles k = first + s * stride of the CURRENT device particles: state kernel +
// derivative kernel; the records land in the window half of a fresh stamp on `n_dest` ranks (1 = this handle only).
// Returns the stamp through *stamp_out.
// `vshards` > 1: the work units are (particle, point shard) pairs, unit u = first + s * stride evaluates shard u % vshards of particle
// u / vshards and stores a raw partial record in window row u; the Stein kernel adds a particle's shards in order.
int svn_dev_stage1(ndtb_handle_s* h, int K, int first, int stride, bool to_peers, unsigned long long* stamp_out, int vshards = 1);

// Stage 2 + 3 (+ mean pose / stop test when `with_mean`) on the records of `stamp`: waits for the K flags when `wait_flags`.
int svn_dev_stein(ndtb_handle_s* h, int K, unsigned long long stamp, bool wait_flags, bool write_poses, bool with_mean, int iter, int vshards = 1);

// Host-staged stage 2 (parity hook: ndtb_svn_stage2 / ndtb_svn_finish_step): poses, grads, hess from the host, updates (and the
// retracted particles when new_poses != nullptr) back to the host.  Same kernel as the device-resident loop.
int svn_stein_update(ndtb_handle_s* h, int K, const double* poses, const double* grads, const double* hess, double* updates,
                     double* new_poses);

}  // namespace ndtb
