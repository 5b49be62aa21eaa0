// K x K RBF-kernel / Stein variational Newton update on the device (svn_stein.cu).
#pragma once
#include "handle.cuh"

namespace ndtb {
// Stage 2 (+ stage 3 when new_poses != nullptr) of svn_ndt::align (svn_ndt_impl.hpp:789-855).
// poses: K x 16 doubles (host), grads/hess in NDT order (host) -> updates[6K] (GTSAM order) and,
// optionally, the retracted particles new_poses[16K].
int svn_stein_update(ndtb_handle_s* h, int K, const double* poses, const double* grads, const double* hess, double* updates,
                     double* new_poses);
}  // namespace ndtb
