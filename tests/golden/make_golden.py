"""Generates tests/golden/ndt_golden.npz from the ORACLE (the reference itself cannot be built or
imported here: C++ templates over PCL/Eigen/GTSAM, none of which are in the image).

    python tests/golden/make_golden.py

Inputs are stored with the outputs so the fixture does not depend on the numpy RNG stream.
Contents (small 16 x 256-ray street scan pair, 4096 points each):
  target, source                      float32 (N,4)
  pclomp_* / svn_*                    voxel tables (idx, n, mean, icov) of both grid flavours
  p_eval, T_eval, score/g/H/nvtl      pclomp computeDerivatives at 3 poses (DIRECT7 and DIRECT1)
  align_*                             Newton trajectory (trace, transformation_array, final T)
  svn_particles0, svn_step_*          one SVN iteration from stored particles (K = 8, GN and full)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as O  # noqa: E402
from slam_sam_b200 import synth  # noqa: E402


def main():
    scene = synth.Scene()
    A = synth.trajectory_pose(40.0)
    B = A @ synth.pose_matrix((0.3, 0.05, 0.01), np.deg2rad((0.3, 0.2, 1.0)))
    tgt = synth.ouster_scan(scene, A, beams=16, cols=256, seed=11)
    src = synth.ouster_scan(scene, B, beams=16, cols=256, seed=12)
    out = dict(target=tgt, source=src)

    for search in ("DIRECT7", "DIRECT1"):
        ndt = O.OracleNdt(resolution=2.0, search=search, max_iterations=35, trans_epsilon=0.01, step_size=0.1, num_threads=1)
        assert ndt.set_target(tgt) == 0
        ndt.set_source(src)
        if search == "DIRECT7":
            g = ndt.grid()
            out.update(pclomp_idx=g["idx"], pclomp_n=g["n"], pclomp_mean=g["mean"], pclomp_icov=g["icov"], pclomp_cov=g["cov"],
                       pclomp_min_b=g["min_b"], pclomp_div_b=g["div_b"])
        r = ndt.align()
        out[f"align_{search}_trace"] = r["trace"]
        out[f"align_{search}_tarray"] = r["transformation_array"]
        out[f"align_{search}_T"] = r["T"]
        out[f"align_{search}_iters"] = np.array([r["iterations"], int(r["converged"])])
        out[f"align_{search}_hessian"] = r["hessian"]
        poses = [np.zeros(6), np.array([0.3, 0.05, 0.01, 0.005, 0.003, 0.017]), r["trace"][-1, :6]]
        res = [ndt.compute_derivatives(p) for p in poses]
        out[f"eval_{search}_p"] = np.array(poses)
        out[f"eval_{search}_score"] = np.array([x[0] for x in res])
        out[f"eval_{search}_g"] = np.array([x[1] for x in res])
        out[f"eval_{search}_H"] = np.array([x[2] for x in res])
        out[f"eval_{search}_nvtl"] = np.array([x[3] for x in res])

    # regulariser (ndt_omp_impl.hpp:349-382)
    ndt = O.OracleNdt(resolution=2.0, reg_scale=10.0)
    ndt.set_target(tgt); ndt.set_source(src)
    regT = synth.pose_matrix((0.5, -0.2, 0.0), (0, 0, 0.02)).astype(np.float32)
    ndt.set_regularization_pose(regT)
    s, g, H, nv = ndt.compute_derivatives(np.array([0.3, 0.05, 0.01, 0.005, 0.003, 0.017]))
    out.update(reg_T=regT, reg_score=np.array([s]), reg_g=g, reg_H=H)

    prior = np.linalg.inv(A) @ B
    K = 8
    P0 = O.OracleSvn.sample_particles(prior, K, seed=7)
    out["svn_prior"] = prior
    out["svn_particles0"] = P0
    for gn in (1, 0):
        svn = O.OracleSvn(resolution=2.0, min_pts=3, K=K, use_gn=bool(gn), kernel_h=1.0, step_size=1.0)
        assert svn.set_target(tgt) == 0
        svn.set_source(src)
        if gn:
            g = svn.grid()
            out.update(svn_idx=g["idx"], svn_n=g["n"], svn_mean=g["mean"], svn_icov=g["icov"])
        P1, s, gr, H, u = svn.step(P0)
        out[f"svn_step_gn{gn}_particles1"] = P1
        out[f"svn_step_gn{gn}_scores"] = s
        out[f"svn_step_gn{gn}_grads"] = gr
        out[f"svn_step_gn{gn}_hess"] = H
        out[f"svn_step_gn{gn}_updates"] = u
    np.savez_compressed(os.path.join(HERE, "ndt_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "ndt_golden.npz"), {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
