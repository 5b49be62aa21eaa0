"""The oracle must reproduce the committed golden vectors bit for bit (CPU).  The goldens were
generated from the oracle by tests/golden/make_golden.py; this guards the oracle against drift, and
the GPU tests compare the CUDA path with the same file."""
import os

import numpy as np
import pytest

import oracle_lib as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ndt_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


def test_voxel_tables(oracle, gold):
    ndt = O.OracleNdt(resolution=2.0)
    assert ndt.set_target(gold["target"]) == 0
    g = ndt.grid()
    assert np.array_equal(g["idx"], gold["pclomp_idx"]) and np.array_equal(g["n"], gold["pclomp_n"])
    assert np.array_equal(g["mean"], gold["pclomp_mean"]) and np.array_equal(g["icov"], gold["pclomp_icov"])
    svn = O.OracleSvn(resolution=2.0, min_pts=3)
    assert svn.set_target(gold["target"]) == 0
    g = svn.grid()
    assert np.array_equal(g["idx"], gold["svn_idx"]) and np.array_equal(g["n"], gold["svn_n"])
    assert np.array_equal(g["mean"], gold["svn_mean"]) and np.array_equal(g["icov"], gold["svn_icov"])
    # svn erases invalid leaves, pclomp keeps them flagged: the valid sets differ by the min-points rule only
    assert (gold["svn_n"] >= 3).all() and (gold["pclomp_n"] < 6).any()


@pytest.mark.parametrize("search", ["DIRECT7", "DIRECT1"])
def test_derivatives_and_align(oracle, gold, search):
    ndt = O.OracleNdt(resolution=2.0, search=search, max_iterations=35, trans_epsilon=0.01, step_size=0.1, num_threads=1)
    ndt.set_target(gold["target"]); ndt.set_source(gold["source"])
    for k, p in enumerate(gold[f"eval_{search}_p"]):
        s, g, H, nv = ndt.compute_derivatives(p)
        assert s == gold[f"eval_{search}_score"][k] and nv == gold[f"eval_{search}_nvtl"][k]
        assert np.array_equal(g, gold[f"eval_{search}_g"][k]) and np.array_equal(H, gold[f"eval_{search}_H"][k])
    r = ndt.align()
    assert [r["iterations"], int(r["converged"])] == list(gold[f"align_{search}_iters"])
    assert np.array_equal(r["T"], gold[f"align_{search}_T"])
    assert np.array_equal(r["trace"], gold[f"align_{search}_trace"])
    assert np.array_equal(r["transformation_array"], gold[f"align_{search}_tarray"])
    # thread count must not change the result: per-point partials + serial in-order sum (ndt_omp_impl.hpp:340-347)
    ndt8 = O.OracleNdt(resolution=2.0, search=search, max_iterations=35, trans_epsilon=0.01, step_size=0.1, num_threads=8)
    ndt8.set_target(gold["target"]); ndt8.set_source(gold["source"])
    assert np.array_equal(ndt8.align()["trace"], gold[f"align_{search}_trace"])


def test_regulariser(oracle, gold):
    ndt = O.OracleNdt(resolution=2.0, reg_scale=10.0)
    ndt.set_target(gold["target"]); ndt.set_source(gold["source"])
    ndt.set_regularization_pose(gold["reg_T"])
    s, g, H, _ = ndt.compute_derivatives(np.array([0.3, 0.05, 0.01, 0.005, 0.003, 0.017]))
    assert s == gold["reg_score"][0] and np.array_equal(g, gold["reg_g"]) and np.array_equal(H, gold["reg_H"])
    ndt.set_regularization_pose(None)
    s0, g0, H0, _ = ndt.compute_derivatives(np.array([0.3, 0.05, 0.01, 0.005, 0.003, 0.017]))
    assert s0 == gold["eval_DIRECT7_score"][1]
    assert s != s0 and not np.array_equal(g[:2], g0[:2]) and np.array_equal(g[2:], g0[2:])


@pytest.mark.parametrize("gn", [1, 0])
def test_svn_step(oracle, gold, gn):
    svn = O.OracleSvn(resolution=2.0, min_pts=3, K=8, use_gn=bool(gn), kernel_h=1.0, step_size=1.0)
    svn.set_target(gold["target"]); svn.set_source(gold["source"])
    P1, s, g, H, u = svn.step(gold["svn_particles0"])
    assert np.array_equal(P1, gold[f"svn_step_gn{gn}_particles1"])
    assert np.array_equal(s, gold[f"svn_step_gn{gn}_scores"]) and np.array_equal(g, gold[f"svn_step_gn{gn}_grads"])
    assert np.array_equal(H, gold[f"svn_step_gn{gn}_hess"]) and np.array_equal(u, gold[f"svn_step_gn{gn}_updates"])
    # stage 2 alone reproduces the same updates from the stored stage-1 outputs
    assert np.array_equal(svn.stage2(gold["svn_particles0"], g, H), u)
