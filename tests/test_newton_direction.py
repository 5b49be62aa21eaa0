"""Newton direction H^-1(-g) of ndtb_align (ndt_omp_impl.hpp:142-145: Eigen::JacobiSVD<6x6>::solve): the
pivoted-elimination path against the Jacobi SVD restatement and against numpy, and the fallback for a
(nearly) singular Hessian.  Host code of libndtb.so -- runs without a GPU."""
import numpy as np
import pytest

from slam_sam_b200 import ndtb


def ndt_like_hessian(rng, asym=1e-7):
    B = rng.standard_normal((6, 6))
    sc = np.array([1e2, 1e2, 1e2, 2e3 * (1 + rng.random()), 2e3, 3e3])
    H = -(B @ B.T + 0.3 * np.eye(6)) * np.outer(sc, sc)
    H = H * (1 + asym * rng.standard_normal((6, 6)))  # float-level asymmetry, as accumulated float terms have
    g = rng.standard_normal(6) * 1e3 * sc
    return H, g


def test_direct_matches_svd_and_numpy():
    rng = np.random.default_rng(5)
    worst_svd, worst_np, n_direct = 0.0, 0.0, 0
    for _ in range(500):
        H, g = ndt_like_hessian(rng)
        xs, used_s = ndtb.newton_direction(H, -g, method=0)
        xd, used_d = ndtb.newton_direction(H, -g, method=1)
        assert not used_s
        n_direct += used_d
        ref = np.linalg.solve(H.astype(np.longdouble).astype(np.float64), -g)
        worst_svd = max(worst_svd, np.linalg.norm(xs - xd) / np.linalg.norm(xs))
        worst_np = max(worst_np, np.linalg.norm(ref - xd) / np.linalg.norm(ref))
    assert n_direct >= 450          # the well-conditioned case is the common one
    assert worst_svd < 1e-10        # far inside the 1e-9 per-iteration parity bar
    assert worst_np < 1e-10


def test_rank_deficient_falls_back_to_svd():
    rng = np.random.default_rng(6)
    B = rng.standard_normal((6, 4))
    H = -(B @ B.T)                      # rank 4: JacobiSVD::solve gives the minimum-norm solution
    g = H @ rng.standard_normal(6)      # consistent right-hand side
    x, used = ndtb.newton_direction(H, g, method=1)
    assert not used
    np.testing.assert_allclose(x, np.linalg.pinv(H) @ g, rtol=1e-8, atol=1e-10)
    x0, _ = ndtb.newton_direction(H, g, method=0)
    np.testing.assert_array_equal(x, x0)


def test_zero_and_nan_hessian():
    x, used = ndtb.newton_direction(np.zeros((6, 6)), np.ones(6), method=1)
    assert not used and np.all(x == 0)        # rank 0: zero step, ndtb_align then stops (ndt_omp_impl.hpp:150-159)
    H = np.eye(6); H[2, 3] = np.nan
    x, used = ndtb.newton_direction(H, np.ones(6), method=1)
    x0, _ = ndtb.newton_direction(H, np.ones(6), method=0)
    assert not used                            # a non-finite H is never certified: same answer as the SVD path
    np.testing.assert_array_equal(x, x0)


def test_ill_conditioned_uses_svd():
    H = -np.diag([1.0, 1.0, 1.0, 1.0, 1.0, 1e-9])
    x, used = ndtb.newton_direction(H, np.ones(6), method=1)
    assert not used
    np.testing.assert_allclose(x, -1.0 / np.diag(-H) * 1.0, rtol=1e-12)


def test_hessian_to_covariance_matches_reference_formula():
    """run/pipeline_ligo_tc.cpp:536-542 + RegisterCallback::reorderCovarianceForGTSAM (src/registercallback.cpp:170-186)."""
    rng = np.random.default_rng(9)
    for _ in range(50):
        H, _ = ndt_like_hessian(rng)
        inv = -np.linalg.inv(H + 1e-6 * np.eye(6))
        want = np.zeros((6, 6))
        want[3:, 3:] = inv[:3, :3]; want[:3, :3] = inv[3:, 3:]; want[:3, 3:] = inv[:3, 3:]; want[3:, :3] = inv[3:, :3]
        got = ndtb.hessian_to_covariance(H)
        assert np.abs(got - want).max() <= 1e-9 * np.abs(want).max()
    assert np.isnan(ndtb.hessian_to_covariance(-1e-6 * np.eye(6))).all()   # H + 1e-6 I singular
