"""The oracle's analytic gradient / Hessian against finite differences of its own score (CPU).
This ties the restated float pair maths (ndt_omp_impl.hpp:640-697) to the objective it
differentiates (Magnusson 2009, eq 6.9-6.13) independently of any golden vector."""
import numpy as np

import oracle_lib as O


def fd_setup(street, search, res=2.0):
    ndt = O.OracleNdt(resolution=res, search=search, num_threads=8)
    assert ndt.set_target(street["target"]) == 0
    ndt.set_source(street["source"])
    return ndt


P0 = np.array([0.45, 0.0, 0.3, 0.01, 0.08, 0.25])   # near the two-plane ground truth


def test_gradient_matches_finite_differences(oracle, two_plane):
    ndt = fd_setup(two_plane, "DIRECT7", 1.0)
    p0 = P0
    s0, g, H, _ = ndt.compute_derivatives(p0)
    # the score is piecewise smooth (points change voxels) and evaluated in float: use a step large
    # enough to beat float noise and compare to a few percent
    for i, h in enumerate([1e-3] * 6):
        dp = np.zeros(6); dp[i] = h
        sp = ndt.compute_derivatives(p0 + dp)[0]
        sm = ndt.compute_derivatives(p0 - dp)[0]
        fd = (sp - sm) / (2 * h)
        assert abs(fd - g[i]) <= 0.04 * abs(g).max(), (i, fd, g[i])


def test_hessian_matches_gradient_differences(oracle, two_plane):
    ndt = fd_setup(two_plane, "DIRECT7", 1.0)
    p0 = P0
    _, g0, H, _ = ndt.compute_derivatives(p0)
    Hs = (H + H.T) / 2
    assert np.abs(H - H.T).max() <= 1e-5 * np.abs(H).max()     # symmetric up to float rounding only
    for i, h in enumerate([1e-3] * 6):
        dp = np.zeros(6); dp[i] = h
        gp = ndt.compute_derivatives(p0 + dp)[1]
        gm = ndt.compute_derivatives(p0 - dp)[1]
        fd = (gp - gm) / (2 * h)
        assert np.abs(fd - Hs[i]).max() <= 0.1 * np.abs(Hs).max(), (i, fd, Hs[i])


def test_direct1_is_subset_of_direct7(oracle, street):
    n7 = fd_setup(street, "DIRECT7"); n1 = fd_setup(street, "DIRECT1")
    p = np.zeros(6)
    s7, _, _, nv7 = n7.compute_derivatives(p)
    s1, _, _, nv1 = n1.compute_derivatives(p)
    assert 0 < s1 < s7            # every DIRECT1 pair is also a DIRECT7 pair and each pair's score is >= 0
    assert nv1 > 0 and nv7 > 0    # NVTL is a mean over points WITH neighbours, so it is not monotone in the probe count


def test_no_hessian_mode_keeps_score_and_gradient(oracle, street):
    ndt = fd_setup(street, "DIRECT7")
    p = np.array([0.1, 0.0, 0.0, 0.0, 0.0, 0.01])
    s, g, H, _ = ndt.compute_derivatives(p, compute_hessian=True)
    s2, g2, H2, _ = ndt.compute_derivatives(p, compute_hessian=False)
    assert s == s2 and np.array_equal(g, g2) and np.all(H2 == 0) and np.any(H != 0)


def test_gauss_constants(oracle):
    # Appendix A of SURVEY.md / ndt_omp_impl.hpp:95-99 with r = 0.55, res = 1: c1 = 4.5, c2 = 0.55
    d1, d2, d3 = O.OracleNdt(resolution=1.0).gauss()
    c1, c2 = 4.5, 0.55
    assert np.isclose(d3, -np.log(c2)) and np.isclose(d1, -np.log(c1 + c2) - d3) and d1 < 0
    assert np.isclose(d2, -2 * np.log((-np.log(c1 * np.exp(-0.5) + c2) - d3) / d1))
    assert np.allclose(O.OracleSvn(resolution=1.0).gauss(), [d1, d2, d3])
