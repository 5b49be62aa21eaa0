"""GPU parity, SVN flavour + full-size properties (through the C ABI)."""
import os

import numpy as np
import pytest

import oracle_lib as O
from conftest import rel_inf
from slam_sam_b200 import ndtb, registration, synth

pytestmark = pytest.mark.gpu
TOL = 1e-9
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ndt_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


def svn_handle(gold, gn=1, K=8, res=2.0, min_pts=3):
    h = ndtb.Handle(0)
    h.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=res, min_points_per_voxel=min_pts)
    h.set_svn_params(particle_count=K, use_gauss_newton=gn, kernel_bandwidth=1.0, step_size=1.0)
    assert h.set_target(gold["target"]) == 0 and h.set_source(gold["source"]) == 0
    return h


def test_svn_voxel_map_bit_exact(oracle, gold, street):
    h = svn_handle(gold)
    v = h.voxels()
    assert np.array_equal(v["idx"], gold["svn_idx"]) and np.array_equal(v["n"], gold["svn_n"])
    assert np.array_equal(v["mean"], gold["svn_mean"]) and np.array_equal(v["icov"], gold["svn_icov"])
    for min_pts in (3, 6):
        h.set_params(resolution=1.0, min_points_per_voxel=min_pts)
        assert h.set_target(street["target"]) == 0
        o = O.OracleGrid(1, 1.0, min_pts); o.build(street["target"])
        e, v = o.export(), h.voxels()
        assert np.array_equal(v["idx"], e["idx"]) and np.array_equal(v["n"], e["n"])
        assert np.array_equal(v["mean"], e["mean"]) and np.array_equal(v["icov"], e["icov"]) and np.array_equal(v["cov"], e["cov"])


@pytest.mark.parametrize("gn", [1, 0])
def test_svn_step_vs_golden(gold, gn):
    h = svn_handle(gold, gn)
    P0 = gold["svn_particles0"]
    s, g, H = h.svn_particle_derivatives(P0)
    assert rel_inf(s, gold[f"svn_step_gn{gn}_scores"]) <= TOL
    for k in range(len(P0)):
        assert rel_inf(g[k], gold[f"svn_step_gn{gn}_grads"][k]) <= TOL
        assert rel_inf(H[k], gold[f"svn_step_gn{gn}_hess"][k]) <= TOL
    u = h.svn_stage2(P0, gold[f"svn_step_gn{gn}_grads"], gold[f"svn_step_gn{gn}_hess"])
    assert rel_inf(u, gold[f"svn_step_gn{gn}_updates"]) <= TOL
    P1, s2, g2, H2, u2 = h.svn_step(P0)
    assert rel_inf(u2, gold[f"svn_step_gn{gn}_updates"]) <= 1e-8
    assert np.abs(P1 - gold[f"svn_step_gn{gn}_particles1"]).max() <= 1e-9
    # particle sharding (rank r evaluates particles k % world == r): the union equals the full stage 1
    rec = np.zeros((len(P0), 43))
    for r in range(3):
        part = h.svn_stage1_local(P0, r, 3)
        mine = np.arange(len(P0)) % 3 == r
        assert np.all(part[~mine] == 0)
        rec[mine] = part[mine]
    assert np.array_equal(rec[:, 0], s) and np.array_equal(rec[:, 1:7], g) and np.array_equal(rec[:, 7:].reshape(-1, 6, 6), H)
    P1b, ub = h.svn_finish_step(P0, rec)
    assert np.array_equal(P1b, P1) and np.array_equal(ub, u2)


def test_reference_gtest_case_svn_k10(oracle, two_plane):
    """Reads like TEST(ConvergenceComparison, SvnNdtK10), test_svn_ndt.cpp:205-257."""
    ndt = registration.SvnNormalDistributionsTransform()
    ndt.setResolution(1.0)
    ndt.setMinPointPerVoxel(3)
    ndt.setNeighborhoodSearchMethod(registration.DIRECT7)
    ndt.setNumThreads(20)
    ndt.setParticleCount(10)
    ndt.setMaxIterations(100)
    ndt.setKernelBandwidth(1.0)
    ndt.setEarlyStopThreshold(1e-4)
    ndt.setStepSize(1.0)
    ndt.setInputTarget(two_plane["target"])
    P0 = O.OracleSvn.sample_particles(two_plane["guess"], 10, seed=3)   # the reference seeds from the clock (:712); 63-108 iterations over seeds
    result = ndt.align(two_plane["source"], two_plane["guess"], init_particles=P0)
    assert result.converged
    assert result.iterations < 100
    err = np.ascontiguousarray(np.linalg.inv(result.final_pose) @ two_plane["gt"])
    xi = np.zeros(6); oracle.lib().orc_pose_logmap(O._d(err), O._d(xi))
    assert np.linalg.norm(xi[3:]) < 0.05 and np.linalg.norm(xi[:3]) < 0.035
    o = O.OracleSvn(resolution=1.0, min_pts=3, num_threads=8, K=10, max_iter=100, kernel_h=1.0, stop_thresh=1e-4, step_size=1.0)
    o.set_target(two_plane["target"]); o.set_source(two_plane["source"])
    ro = o.align(two_plane["guess"], P0)
    assert result.iterations == ro["iterations"]
    assert np.abs(result.final_pose - ro["final_pose"]).max() <= 1e-7
    assert np.abs(result.final_covariance - ro["final_covariance"]).max() <= 1e-9 + 1e-6 * np.abs(ro["final_covariance"]).max()
    # seeded sampling path (no particles given) also converges to the same basin
    r2 = ndt.align(two_plane["source"], two_plane["guess"], seed=42)
    # (the mean-pose stop test may or may not fire within max_iterations for an arbitrary seed: the reference's
    #  own log never converges in 100 iterations; the basin is what is checked)
    assert r2.iterations <= 100 and np.abs(r2.final_pose[:3, 3] - two_plane["gt"][:3, 3]).max() < 0.05


def test_svn_invalid_inputs_follow_reference(two_plane):
    """svn_ndt_impl.hpp:682-702: invalid inputs -> converged=false, final_pose=prior, covariance=I."""
    ndt = registration.SvnNormalDistributionsTransform()
    prior = two_plane["guess"]
    r = ndt.align(two_plane["source"], prior)                 # no target set
    assert not r.converged and np.array_equal(r.final_pose, prior) and np.array_equal(r.final_covariance, np.eye(6))
    ndt.setInputTarget(two_plane["target"])
    r = ndt.align(np.zeros((0, 4), np.float32), prior)        # empty source
    assert not r.converged and np.array_equal(r.final_pose, prior)
    ndt.setParticleCount(0)
    r = ndt.align(two_plane["source"], prior)
    assert not r.converged and np.array_equal(r.final_covariance, np.eye(6))


def test_full_size_properties():
    """BASELINE sizes (131 072-point scan vs 1 M-point map): size-independent properties."""
    scene = synth.Scene()
    tgt = synth.surface_map(scene, 1_000_000, seed=7)
    pose = synth.trajectory_pose(60.0)
    src = synth.ouster_scan(scene, pose, seed=3)
    assert src.shape == (131072, 4)
    h = ndtb.Handle(0)
    h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01)
    assert h.set_target(tgt) == 0
    v = h.voxels()
    assert np.all(np.diff(v["idx"]) > 0)                                   # sortedness / uniqueness of linear indices
    assert 0.99 * len(tgt) <= int(v["n"].clip(0).sum()) <= len(tgt)         # every point lands in exactly one leaf
    guess = (pose @ synth.pose_matrix((0.15, -0.1, 0.02), np.deg2rad((0.2, -0.2, 0.8)))).astype(np.float32)
    p = np.array([guess[0, 3], guess[1, 3], guess[2, 3], 0.003, -0.003, 0.014 + 0.03 * np.sin(0.05 * 60.0)])
    # (1) determinism: same inputs -> same bits
    h.set_source(src)
    a = h.compute_derivatives(p)
    b = h.compute_derivatives(p)
    assert a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    # (2) additivity over the source cloud: f(A u B) = f(A) + f(B)
    h.set_source(src[:70000]); s1, g1, H1, _ = h.compute_derivatives(p)
    h.set_source(src[70000:]); s2, g2, H2, _ = h.compute_derivatives(p)
    assert abs((s1 + s2) - a[0]) <= 1e-11 * abs(a[0])
    assert rel_inf(g1 + g2, a[1]) <= 1e-10 and rel_inf(H1 + H2, a[2]) <= 1e-10
    # (3) permutation invariance of the source up to FP64 summation order
    rng = np.random.default_rng(0)
    h.set_source(src[rng.permutation(len(src))]); s3, g3, H3, _ = h.compute_derivatives(p)
    assert abs(s3 - a[0]) <= 1e-11 * abs(a[0]) and rel_inf(g3, a[1]) <= 1e-10 and rel_inf(H3, a[2]) <= 1e-10
    # (4) the map does not depend on the order of the target points either (occupancy exact)
    h2 = ndtb.Handle(0); h2.set_params(resolution=1.0)
    assert h2.set_target(tgt[rng.permutation(len(tgt))]) == 0
    v2 = h2.voxels()
    assert np.array_equal(v2["idx"], v["idx"]) and np.array_equal(v2["n"], v["n"])
    assert np.abs(v2["mean"] - v["mean"]).max() <= 1e-9
    # (5) align recovers the pose, twice with the same bits
    h.set_source(src)
    r1 = h.align(guess); r2 = h.align(guess)
    assert r1["converged"] and np.array_equal(r1["T"], r2["T"]) and r1["iterations"] == r2["iterations"]
    assert np.abs(r1["T"][:3, 3] - pose[:3, 3]).max() < 0.05


def test_svn_kdtree_matches_oracle(oracle, street, gold):
    """svn_ndt::NeighborSearchMethod::KDTREE (svn_ndt_impl.hpp:576-579): radiusSearch(x_trans, resolution) over the cloud of
    float(leaf.mean_) of the valid leaves (voxel_grid_covariance_impl.hpp:383-436, 505-555) = 27 hash probes + FLANN's float
    distance test on the device.  Per-particle derivatives and a whole K = 6 align against the oracle."""
    for tgt, src, res in ((gold["target"], gold["source"], 2.0), (street["target"], street["source"], 1.0)):
        h = ndtb.Handle(0)
        assert h.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=res, min_points_per_voxel=3, search_method=ndtb.KDTREE) == 0
        h.set_svn_params(particle_count=6, max_iterations=8, use_gauss_newton=1, kernel_bandwidth=1.0, step_size=1.0)
        assert h.set_target(tgt) == 0 and h.set_source(src) == 0
        o = O.OracleSvn(resolution=res, min_pts=3, search="KDTREE", num_threads=8, K=6, max_iter=8, kernel_h=1.0, step_size=1.0)
        o.set_target(tgt); o.set_source(src)
        P0 = O.OracleSvn.sample_particles(np.eye(4), 6, seed=11)
        s, g, H = h.svn_particle_derivatives(P0)
        hits_kd = h.last_hits()[0]
        for k in range(6):
            so, go, Ho = o.particle_derivatives(P0[k])
            assert abs(s[k] - so) <= TOL * abs(so) and rel_inf(g[k], go) <= TOL and rel_inf(H[k], Ho) <= TOL
        h7 = ndtb.Handle(0)
        h7.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=res, min_points_per_voxel=3, search_method=ndtb.DIRECT7)
        h7.set_target(tgt); h7.set_source(src); h7.svn_particle_derivatives(P0)
        assert hits_kd != h7.last_hits()[0]                       # a different neighbourhood than DIRECT7
        rg, ro = h.svn_align(np.eye(4), P0), o.align(np.eye(4), P0)
        assert rg["iterations"] == ro["iterations"]
        assert np.abs(rg["particles"] - ro["particles"]).max() <= 1e-8
        assert np.abs(rg["final_pose"] - ro["final_pose"]).max() <= 1e-8
