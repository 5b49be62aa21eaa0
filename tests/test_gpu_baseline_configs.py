"""GPU parity AT THE SIZES BASELINE.json NAMES (configs[0], [1], [2]) -- the CUDA path through the C ABI against the CPU oracle.

The small-cloud tests (test_gpu_pclomp.py, test_gpu_svn.py) pin the arithmetic; these pin the same bars on the workloads the
bench measures: 131 072-point OS1-128-shaped scans, the 1 M-point map of bench.make_workload, K = 100 SVN particles.
Bars (BASELINE.json north_star): voxel occupancy / indices bit-exact (means, covariances, eigen-systems too), per-evaluation
score / gradient / Hessian <= 1e-9 relative (inf-norm), iteration counts equal, final poses within 1e-4 m and 1e-5 rad.
The oracle restates the reference with left-to-right float sums; Eigen's SSE packet reductions associate (a0+a2)+(a1+a3)
(oracle/ndt_pclomp.hpp:12-13), so "1e-9" is CUDA == oracle, and oracle == reference is pinned by the README fitness values."""
import numpy as np
import pytest

import oracle_lib as O
from conftest import rel_inf
from slam_sam_b200 import ndtb, synth

pytestmark = pytest.mark.gpu
TOL = 1e-9


def pose_delta(Ta, Tb):
    d = np.linalg.inv(np.asarray(Ta, float)) @ np.asarray(Tb, float)
    ang = np.arccos(np.clip((np.trace(d[:3, :3]) - 1) / 2, -1, 1))
    return np.linalg.norm(d[:3, 3]), ang


def assert_maps_identical(v, e):
    for k in ("min_b", "max_b", "div_b", "idx", "n"):
        assert np.array_equal(v[k], e[k]), k
    ok = e["n"] >= 1
    assert np.array_equal(v["mean"][ok], e["mean"][ok])
    for k in ("cov", "icov", "evals", "evecs"):
        assert np.array_equal(v[k], e[k]), k


def assert_traces_match(tg, to):
    assert tg.shape == to.shape
    worst = 0.0
    for a, b in zip(tg, to):
        assert np.abs(a[:6] - b[:6]).max() <= 1e-9                      # pose vector of the evaluation
        es = abs(a[6] - b[6]) / abs(b[6])
        eg = rel_inf(a[7:13], b[7:13]); eh = rel_inf(a[13:49], b[13:49])
        worst = max(worst, es, eg, eh)
        assert es <= TOL and eg <= TOL and eh <= TOL
        assert abs(a[49] - b[49]) <= TOL * abs(b[49])                   # NVTL
    return worst


@pytest.fixture(scope="module")
def scan_pair():
    """configs[0]: one synthetic OS1-128 scan pair, 131 072 points each."""
    scene = synth.Scene()
    A = synth.trajectory_pose(40.0)
    B = A @ synth.pose_matrix((0.5, 0.1, 0.02), np.deg2rad((0.5, 0.3, 2.0)))
    return dict(target=synth.ouster_scan(scene, A, seed=1), source=synth.ouster_scan(scene, B, seed=2), rel=np.linalg.inv(A) @ B)


@pytest.mark.parametrize("search", ["DIRECT7", "DIRECT1"])
def test_config0_scan_pair_131k(oracle, scan_pair, search):
    """configs[0]: ndt_omp align of one 131k scan pair, 1.0 m voxels, identity guess."""
    tgt, src = scan_pair["target"], scan_pair["source"]
    assert tgt.shape == (131072, 4) and src.shape == (131072, 4)
    kw = dict(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
    h = ndtb.Handle(0)
    h.set_params(search_method=getattr(ndtb, search), **kw)
    assert h.set_target(tgt) == 0 and h.set_source(src) == 0
    o = O.OracleNdt(search=search, num_threads=16, **kw)
    assert o.set_target(tgt) == 0
    o.set_source(src)
    assert_maps_identical(h.voxels(), o.grid())
    rg, ro = h.align(), o.align()
    assert (rg["iterations"], rg["converged"]) == (ro["iterations"], ro["converged"])
    assert rg["iterations"] >= 5                                         # 0.5 m offset, 0.1 m step clamp
    assert_traces_match(h.trace(), ro["trace"])
    dt, da = pose_delta(rg["T"], ro["T"])
    assert dt <= 1e-4 and da <= 1e-5
    assert rel_inf(rg["hessian"], ro["hessian"]) <= TOL
    dt, da = pose_delta(rg["T"], scan_pair["rel"])                       # and it is the right answer
    assert dt < 0.05 and da < 0.01


def test_config1_map_align_at_bench_size(oracle):
    """configs[1], the bench workload itself: 1 M-point map build + align of 131 072-point scans of bench.make_workload(0)."""
    import bench

    tgt, scans, guesses, truths = bench.make_workload(0)
    assert tgt.shape[0] == 1_000_000 and scans[0].shape == (131072, 4)
    kw = dict(resolution=bench.RESOLUTION, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
    h = ndtb.Handle(0)
    h.set_params(search_method=ndtb.DIRECT7, **kw)
    assert h.set_target(tgt) == 0
    o = O.OracleNdt(num_threads=16, **kw)
    assert o.set_target(tgt) == 0
    assert_maps_identical(h.voxels(), o.grid())                          # ~84 k leaves, every field bit for bit
    for k in (0, 3):
        assert h.set_source(scans[k]) == 0
        o.set_source(scans[k])
        rg, ro = h.align(guesses[k]), o.align(guesses[k])
        assert (rg["iterations"], rg["converged"]) == (ro["iterations"], ro["converged"]) and rg["converged"]
        assert_traces_match(h.trace(), ro["trace"])                      # EVERY evaluation of the Newton loop
        dt, da = pose_delta(rg["T"], ro["T"])
        assert dt <= 1e-4 and da <= 1e-5
        assert rel_inf(rg["hessian"], ro["hessian"]) <= TOL
        assert np.abs(rg["T"][:3, 3] - truths[k][:3, 3]).max() < 0.05


def test_config2_svn_k100_131k(oracle, scan_pair):
    """configs[2]: SVN-NDT, K = 100 particles on the 131k scan pair, min 3 points per voxel, Gauss-Newton Hessian, particles from
    seed 7 with the sigmas of svn_ndt_impl.hpp:710.  Three full iterations (stages 1-3, :758-855) against the oracle."""
    import bench

    K = 100
    tgt, src = scan_pair["target"], scan_pair["source"]
    h = ndtb.Handle(0)
    h.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=1.0, min_points_per_voxel=3, search_method=ndtb.DIRECT7)
    h.set_svn_params(particle_count=K, max_iterations=50, kernel_bandwidth=1.0, step_size=1.0, stop_threshold=1e-4, use_gauss_newton=1)
    assert h.set_target(tgt) == 0 and h.set_source(src) == 0
    o = O.OracleSvn(resolution=1.0, min_pts=3, num_threads=16, K=K, max_iter=50, kernel_h=1.0, stop_thresh=1e-4, step_size=1.0)
    assert o.set_target(tgt) == 0
    o.set_source(src)
    e, v = o.grid(), h.voxels()
    for k in ("idx", "n", "mean", "cov", "icov"):
        assert np.array_equal(v[k], e[k]), k
    P0 = bench.svn_initial_particles(K)
    Pg, Po = P0.copy(), P0.copy()
    for it in range(3):
        Pg1, sg, gg, Hg, ug = h.svn_step(Pg)
        Po1, so, go, Ho, uo = o.step(Po)
        for k in range(K):
            assert abs(sg[k] - so[k]) <= TOL * abs(so[k])
            assert rel_inf(gg[k], go[k]) <= TOL and rel_inf(Hg[k], Ho[k]) <= TOL
        assert rel_inf(ug, uo) <= TOL
        assert np.abs(Pg1 - Po1).max() <= 1e-9
        Pg, Po = Po1, Po1                                               # both sides continue from identical particles
