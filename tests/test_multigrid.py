"""Multigrid NDT (pclomp::MultiGridNormalDistributionsTransform, extern/ndt_omp/include/multigrid_pclomp/).

CPU: the restatement (oracle/multigrid_ndt.hpp) against the facts the reference fixes --
  * one grid through setInputTarget lands on the README's KDTREE / pcl::NDT fitness of the bundled scan pair
    (0.213937, extern/ndt_omp/README.md:19; the README states KDTREE == pcl::NDT, and the multigrid class is the
    same radius-search NDT in FP64),
  * the radius search is exactly "all searchable leaves whose float centroid is closer than the resolution",
  * the analytic derivatives are the derivatives of its own score on frozen associations,
  * addTarget / removeTarget / createVoxelKdtree semantics.
GPU: the CUDA path (ndtb_mg_*, slam_sam_b200/csrc/multigrid.cu) against that restatement: leaves bit-exact,
score / gradient / Hessian <= 1e-9 relative per evaluation, same trajectory and pose."""
import os

import numpy as np
import pytest

import oracle_lib as O
from conftest import rel_inf
from slam_sam_b200 import ndtb

FIX = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ndt_omp_pair.npz")
TOL = 1e-9


@pytest.fixture(scope="module")
def pair():
    return dict(np.load(FIX))


def fitness(pair, T):
    from scipy.spatial import cKDTree

    tree = cKDTree(pair["target"].astype(np.float64))
    T = np.asarray(T, np.float32)
    q = (pair["source"] @ T[:3, :3].T + T[:3, 3]).astype(np.float64)
    dist, _ = tree.query(q)
    return float((dist ** 2).mean())


def tiles(cloud, nx=2, ny=2):
    """Split a cloud into nx x ny map tiles by x / y quantiles (how a dynamic map loader hands the map over)."""
    c = np.asarray(cloud)
    qx = np.quantile(c[:, 0], np.linspace(0, 1, nx + 1)); qy = np.quantile(c[:, 1], np.linspace(0, 1, ny + 1))
    qx[-1] += 1; qy[-1] += 1
    out = {}
    for i in range(nx):
        for j in range(ny):
            m = (c[:, 0] >= qx[i]) & (c[:, 0] < qx[i + 1]) & (c[:, 1] >= qy[j]) & (c[:, 1] < qy[j + 1])
            if m.sum():
                out["tile_%d_%d" % (i, j)] = np.ascontiguousarray(c[m])
    return out


@pytest.mark.parametrize("threads", [1, 8])
def test_oracle_single_grid_reproduces_the_readme_kdtree_fitness(oracle, pair, threads):
    o = O.OracleMultigrid(resolution=1.0, num_threads=threads)       # apps/align.cpp:107-109: defaults + setResolution(1.0)
    assert o.set_input_target(pair["target"]) == 0
    o.set_source(pair["source"])
    r = o.align(None)
    assert r["converged"] and r["iterations"] < 35
    assert abs(fitness(pair, r["T"]) - 0.213937) < 5e-7


def test_oracle_radius_search_is_exhaustive(oracle, pair):
    o = O.OracleMultigrid(resolution=1.0)
    for k, c in tiles(pair["target"]).items():
        assert o.add_target(c, k) == 0
    o.create_kdtree()
    L = o.leaves()
    assert (L["n"] >= 6).all() and len(L["n"]) > 500
    r2 = np.float32(1.0)
    rng = np.random.default_rng(0)
    for q in pair["source"][rng.integers(0, len(pair["source"]), 300), :3]:
        q = q.astype(np.float32)
        d = ((q[None, :] - L["centroid"]) ** 2).astype(np.float32)
        dist = (d[:, 0] + d[:, 1]) + d[:, 2]                       # L2_Simple<float>: x, y, z accumulated in order
        want = np.flatnonzero(dist < r2)
        means, got_d = o.radius_search(q)
        assert len(means) == len(want)
        order = want[np.argsort(dist[want], kind="stable")]
        assert np.array_equal(np.sort(got_d), got_d)               # FLANN returns the hits sorted by distance
        assert np.array_equal(means, L["mean"][order]) or set(map(tuple, means)) == set(map(tuple, L["mean"][want]))


def test_oracle_tiles_share_no_leaves_and_change_little(oracle, pair):
    whole = O.OracleMultigrid(resolution=1.0, num_threads=8); whole.set_input_target(pair["target"]); whole.set_source(pair["source"])
    parts = O.OracleMultigrid(resolution=1.0, num_threads=8)
    for k, c in tiles(pair["target"]).items():
        parts.add_target(c, k)
    with pytest.raises(RuntimeError):
        parts.set_source(pair["source"]); parts.align(None)         # kd-tree not built yet
    parts.create_kdtree()
    # a voxel cut by a tile border becomes two leaves (or drops below 6 points): counts differ, the basin does not
    assert abs(len(parts.leaves()["n"]) - len(whole.leaves()["n"])) < 0.1 * len(whole.leaves()["n"])
    a, b = whole.align(None), parts.align(None)
    assert a["converged"] and b["converged"]
    assert np.abs(a["T"][:3, 3] - b["T"][:3, 3]).max() < 0.05
    # removing a tile and rebuilding the tree really removes its leaves
    n_before = len(parts.leaves()["n"])
    parts.remove_target("tile_0_0"); parts.create_kdtree()
    assert len(parts.leaves()["n"]) < n_before
    parts.remove_target("no_such_tile")                             # removeCloud ignores unknown ids (:183-185)


def test_oracle_derivatives_match_finite_differences(oracle, pair):
    """The analytic FP64 gradient / Hessian against finite differences of the score, on the source points whose
    neighbourhood is frozen over the stencil: a leaf enters a point's neighbourhood when its centroid comes within one
    resolution, where its (identity-inflated) Gaussian is far from zero -- a jump the analytic derivatives do not
    contain (13 % of the z gradient on the full cloud).  Points with a centroid within 1 cm of the search sphere are
    left out (h = 1e-4 moves a point by < 5 mm)."""
    from scipy.spatial import cKDTree

    from slam_sam_b200 import synth

    o = O.OracleMultigrid(resolution=1.0, num_threads=8)
    o.set_input_target(pair["target"])
    p0 = np.array([0.3, -0.2, 0.05, 0.004, -0.006, 0.02])
    T = synth.pose_matrix(p0[:3], p0[3:])
    q = pair["source"][:, :3].astype(float) @ T[:3, :3].T + T[:3, 3]
    cen = o.leaves()["centroid"].astype(float)
    tree = cKDTree(cen)
    keep = np.array([not any(abs(np.linalg.norm(cen[j] - x) - 1.0) < 0.01 for j in tree.query_ball_point(x, 1.02)) for x in q])
    assert keep.sum() > 0.8 * len(q)
    o.set_source(np.ascontiguousarray(pair["source"][keep]))
    s0, g, H, _ = o.compute_derivatives(p0)
    Hs = (H + H.T) / 2
    assert np.abs(H - H.T).max() <= 1e-9 * np.abs(H).max()
    h = 1e-4
    for i in range(6):
        dp = np.zeros(6); dp[i] = h
        sp, gp, _, _ = o.compute_derivatives(p0 + dp)
        sm, gm, _, _ = o.compute_derivatives(p0 - dp)
        assert abs((sp - sm) / (2 * h) - g[i]) <= 2e-3 * np.abs(g).max(), i
        assert np.abs((gp - gm) / (2 * h) - Hs[i]).max() <= 5e-3 * np.abs(Hs).max(), i


# ------------------------------------------------------------------------------------------------ GPU
def gpu_handle(**kw):
    h = ndtb.Handle(0)
    h.set_params(resolution=1.0, step_size=0.1, trans_epsilon=0.1, max_iterations=35, **kw)
    return h


@pytest.mark.gpu
def test_gpu_multigrid_leaves_bit_exact(oracle, pair):
    h = gpu_handle()
    o = O.OracleMultigrid(resolution=1.0)
    for k, c in tiles(pair["target"], 3, 2).items():
        assert h.mg_add_target(c, k) == 0 and o.add_target(c, k) == 0
    assert not h.mg_info()["current"]
    assert h.mg_create_kdtree() == 0
    o.create_kdtree()
    info = h.mg_info()
    assert info["current"] and info["grids"] == 6 and info["grids_in_tree"] == 6
    assert h.mg_grid_ids() == sorted(tiles(pair["target"], 3, 2).keys())
    a, b = h.mg_leaves(), o.leaves()
    assert info["leaves"] == len(b["n"])
    assert np.array_equal(a["n"], b["n"]) and np.array_equal(a["mean"], b["mean"]) and np.array_equal(a["icov"], b["icov"])
    assert np.array_equal(a["centroid"], b["centroid"])


@pytest.mark.gpu
@pytest.mark.parametrize("split", [(1, 1), (2, 2), (4, 3)])
def test_gpu_multigrid_derivatives_match_oracle(oracle, pair, split):
    h = gpu_handle(); o = O.OracleMultigrid(resolution=1.0, num_threads=8)
    for k, c in tiles(pair["target"], *split).items():
        assert h.mg_add_target(c, k) == 0 and o.add_target(c, k) == 0
    h.mg_create_kdtree(); o.create_kdtree()
    assert h.set_source(pair["source"]) == 0
    o.set_source(pair["source"])
    for p in (np.zeros(6), np.array([0.3, -0.2, 0.05, 0.004, -0.006, 0.02]), np.array([1.5, 0.4, -0.1, 0.01, 0.02, -0.05])):
        for hess in (True, False):
            s, g, H, nv = h.mg_compute_derivatives(p, compute_hessian=hess)
            so, go, Ho, nvo = o.compute_derivatives(p, compute_hessian=hess)
            assert abs(s - so) <= TOL * abs(so) and rel_inf(g, go) <= TOL and abs(nv - nvo) <= TOL * abs(nvo)
            if hess:
                assert rel_inf(H, Ho) <= TOL
            else:
                assert np.all(H == 0) and np.all(Ho == 0)


@pytest.mark.gpu
@pytest.mark.parametrize("split,ls", [((1, 1), False), ((2, 2), False), ((2, 2), True)])
def test_gpu_multigrid_align_matches_oracle(oracle, pair, split, ls):
    h = gpu_handle(use_line_search=int(ls)); o = O.OracleMultigrid(resolution=1.0, num_threads=8, use_line_search=ls)
    for k, c in tiles(pair["target"], *split).items():
        assert h.mg_add_target(c, k) == 0 and o.add_target(c, k) == 0
    h.mg_create_kdtree(); o.create_kdtree()
    h.set_source(pair["source"]); o.set_source(pair["source"])
    guess = np.eye(4, dtype=np.float32)
    if ls:
        guess[0, 3] = 1.0; guess[1, 3] = -0.6
        h.set_params(step_size=1.0); o = O.OracleMultigrid(resolution=1.0, num_threads=8, use_line_search=True, step_size=1.0)
        for k, c in tiles(pair["target"], *split).items():
            o.add_target(c, k)
        o.create_kdtree(); o.set_source(pair["source"])
    r, ro = h.mg_align(guess), o.align(guess)
    assert (r["iterations"], r["converged"]) == (ro["iterations"], ro["converged"])
    tr, tg = h.trace(), ro["trace"]
    assert tr.shape == tg.shape
    for a, b in zip(tr, tg):
        assert np.abs(a[:6] - b[:6]).max() <= 1e-9
        assert abs(a[6] - b[6]) <= TOL * abs(b[6]) and rel_inf(a[7:13], b[7:13]) <= TOL and rel_inf(a[13:49], b[13:49]) <= TOL
    d = np.linalg.inv(r["T"].astype(float)) @ ro["T"].astype(float)
    assert np.linalg.norm(d[:3, 3]) <= 1e-4 and np.arccos(np.clip((np.trace(d[:3, :3]) - 1) / 2, -1, 1)) <= 1e-5
    assert rel_inf(r["hessian"], ro["hessian"]) <= TOL
    if split == (1, 1) and not ls:
        assert abs(fitness(pair, r["T"]) - 0.213937) < 5e-7        # the README's KDTREE / pcl::NDT value, on the device


@pytest.mark.gpu
def test_gpu_multigrid_dynamic_map_update(oracle, pair):
    """The dynamic-map-loading cycle: add tiles, build the tree, align; drop one tile and add another, rebuild, align."""
    T = tiles(pair["target"], 2, 2)
    keys = sorted(T)
    h = gpu_handle(); o = O.OracleMultigrid(resolution=1.0, num_threads=8)
    h.set_source(pair["source"]); o.set_source(pair["source"])
    for k in keys[:3]:
        h.mg_add_target(T[k], k); o.add_target(T[k], k)
    h.mg_create_kdtree(); o.create_kdtree()
    a, b = h.mg_align(None), o.align(None)
    assert a["iterations"] == b["iterations"] and np.abs(a["T"] - b["T"]).max() <= 1e-6
    # stale tree: evaluation refused until the tree is rebuilt
    assert h.mg_remove_target(keys[0]) == 0 and h.mg_add_target(T[keys[3]], keys[3]) == 0
    with pytest.raises(ndtb.NdtbError):
        h.mg_align(None)
    o.remove_target(keys[0]); o.add_target(T[keys[3]], keys[3])
    h.mg_create_kdtree(); o.create_kdtree()
    assert h.mg_grid_ids() == keys[1:]
    a2, b2 = h.mg_align(None), o.align(None)
    assert a2["iterations"] == b2["iterations"] and np.abs(a2["T"] - b2["T"]).max() <= 1e-6
    assert np.array_equal(h.mg_leaves()["mean"], o.leaves()["mean"])
    # replacing a tile under the same id (the reference orphans the old node)
    half = T[keys[1]][::2]
    h.mg_add_target(half, keys[1]); o.add_target(half, keys[1]); h.mg_create_kdtree(); o.create_kdtree()
    assert np.array_equal(h.mg_leaves()["n"], o.leaves()["n"])
    # regularisation pose goes through the same host epilogue
    reg = np.eye(4, dtype=np.float32); reg[0, 3] = 0.4
    h.set_params(regularization_scale_factor=0.01); h.set_regularization_pose(reg)
    o2 = O.OracleMultigrid(resolution=1.0, num_threads=8, reg_scale=0.01)
    for k in keys[1:]:
        o2.add_target(half if k == keys[1] else T[k], k)
    o2.create_kdtree(); o2.set_source(pair["source"]); o2.set_regularization_pose(reg)
    p = np.array([0.2, 0.1, 0.0, 0.0, 0.0, 0.3])
    s, g, H, _ = h.mg_compute_derivatives(p); so, go, Ho, _ = o2.compute_derivatives(p)
    assert abs(s - so) <= TOL * abs(so) and rel_inf(g, go) <= TOL and rel_inf(H, Ho) <= TOL


@pytest.mark.gpu
def test_gpu_multigrid_facade_reads_like_the_reference(oracle, pair):
    """The call sequence of a dynamic-map localiser on the reference class, on the Python mirror of it."""
    from slam_sam_b200 import registration

    ndt = registration.MultiGridNormalDistributionsTransform()
    ndt.setResolution(1.0); ndt.setStepSize(0.1); ndt.setTransformationEpsilon(0.01); ndt.setMaximumIterations(30)
    ndt.setNumThreads(8)
    with pytest.raises(ndtb.NdtbError):
        ndt.setNeighborhoodSearchMethod(registration.DIRECT7)
    T = tiles(pair["target"], 2, 2)
    for k, c in T.items():
        ndt.addTarget(c, k)
    ndt.createVoxelKdtree()
    assert ndt.getCurrentMapIDs() == sorted(T)
    ndt.setInputSource(pair["source"])
    ndt.align(None)
    assert ndt.hasConverged() and ndt.getFinalNumIteration() < 30
    o = O.OracleMultigrid(resolution=1.0, step_size=0.1, trans_epsilon=0.01, max_iterations=30, num_threads=8)
    for k, c in T.items():
        o.add_target(c, k)
    o.create_kdtree(); o.set_source(pair["source"])
    ro = o.align(None)
    assert ndt.getFinalNumIteration() == ro["iterations"]
    assert np.abs(ndt.getFinalTransformation() - ro["T"]).max() <= 1e-6
    assert abs(ndt.getTransformationProbability() - ro["trans_probability"]) <= TOL * abs(ro["trans_probability"])
    assert abs(ndt.getNearestVoxelTransformationLikelihood() - ro["nvtl"]) <= TOL * abs(ro["nvtl"])
    assert rel_inf(ndt.getHessian(), ro["hessian"]) <= TOL
    assert len(ndt.getFinalTransformationArray()) == len(ro["transformation_array"])
    tp, nv = o.calculate_scores(ro["T"])
    assert abs(ndt.calculateTransformationProbability(ro["T"]) - tp) <= TOL * abs(tp)
    assert abs(ndt.calculateNearestVoxelTransformationLikelihood(ro["T"]) - nv) <= TOL * abs(nv)
    assert ndt.getVoxelPCD().shape == (len(o.leaves()["n"]), 3)
    ndt.removeTarget(sorted(T)[0]); ndt.createVoxelKdtree()
    assert ndt.getCurrentMapIDs() == sorted(T)[1:]
    # setInputTarget == addTarget("default") + createVoxelKdtree
    one = registration.MultiGridNormalDistributionsTransform(); one.setResolution(1.0)
    one.setInputTarget(pair["target"]); one.setInputSource(pair["source"])
    assert one.getCurrentMapIDs() == ["default"]
    one.align(None)
    assert abs(fitness(pair, one.getFinalTransformation()) - 0.213937) < 5e-7
