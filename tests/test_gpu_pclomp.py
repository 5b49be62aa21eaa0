"""GPU parity, pclomp flavour: CUDA path (through the C ABI) vs the CPU oracle and the golden file.

Bars (BASELINE.json north_star): voxel occupancy and indices bit-exact (we also get means / inverse
covariances bit-exact); per-evaluation score, gradient, Hessian <= 1e-9 relative (to the inf-norm);
final poses within 1e-4 m and 1e-5 rad."""
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


def assert_grid_equal(a, b, exact_stats=True):
    assert np.array_equal(a["min_b"], b["min_b"]) and np.array_equal(a["max_b"], b["max_b"]) and np.array_equal(a["div_b"], b["div_b"])
    assert np.array_equal(a["idx"], b["idx"])      # occupancy + linear indices: bit-exact
    assert np.array_equal(a["n"], b["n"])
    valid = b["n"] >= 1
    if exact_stats:
        assert np.array_equal(a["mean"][valid], b["mean"][valid])
        assert np.array_equal(a["icov"], b["icov"]) and np.array_equal(a["cov"], b["cov"])
        assert np.array_equal(a["evals"], b["evals"])
        assert np.array_equal(a["evecs"], b["evecs"])      # Leaf::evecs_ (exported to ndt_ellipsoids.txt, include/pipeline.hpp:163-222): same Jacobi sweeps, same signs


def pose_delta(Ta, Tb):
    d = np.linalg.inv(np.asarray(Ta, float)) @ np.asarray(Tb, float)
    ang = np.arccos(np.clip((np.trace(d[:3, :3]) - 1) / 2, -1, 1))
    return np.linalg.norm(d[:3, 3]), ang


@pytest.mark.parametrize("res", [0.5, 1.0, 2.0])
def test_voxel_map_bit_exact(oracle, street, two_plane, res):
    for cloud in (street["target"], two_plane["target"]):
        h = ndtb.Handle(0)
        h.set_params(resolution=res)
        assert h.set_target(cloud) == 0
        o = O.OracleGrid(0, res, 6); assert o.build(cloud) == 0
        assert_grid_equal(h.voxels(), o.export())


def test_voxel_map_golden(gold):
    h = ndtb.Handle(0); h.set_params(resolution=2.0)
    assert h.set_target(gold["target"]) == 0
    v = h.voxels()
    assert np.array_equal(v["idx"], gold["pclomp_idx"]) and np.array_equal(v["n"], gold["pclomp_n"])
    assert np.array_equal(v["mean"], gold["pclomp_mean"]) and np.array_equal(v["icov"], gold["pclomp_icov"])


def test_voxel_map_edge_cases(oracle):
    h = ndtb.Handle(0)
    assert h.L.ndtb_set_target(h.h, None, 0, 16) == -3                # NDTB_ERR_EMPTY (:60-65)
    big = np.array([[0, 0, 0, 0], [4000, 4000, 4000, 0]], np.float32)
    assert h.set_target(big) == -4                                    # NDTB_ERR_OVERFLOW (:83-94)
    rng = np.random.default_rng(0)
    c = rng.uniform(0, 3, (500, 4)).astype(np.float32)
    c[5, 0] = np.nan; c[17, 2] = np.inf
    h.set_params(is_dense=0)
    assert h.set_target(c) == 0
    o = O.OracleGrid(0, 1.0, 6); o.build(c, is_dense=False)
    assert_grid_equal(h.voxels(), o.export())
    # identical points: zero scatter, but the reference's second moment starts at identity (Leaf::cov_ = I,
    # voxel_grid_covariance_omp.h:113), so the leaf is kept with cov = I (n-1)/n^2
    same = np.tile(np.array([[0.5, 0.5, 0.5, 0]], np.float32), (10, 1))
    h.set_params(is_dense=1)
    assert h.set_target(same) == 0
    o = O.OracleGrid(0, 1.0, 6); o.build(same)
    v = h.voxels(); assert list(v["n"]) == [10]
    assert_grid_equal(v, o.export())
    np.testing.assert_allclose(v["icov"][0].reshape(3, 3), np.eye(3) * 100 / 9, rtol=1e-12)
    # strides: packed xyz (12 B), PointXYZ (16 B), PointXYZI (32 B) give the same map
    base = rng.uniform(-5, 5, (3000, 3)).astype(np.float32)
    ref = None
    for width in (3, 4, 8):
        c = np.zeros((3000, width), np.float32); c[:, :3] = base
        if width > 3:
            c[:, 3:] = 7.0
        assert h.set_target(c) == 0
        v = h.voxels()
        if ref is None:
            ref = v
        else:
            assert_grid_equal(v, ref)
    # single point, single voxel
    assert h.set_target(np.array([[1.5, 2.5, 3.5, 0]], np.float32)) == 0 and list(h.voxels()["n"]) == [1]


@pytest.mark.parametrize("search", ["DIRECT7", "DIRECT1"])
def test_derivatives_vs_golden_and_oracle(oracle, gold, search):
    h = ndtb.Handle(0)
    h.set_params(resolution=2.0, search_method=getattr(ndtb, search))
    assert h.set_target(gold["target"]) == 0 and h.set_source(gold["source"]) == 0
    for k, p in enumerate(gold[f"eval_{search}_p"]):
        s, g, H, nv = h.compute_derivatives(p)
        assert abs(s - gold[f"eval_{search}_score"][k]) <= TOL * abs(gold[f"eval_{search}_score"][k])
        assert rel_inf(g, gold[f"eval_{search}_g"][k]) <= TOL and rel_inf(H, gold[f"eval_{search}_H"][k]) <= TOL
        assert abs(nv - gold[f"eval_{search}_nvtl"][k]) <= TOL * abs(nv)
        s2, g2, H2, _ = h.compute_derivatives(p, compute_hessian=False)
        assert s2 == s and np.array_equal(g2, g) and np.all(H2 == 0)
        # run-to-run bit stability (fixed reduction tree, no FP64 atomics)
        s3, g3, H3, _ = h.compute_derivatives(p)
        assert s3 == s and np.array_equal(g3, g) and np.array_equal(H3, H)


def test_regulariser_and_scores(oracle, gold):
    h = ndtb.Handle(0)
    h.set_params(resolution=2.0, regularization_scale_factor=10.0)
    h.set_target(gold["target"]); h.set_source(gold["source"])
    h.set_regularization_pose(gold["reg_T"])
    s, g, H, _ = h.compute_derivatives(np.array([0.3, 0.05, 0.01, 0.005, 0.003, 0.017]))
    assert abs(s - gold["reg_score"][0]) <= TOL * abs(s) and rel_inf(g, gold["reg_g"]) <= TOL and rel_inf(H, gold["reg_H"]) <= TOL
    h.set_regularization_pose(None)
    o = O.OracleNdt(resolution=2.0); o.set_target(gold["target"]); o.set_source(gold["source"])
    tp, nv = h.calculate_scores(gold["align_DIRECT7_T"])
    tpo, nvo = o.calculate_scores(gold["align_DIRECT7_T"])
    assert abs(tp - tpo) <= 1e-12 * abs(tpo) and abs(nv - nvo) <= 1e-12 * abs(nvo)


@pytest.mark.parametrize("search", ["DIRECT7", "DIRECT1"])
def test_align_vs_golden(gold, search):
    h = ndtb.Handle(0)
    h.set_params(resolution=2.0, search_method=getattr(ndtb, search), max_iterations=35, trans_epsilon=0.01, step_size=0.1)
    h.set_target(gold["target"]); h.set_source(gold["source"])
    r = h.align()
    assert [r["iterations"], int(r["converged"])] == list(gold[f"align_{search}_iters"])
    dt, da = pose_delta(r["T"], gold[f"align_{search}_T"])
    assert dt <= 1e-4 and da <= 1e-5
    tr, tg = h.trace(), gold[f"align_{search}_trace"]
    assert tr.shape == tg.shape
    for a, b in zip(tr, tg):
        assert np.abs(a[:6] - b[:6]).max() <= 1e-9                       # pose vector of the evaluation
        assert abs(a[6] - b[6]) <= TOL * abs(b[6]) and rel_inf(a[7:13], b[7:13]) <= TOL and rel_inf(a[13:49], b[13:49]) <= TOL
    assert np.abs(r["transformation_array"] - gold[f"align_{search}_tarray"]).max() <= 1e-6
    assert rel_inf(r["hessian"], gold[f"align_{search}_hessian"]) <= TOL


@pytest.mark.parametrize("step", [0.1, 0.5, 1.0])
def test_align_with_line_search(oracle, gold, step):
    """NdtParams::use_line_search: More-Thuente step length (ndt_omp_impl.hpp:995-1071) with gradient-only
    evaluations inside the search and the all-FP64 computeHessian pass (:701-800) after a search that iterated."""
    guess = np.eye(4, dtype=np.float32); guess[0, 3] = 2.5; guess[1, 3] = 1.5     # far enough that every step size makes the search iterate
    h = ndtb.Handle(0)
    h.set_params(resolution=2.0, max_iterations=35, trans_epsilon=0.01, step_size=step, use_line_search=1)
    h.set_target(gold["target"]); h.set_source(gold["source"])
    r = h.align(guess)
    o = O.OracleNdt(resolution=2.0, max_iterations=35, trans_epsilon=0.01, step_size=step, use_line_search=True, num_threads=8)
    o.set_target(gold["target"]); o.set_source(gold["source"])
    ro = o.align(guess)
    assert (r["iterations"], r["converged"]) == (ro["iterations"], ro["converged"])
    tr, tg = h.trace(), ro["trace"]
    assert tr.shape == tg.shape and len(tr) > r["iterations"] + 1        # the search iterated: extra FP64-Hessian rows
    for a, b in zip(tr, tg):
        assert np.abs(a[:6] - b[:6]).max() <= 1e-9
        assert abs(a[6] - b[6]) <= TOL * abs(b[6]) and rel_inf(a[7:13], b[7:13]) <= TOL and rel_inf(a[13:49], b[13:49]) <= TOL
    dt, da = pose_delta(r["T"], ro["T"])
    assert dt <= 1e-4 and da <= 1e-5
    assert rel_inf(r["hessian"], ro["hessian"]) <= TOL


def test_reference_gtest_case_pclomp(oracle, two_plane):
    """Reads like TEST(ConvergenceComparison, PclOmp), test_svn_ndt.cpp:138-199."""
    ndt = registration.NormalDistributionsTransform()
    ndt.setResolution(1.0)
    ndt.setNeighborhoodSearchMethod(registration.DIRECT7)
    ndt.setMaximumIterations(50)
    ndt.setTransformationEpsilon(1e-4)
    ndt.setStepSize(0.1)
    ndt.setNumThreads(20)
    ndt.setInputTarget(two_plane["target"])
    ndt.setInputSource(two_plane["source"])
    ndt.computeTransformation(two_plane["guess"].astype(np.float32))
    assert ndt.hasConverged()
    assert ndt.getFinalNumIteration() < 50
    err = np.ascontiguousarray(np.linalg.inv(ndt.getFinalTransformation().astype(float)) @ two_plane["gt"])
    xi = np.zeros(6); oracle.lib().orc_pose_logmap(O._d(err), O._d(xi))
    assert np.linalg.norm(xi[3:]) < 0.05 and np.linalg.norm(xi[:3]) < 0.035
    # and the same trajectory as the oracle
    o = O.OracleNdt(resolution=1.0, max_iterations=50, trans_epsilon=1e-4, step_size=0.1, num_threads=8)
    o.set_target(two_plane["target"]); o.set_source(two_plane["source"])
    ro = o.align(two_plane["guess"].astype(np.float32))
    assert ndt.getFinalNumIteration() == ro["iterations"]
    dt, da = pose_delta(ndt.getFinalTransformation(), ro["T"])
    assert dt <= 1e-4 and da <= 1e-5
    res = ndt.getResult()
    assert abs(res.transform_probability - ro["trans_probability"]) <= 1e-9 * abs(ro["trans_probability"])
    assert len(res.transformation_array) == len(ro["transformation_array"])


def test_setters_rebuild_only_on_change(street):
    ndt = registration.NormalDistributionsTransform()
    ndt.setInputTarget(street["target"])
    v1 = ndt.getTargetCells()
    ndt.setResolution(1.0)          # unchanged -> no rebuild, same map
    assert np.array_equal(ndt.getTargetCells()["idx"], v1["idx"])
    ndt.setResolution(2.0)          # ndt_omp.h:135-142: rebuild with the new leaf size
    v2 = ndt.getTargetCells()
    assert len(v2["idx"]) < len(v1["idx"])
    ndt.setNeighborhoodSearchMethod(registration.KDTREE)   # radius search: generic kernel (off the hot path)
    ndt.setNeighborhoodSearchMethod(registration.DIRECT7)
    empty = registration.NormalDistributionsTransform()
    with pytest.raises(ndtb.NdtbError):
        empty.align()               # NDTB_ERR_NOT_READY: no target / source


@pytest.mark.parametrize("search", ["DIRECT26", "KDTREE"])
def test_direct26_and_kdtree_match_oracle(oracle, street, gold, search):
    """DIRECT26 (the 26 cells of pcl::getAllNeighborCellIndices, centre excluded; voxel_grid_covariance_omp_impl.hpp:366-411) and
    KDTREE (radius search over float(leaf.mean_), voxel_grid_covariance_omp.h:458-485) run on the device through the generic
    two-phase kernel: per-evaluation parity and the whole Newton trajectory."""
    for tgt, src, res in ((gold["target"], gold["source"], 2.0), (street["target"], street["source"], 1.0)):
        h = ndtb.Handle(0)
        h.set_params(resolution=res, search_method=getattr(ndtb, search), max_iterations=35, trans_epsilon=0.01, step_size=0.1)
        assert h.set_target(tgt) == 0 and h.set_source(src) == 0
        o = O.OracleNdt(resolution=res, search=search, max_iterations=35, trans_epsilon=0.01, step_size=0.1, num_threads=4)
        o.set_target(tgt); o.set_source(src)
        for p in (np.zeros(6), np.array([0.3, -0.1, 0.02, 0.004, -0.003, 0.02])):
            s, g, H, nv = h.compute_derivatives(p)
            so, go, Ho, nvo = o.compute_derivatives(p)
            assert abs(s - so) <= TOL * abs(so) and rel_inf(g, go) <= TOL and rel_inf(H, Ho) <= TOL and abs(nv - nvo) <= TOL * abs(nvo)
            hits26 = h.last_hits()[0]
            h7 = ndtb.Handle(0); h7.set_params(resolution=res, search_method=ndtb.DIRECT7); h7.set_target(tgt); h7.set_source(src)
            h7.compute_derivatives(p)
            assert hits26 > h7.last_hits()[0] or search == "KDTREE"   # 26 cells reach more leaves than the 7 of DIRECT7
            s2, g2, H2, _ = h.compute_derivatives(p, compute_hessian=False)
            assert s2 == s and np.array_equal(g2, g) and np.all(H2 == 0)
            s3, g3, H3, _ = h.compute_derivatives(p)
            assert s3 == s and np.array_equal(g3, g) and np.array_equal(H3, H)   # run-to-run bit stable
        rg, ro = h.align(), o.align()
        assert rg["iterations"] == ro["iterations"] and rg["converged"] == ro["converged"]
        tg, to = h.trace(), ro["trace"]
        assert len(tg) == len(to)
        for a, b in zip(tg, to):
            assert abs(a[6] - b[6]) <= TOL * abs(b[6]) and np.abs(a[7:49] - b[7:49]).max() <= TOL * np.abs(b[13:49]).max()
        dt, da = pose_delta(rg["T"], ro["T"])
        assert dt <= 1e-4 and da <= 1e-5
        tp, nv = h.calculate_scores(ro["T"])           # the score-only passes (ndt_omp_impl.hpp:1169-1281) with the same neighbourhood
        tpo, nvo = o.calculate_scores(ro["T"])
        assert abs(tp - tpo) <= TOL * abs(tpo) and abs(nv - nvo) <= TOL * abs(nvo)
    # what stays outside: these two with line search or for the svn flavour
    h = ndtb.Handle(0)
    p = ndtb.Params(); h.L.ndtb_get_params(h.h, p)
    p.search_method = getattr(ndtb, search); p.use_line_search = 1
    assert h.L.ndtb_set_params(h.h, p) == ndtb.NDTB_ERR_INVALID
    p.use_line_search = 0; p.grid_flavour = ndtb.FLAVOUR_SVN      # svn_ndt::NeighborSearchMethod has KDTREE but no DIRECT26 (svn_ndt.h:31-36)
    assert h.L.ndtb_set_params(h.h, p) == (ndtb.NDTB_OK if search == "KDTREE" else ndtb.NDTB_ERR_INVALID)


@pytest.mark.parametrize("search", ["DIRECT7", "DIRECT1", "KDTREE"])
def test_calculate_score_and_maps(oracle, street, search):
    """calculateScore (ndt_omp_impl.hpp:1083-1167): score, getScoreMap() and getEmptyVoxels() against the oracle."""
    h = ndtb.Handle(0)
    h.set_params(resolution=1.0, search_method=getattr(ndtb, search))
    assert h.set_target(street["target"]) == 0 and h.set_source(street["source"]) == 0
    o = O.OracleNdt(resolution=1.0, search=search)
    o.set_target(street["target"]); o.set_source(street["source"])
    T = street["rel"].astype(np.float32)
    sg, mg, eg = h.calculate_score(T)
    so, mo, eo = o.calculate_score(T)
    assert abs(sg - so) <= 1e-12 * abs(so)
    assert sg == pytest.approx(h.calculate_scores(T)[0], rel=1e-12)       # same number as calculateTransformationProbability
    assert list(mg.keys()) == list(mo.keys()) and len(mg) > 100
    a, b = np.array(list(mg.values())), np.array(list(mo.values()))
    assert np.abs(a - b).max() <= 1e-12 * np.abs(b).max()
    assert np.array_equal(eg, eo) and len(eg) > 0


def test_ndt_export_files_byte_for_byte(oracle, street, tmp_path):
    """extractNdtData + writeNdtDataToFiles (include/pipeline.hpp:163-274): ndt_ellipsoids.txt and ndt_voxels.txt from the device map,
    written by the library (iostream, std::fixed, 8 decimals like the reference) == the oracle's writer (printf), byte for byte."""
    h = ndtb.Handle(0)
    h.set_params(resolution=1.0)
    assert h.set_target(street["target"]) == 0
    o = O.OracleNdt(resolution=1.0); o.set_target(street["target"])
    pts = np.zeros((5, 8), np.float32); pts[:, :3] = street["target"][:5, :3]; pts[:, 4] = np.arange(5) + 0.5
    h.write_ndt_files(tmp_path / "ge.txt", tmp_path / "gv.txt", tmp_path / "gp.txt", pts)
    n = o.write_files(tmp_path / "oe.txt", tmp_path / "ov.txt")
    assert n == int((h.voxels()["n"] >= 6).sum()) and n > 100
    assert (tmp_path / "ge.txt").read_bytes() == (tmp_path / "oe.txt").read_bytes()
    assert (tmp_path / "gv.txt").read_bytes() == (tmp_path / "ov.txt").read_bytes()
    first = (tmp_path / "ge.txt").read_text().splitlines()[0].split()
    assert len(first) == 16 and int(first[15]) >= 6
    lines = (tmp_path / "gp.txt").read_text().splitlines()
    assert len(lines) == 5 and lines[2].split()[3] == "2.50000000"


def test_staged_uploads_equal_the_direct_ones(street):
    """ndtb_stage_* / ndtb_commit_*: the pipelined setInputTarget / setInputSource give the same map and the same align, in
    FIFO order, and refuse a third staged cloud / a commit with nothing staged."""
    tgt, src = street["target"], street["source"]
    tgt2 = np.ascontiguousarray(tgt[: tgt.shape[0] // 2])
    direct = ndtb.Handle(0); direct.set_params(resolution=1.0)
    assert direct.set_target(tgt) == 0 and direct.set_source(src) == 0
    want_map, want = direct.voxels(), direct.align()
    h = ndtb.Handle(0); h.set_params(resolution=1.0)
    assert h.commit_target() == ndtb.NDTB_ERR_NOT_READY and h.commit_source() == ndtb.NDTB_ERR_NOT_READY
    assert h.stage_target(tgt2) == 0 and h.stage_target(tgt) == 0        # two in flight
    assert h.stage_target(tgt) == ndtb.NDTB_ERR_NOT_READY                # a third is refused
    assert h.stage_source(src[:, :3]) == 0                               # 12-byte stride
    assert h.staged_count() == (2, 1)
    assert h.commit_target() == 0                                        # oldest first: the half cloud
    direct2 = ndtb.Handle(0); direct2.set_params(resolution=1.0)
    assert direct2.set_target(tgt2) == 0
    assert_grid_equal(h.voxels(), direct2.voxels())
    assert h.commit_target() == 0 and h.commit_source() == 0
    assert h.staged_count() == (0, 0)
    assert_grid_equal(h.voxels(), want_map)
    got = h.align()
    assert got["iterations"] == want["iterations"] and np.array_equal(got["T"], want["T"])
    # keep going: the buffers rotate (the former target buffer is the next staging slot)
    for k in range(4):
        assert h.stage_target(tgt if k % 2 else tgt2) == 0 and h.stage_source(src) == 0
        assert h.commit_target() == 0 and h.commit_source() == 0
        assert_grid_equal(h.voxels(), want_map if k % 2 else direct2.voxels())
    assert np.array_equal(h.align()["T"], want["T"])
    # after a caller-owned device buffer the commit must not recycle that buffer
    import torch
    d = torch.from_numpy(np.ascontiguousarray(tgt)).cuda(); torch.cuda.synchronize()
    assert h.set_target_device(d.data_ptr(), tgt.shape[0]) == 0
    assert h.stage_target(tgt2) == 0 and h.commit_target() == 0 and h.stage_target(tgt) == 0 and h.commit_target() == 0
    assert_grid_equal(h.voxels(), want_map)
    assert np.array_equal(d.cpu().numpy(), tgt)


@pytest.mark.parametrize("scale", [1e8, 1e12, 1e20])
def test_flagged_leaves_stay_reachable_by_the_radius_search(oracle, scale):
    """Leaves the reference flags nr_points_ = -1 AFTER they entered its centroid cloud (negative eigenvalue / infinite inverse,
    voxel_grid_covariance_omp_impl.hpp:335-338, 355-359) are still returned by its kd-tree, with the icov_ the leaf holds.
    Clusters of points one float ulp apart along x, at coordinates where x^2 swamps the FP64 second moment, produce such leaves:
    the map must flag the same leaves, KDTREE must see them (row code -2 - v in the table), DIRECT7 must not."""
    rng = np.random.default_rng(3)
    leaf = scale / 8
    centers = (rng.uniform(-4, 4, (60, 3)) * leaf).astype(np.float32)
    pts = []
    for c in centers:
        k = rng.integers(6, 12)
        p = np.tile(c, (k, 1))
        p[:, 0] = c[0] + rng.integers(0, 3, k).astype(np.float32) * np.spacing(np.abs(c[0]))
        pts.append(p)
    cloud = np.zeros((sum(len(p) for p in pts), 4), np.float32)
    cloud[:, :3] = np.concatenate(pts)
    src = cloud[::3].copy()
    g = O.OracleGrid(0, float(leaf), 6); g.build(cloud); e = g.export()
    assert (e["n"] == -1).sum() >= 5 and (e["n"] >= 6).sum() >= 5
    scores = {}
    for search in ("KDTREE", "DIRECT7"):
        h = ndtb.Handle(0)
        h.set_params(resolution=float(leaf), search_method=getattr(ndtb, search))
        assert h.set_target(cloud) == 0 and h.set_source(src) == 0
        assert_grid_equal(h.voxels(), e)
        o = O.OracleNdt(resolution=float(leaf), search=search, num_threads=2)
        assert o.set_target(cloud) == 0
        o.set_source(src)
        s, gr, H, nv = h.compute_derivatives(np.zeros(6))
        so, go, Ho, nvo = o.compute_derivatives(np.zeros(6))
        assert abs(s - so) <= TOL * abs(so) and rel_inf(gr, go) <= TOL and rel_inf(H, Ho) <= TOL and abs(nv - nvo) <= TOL * abs(nvo)
        scores[search] = s
    assert scores["KDTREE"] > scores["DIRECT7"] * 1.1                     # the flagged leaves contribute to the radius search only
