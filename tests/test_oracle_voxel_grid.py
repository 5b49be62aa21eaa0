"""Oracle voxel grids against an independent numpy restatement + the reference's edge cases (CPU)."""
import numpy as np
import pytest

import oracle_lib as O


def numpy_grid(cloud, leaf, flavour, min_pts, eig_mult=0.01):
    """Independent restatement (numpy, float32 index maths) of
    voxel_grid_covariance_omp_impl.hpp:80-113,211-361 / voxel_grid_covariance_impl.hpp:103-343."""
    xyz = cloud[:, :3].astype(np.float32)
    inv = np.float32(1.0) / np.float32(leaf)
    mn, mx = xyz.min(0), xyz.max(0)
    min_b = np.floor(mn * inv).astype(np.int64)
    max_b = np.floor(mx * inv).astype(np.int64)
    div_b = max_b - min_b + 1
    ijk = (np.floor(xyz * inv) - min_b.astype(np.float32)).astype(np.int64)
    idx = ijk[:, 0] + ijk[:, 1] * div_b[0] + ijk[:, 2] * div_b[0] * div_b[1]
    out = {}
    for key in np.unique(idx):
        pts = xyz[idx == key].astype(np.float64)
        n = len(pts)
        S = pts.sum(0); SS = np.eye(3) + pts.T @ pts      # Leaf::cov_ starts at identity (voxel_grid_covariance_omp.h:113, svn .h:109)
        mean = S / n
        if n < min_pts:
            if flavour == 0:
                out[int(key)] = dict(n=n, mean=mean)
            continue
        if flavour == 0:
            cov = (SS - 2 * np.outer(S, mean)) / n + np.outer(mean, mean)
            cov *= (n - 1.0) / n
        else:
            cov = SS / n - np.outer(mean, mean)
            cov *= n / (n - 1.0)
        w, V = np.linalg.eigh((cov + cov.T) / 2)
        if flavour == 0:
            if w[0] < 0 or w[1] < 0 or w[2] <= 0:
                out[int(key)] = dict(n=-1, mean=mean); continue
            lim = eig_mult * w[2]
            if w[0] < lim:
                w[0] = lim
                if w[1] < lim:
                    w[1] = lim
                cov = V @ np.diag(w) @ np.linalg.inv(V)
        else:
            if w[0] < 0 or w[1] < 0 or w[2] < 1e-12:
                continue
            lim = max(1e-12, eig_mult * w[2])
            if w[0] < lim or w[1] < lim:
                w = np.maximum(w, [lim, lim, -np.inf])
                cov = V @ np.diag(w) @ V.T
        out[int(key)] = dict(n=n, mean=mean, cov=cov, icov=np.linalg.inv(cov))
    return out, min_b, max_b, div_b


@pytest.mark.parametrize("flavour,min_pts", [(0, 6), (1, 3), (1, 6)])
def test_grid_matches_numpy(oracle, street, flavour, min_pts):
    cloud = street["target"]
    ref, min_b, max_b, div_b = numpy_grid(cloud, 1.0, flavour, min_pts)
    g = O.OracleGrid(flavour, 1.0, min_pts)
    assert g.build(cloud) == 0
    e = g.export()
    assert np.array_equal(e["min_b"], min_b) and np.array_equal(e["max_b"], max_b) and np.array_equal(e["div_b"], div_b)
    assert list(e["idx"]) == sorted(ref.keys())          # occupancy + linear indices: exact
    for k, key in enumerate(e["idx"]):
        r = ref[int(key)]
        assert e["n"][k] == r["n"]
        np.testing.assert_allclose(e["mean"][k], r["mean"], rtol=1e-13)
        if "icov" in r:
            np.testing.assert_allclose(e["cov"][k].reshape(3, 3), r["cov"], rtol=1e-7, atol=1e-12)
            np.testing.assert_allclose(e["icov"][k].reshape(3, 3), r["icov"], rtol=1e-6, atol=1e-6)
            ev = e["evals"][k]
            assert ev[0] >= 0.01 * ev[2] * (1 - 1e-12)     # eigenvalue inflation (eq 6.11)


def test_neighbourhood_order_and_bounds(oracle, street):
    cloud = street["target"]
    g = O.OracleGrid(0, 1.0, 6)
    g.build(cloud)
    e = g.export()
    valid = e["n"] >= 6
    key_of_mean = {tuple(m): int(i) for m, i in zip(map(tuple, e["mean"][valid]), e["idx"][valid])}
    div = e["div_b"]; min_b = e["min_b"]
    rng = np.random.default_rng(0)
    for p in cloud[rng.integers(0, len(cloud), 200), :3]:
        nb = g.neighbours(p, 7)
        ijk = np.floor(p / np.float32(1.0)).astype(int)
        expect = []
        for d in [(0, 0, 0), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]:   # reference probe order
            q = ijk + d - min_b
            if (q < 0).any() or (q >= div).any():
                continue
            lin = q[0] + q[1] * div[0] + q[2] * div[0] * div[1]
            if lin in set(e["idx"][valid]):
                expect.append(int(lin))
        got = [key_of_mean[tuple(m)] for m in nb]
        assert got == expect
        assert len(g.neighbours(p, 1)) == (1 if expect and expect[0] == (ijk - min_b) @ [1, div[0], div[0] * div[1]] else 0)
    # far outside the bounding box: nothing, for both flavours
    assert len(g.neighbours((1e4, 0, 0), 7)) == 0
    gs = O.OracleGrid(1, 1.0, 3); gs.build(cloud)
    assert len(gs.neighbours((1e4, 0, 0), 7)) == 0


def test_edge_cases(oracle):
    g = O.OracleGrid(0, 1.0, 6)
    assert g.build(np.zeros((0, 4), np.float32)) == -1                      # empty input
    big = np.array([[0, 0, 0, 0], [4000, 4000, 4000, 0]], np.float32)
    assert O.OracleGrid(0, 1.0, 6).build(big) == -2                          # dx*dy*dz > INT32_MAX (:83-94)
    assert O.OracleGrid(1, 1.0, 6).build(big) == -2
    # non-finite points: skipped when !is_dense (pclomp) / always (svn)
    rng = np.random.default_rng(0)
    c = rng.uniform(0, 2, (200, 4)).astype(np.float32)
    c[5, 0] = np.nan; c[17, 2] = np.inf
    g = O.OracleGrid(0, 1.0, 6); assert g.build(c, is_dense=False) == 0
    assert g.export()["n"].sum() == 198
    g = O.OracleGrid(1, 1.0, 3); assert g.build(c, is_dense=False) == 0
    assert g.export()["n"].sum() <= 198
    # a cloud flagged dense that still holds inf: pcl::getMinMax3D takes every point -> overflow guard trips
    assert O.OracleGrid(1, 1.0, 3).build(c, is_dense=True) == -2
    # a voxel with identical points: the scatter is zero, but the reference's second moment starts at identity
    # (Leaf::cov_ = I, voxel_grid_covariance_omp.h:113 / svn voxel_grid_covariance.h:109), so the covariance is
    # I/n scaled by Bessel's factor -- the leaf is KEPT, the lambda2 <= 0 rejection never fires on real clouds
    same = np.tile(np.array([[0.5, 0.5, 0.5, 0]], np.float32), (10, 1))
    g = O.OracleGrid(0, 1.0, 6); g.build(same); e = g.export(); assert list(e["n"]) == [10]
    np.testing.assert_allclose(e["cov"][0].reshape(3, 3), np.eye(3) * (1 / 10) * (9 / 10), atol=1e-12)
    g = O.OracleGrid(1, 1.0, 6); g.build(same); e = g.export(); assert list(e["n"]) == [10]
    np.testing.assert_allclose(e["icov"][0].reshape(3, 3), np.eye(3) * 9.0, atol=1e-9)
    # exactly min_points_per_voxel points is enough, one fewer is not
    pts = rng.uniform(0, 1, (6, 4)).astype(np.float32)
    g = O.OracleGrid(0, 1.0, 6); g.build(pts); assert len(g.neighbours((0.5, 0.5, 0.5), 1)) == 1
    g = O.OracleGrid(0, 1.0, 7); g.build(pts); assert len(g.neighbours((0.5, 0.5, 0.5), 1)) == 0
