"""Oracle building blocks against numpy (CPU).  These pin the third-party arithmetic the reference
takes from Eigen / GTSAM: 3x3 eigen / inverse, 6x6 SVD solve, LDLT, SE(3) maps."""
import numpy as np

import oracle_lib as O
from slam_sam_b200 import synth


def test_eig3_and_inverse(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(0)
    for k in range(200):
        a = rng.normal(size=(3, 3)) * 10 ** rng.uniform(-3, 2)
        A = a @ a.T
        ev = np.zeros(3); V = np.zeros((3, 3)); inv = np.zeros((3, 3))
        L.orc_eig3_sym(O._d(A), O._d(ev), O._d(V))
        w = np.linalg.eigvalsh(A)
        assert np.allclose(ev, w, rtol=1e-11, atol=1e-13 * abs(w).max())
        assert np.all(np.diff(ev) >= 0)
        assert np.allclose(V @ np.diag(ev) @ V.T, A, atol=1e-12 * abs(A).max())
        if np.linalg.cond(A) < 1e8:
            L.orc_mat3_inverse(O._d(A), O._d(inv))
            assert np.allclose(inv @ A, np.eye(3), atol=1e-7)
    # planar (rank-2) covariance, the usual LiDAR voxel: smallest eigenvalue ~ 0
    pts = rng.normal(size=(50, 3)) * [1.0, 1.0, 0.0]
    A = np.cov(pts.T)
    ev = np.zeros(3); V = np.zeros((3, 3))
    L.orc_eig3_sym(O._d(np.ascontiguousarray(A)), O._d(ev), O._d(V))
    assert abs(ev[0]) < 1e-15 and ev[2] > 0.1


def test_svd6_solve(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(1)
    for k in range(200):
        A = rng.normal(size=(6, 6)); b = rng.normal(size=6); x = np.zeros(6); S = np.zeros(6)
        L.orc_svd6_solve(O._d(A), O._d(b), O._d(x), O._d(S))
        assert np.allclose(S, np.linalg.svd(A, compute_uv=False), rtol=1e-11)
        assert np.allclose(x, np.linalg.solve(A, b), rtol=1e-7, atol=1e-9)
    # rank-deficient: Eigen's JacobiSVD::solve gives the minimum-norm (pseudo-inverse) solution
    B = rng.normal(size=(6, 3)); A = np.ascontiguousarray(B @ B.T); b = rng.normal(size=6); x = np.zeros(6); S = np.zeros(6)
    L.orc_svd6_solve(O._d(A), O._d(b), O._d(x), O._d(S))
    assert np.allclose(x, np.linalg.pinv(A) @ b, atol=1e-9)
    # zero matrix -> zero step (computeTransformation's delta_p_norm == 0 exit)
    A = np.zeros((6, 6)); L.orc_svd6_solve(O._d(A), O._d(b), O._d(x), O._d(S))
    assert np.all(x == 0)


def test_ldlt6_and_eig6(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(2)
    for k in range(100):
        a = rng.normal(size=(6, 6)); A = a @ a.T + 1e-3 * np.eye(6)
        if k % 2:
            A = -A
        A = np.ascontiguousarray(A); b = rng.normal(size=6); x = np.zeros(6)
        assert L.orc_ldlt6_solve(O._d(A), O._d(b), O._d(x)) == 1
        assert np.allclose(x, np.linalg.solve(A, b), rtol=1e-8, atol=1e-9)
    a = rng.normal(size=(6, 6)); A = np.ascontiguousarray(a @ a.T); ev = np.zeros(6); V = np.zeros((6, 6))
    L.orc_eig6_sym(O._d(A), O._d(ev), O._d(V))
    assert np.allclose(ev, np.linalg.eigvalsh(A)) and np.allclose(V @ np.diag(ev) @ V.T, A)


def test_se3_maps(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(3)
    for k in range(100):
        xi = rng.normal(size=6) * (0.5 if k % 4 else 1e-9)
        M = np.zeros((4, 4)); L.orc_pose_expmap(O._d(xi), O._d(M))
        assert np.allclose(M, synth.se3_exp(xi), atol=1e-14)
        assert np.allclose(M[:3, :3] @ M[:3, :3].T, np.eye(3), atol=1e-13)
        xi2 = np.zeros(6); L.orc_pose_logmap(O._d(M), O._d(xi2))
        assert np.allclose(xi, xi2, atol=1e-10)
        rpy = np.zeros(3); L.orc_rot_rpy(O._d(M), O._d(rpy))
        R = synth.rot_z(rpy[2]) @ synth.rot_y(rpy[1]) @ synth.rot_x(rpy[0])  # GTSAM: R = Rz Ry Rx
        assert np.allclose(R, M[:3, :3], atol=1e-12)


def test_pose_matrix_and_euler(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(4)
    for k in range(100):
        p = np.concatenate([rng.normal(size=3), rng.uniform(-1, 1, 3)])
        T = np.zeros(16, np.float32); L.orc_pose_to_matrix_f(O._d(p), O._f(T))
        assert np.allclose(T.reshape(4, 4), synth.pose_matrix(p[:3], p[3:]), atol=1e-6)
        e = np.zeros(3, np.float32); L.orc_euler_angles_012_f(O._f(T), O._f(e))
        # whatever branch Eigen 3.4 takes, the angles must reproduce the rotation
        assert np.allclose(synth.pose_matrix(p[:3], e.astype(float)), T.reshape(4, 4), atol=2e-5)
        assert 0.0 <= e[0] <= np.pi + 1e-6  # Eigen 3.4: first angle folded into [0, pi]
    T = np.eye(4, dtype=np.float32).reshape(16); e = np.ones(3, np.float32)
    L.orc_euler_angles_012_f(O._f(T), O._f(e))
    assert np.all(e == 0)
