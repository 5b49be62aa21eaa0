"""ctypes binding of the CPU oracle (oracle/libndt_oracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module; the product package (slam_sam_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_LIB = None

c_f = C.POINTER(C.c_float)
c_d = C.POINTER(C.c_double)
c_i = C.POINTER(C.c_int)
c_ll = C.POINTER(C.c_longlong)


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(ORACLE_DIR, "libndt_oracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_grid_create.restype = C.c_void_p
        L.orc_grid_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_double]
        L.orc_grid_destroy.argtypes = [C.c_void_p]
        L.orc_grid_build.argtypes = [C.c_void_p, c_f, C.c_size_t, C.c_size_t, C.c_int]
        L.orc_grid_meta.argtypes = [C.c_void_p, c_i, c_i, c_i]
        L.orc_grid_count.restype = C.c_size_t
        L.orc_grid_count.argtypes = [C.c_void_p]
        L.orc_grid_export.argtypes = [C.c_void_p, c_ll, c_i, c_d, c_d, c_d, c_d, c_d]
        L.orc_grid_neighbours.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, c_d]
        L.orc_ndt_create.restype = C.c_void_p
        L.orc_ndt_destroy.argtypes = [C.c_void_p]
        L.orc_ndt_grid.restype = C.c_void_p
        L.orc_ndt_grid.argtypes = [C.c_void_p]
        L.orc_ndt_set_params.argtypes = [C.c_void_p, C.c_float, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int,
                                         C.c_float, C.c_int, C.c_double, C.c_int, C.c_double]
        L.orc_ndt_gauss.argtypes = [C.c_void_p, c_d]
        L.orc_ndt_set_target.argtypes = [C.c_void_p, c_f, C.c_size_t, C.c_size_t, C.c_int]
        L.orc_ndt_set_source.argtypes = [C.c_void_p, c_f, C.c_size_t, C.c_size_t]
        L.orc_ndt_set_regularization_pose.argtypes = [C.c_void_p, c_f]
        L.orc_ndt_compute_derivatives.restype = C.c_double
        L.orc_ndt_compute_derivatives.argtypes = [C.c_void_p, c_d, C.c_int, c_d, c_d, c_d]
        L.orc_ndt_compute_derivatives_T.restype = C.c_double
        L.orc_ndt_compute_derivatives_T.argtypes = [C.c_void_p, c_f, c_d, C.c_int, c_d, c_d, c_d]
        L.orc_ndt_calculate_scores.argtypes = [C.c_void_p, c_f, c_d, c_d]
        L.orc_ndt_calculate_score.restype = C.c_double
        L.orc_ndt_calculate_score.argtypes = [C.c_void_p, c_f, C.c_size_t, c_ll, c_d, C.POINTER(C.c_size_t), C.c_size_t, c_i, C.POINTER(C.c_size_t)]
        L.orc_ndt_write_files.restype = C.c_long
        L.orc_ndt_write_files.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.orc_ndt_align.argtypes = [C.c_void_p, c_f, c_f, c_d, c_d, c_d, c_i, c_i, c_d, C.c_int, c_i, c_f, C.c_int, c_i]
        L.orc_assemble_window.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_size_t, c_d, c_f]
        L.orc_assemble_window.restype = None
        L.orc_voxelgrid_filter.argtypes = [c_f, C.c_size_t, C.c_size_t, C.c_float, c_f]
        L.orc_mg_create.restype = C.c_void_p
        L.orc_mg_destroy.argtypes = [C.c_void_p]
        L.orc_mg_set_params.argtypes = [C.c_void_p, C.c_float, C.c_double, C.c_double, C.c_int, C.c_int, C.c_float, C.c_int, C.c_double,
                                        C.c_int, C.c_double]
        L.orc_mg_add_target.argtypes = [C.c_void_p, c_f, C.c_size_t, C.c_size_t, C.c_int, C.c_char_p]
        L.orc_mg_remove_target.argtypes = [C.c_void_p, C.c_char_p]
        L.orc_mg_create_kdtree.argtypes = [C.c_void_p]
        L.orc_mg_leaf_count.restype = C.c_long
        L.orc_mg_leaf_count.argtypes = [C.c_void_p]
        L.orc_mg_export.restype = C.c_long
        L.orc_mg_export.argtypes = [C.c_void_p, C.c_long, c_i, c_d, c_d, c_f]
        L.orc_mg_radius_search.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, c_d, c_f]
        L.orc_mg_set_source.argtypes = [C.c_void_p, c_f, C.c_size_t, C.c_size_t]
        L.orc_mg_set_regularization_pose.argtypes = [C.c_void_p, c_f]
        L.orc_mg_compute_derivatives.restype = C.c_double
        L.orc_mg_compute_derivatives.argtypes = [C.c_void_p, c_d, C.c_int, c_d, c_d, c_d]
        L.orc_mg_calculate_scores.argtypes = [C.c_void_p, c_f, c_d, c_d]
        L.orc_mg_align.argtypes = [C.c_void_p, c_f, c_f, c_d, c_d, c_d, c_i, c_i, c_d, C.c_int, c_i, c_f, C.c_int, c_i]
        L.orc_voxelgrid_filter.restype = C.c_long
        L.orc_svn_create.restype = C.c_void_p
        L.orc_svn_destroy.argtypes = [C.c_void_p]
        L.orc_svn_grid.restype = C.c_void_p
        L.orc_svn_grid.argtypes = [C.c_void_p]
        L.orc_svn_set_params.argtypes = [C.c_void_p, C.c_float, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_double]
        L.orc_svn_gauss.argtypes = [C.c_void_p, c_d]
        L.orc_svn_set_target.argtypes = [C.c_void_p, c_f, C.c_size_t, C.c_size_t]
        L.orc_svn_set_source.argtypes = [C.c_void_p, c_f, C.c_size_t, C.c_size_t]
        L.orc_svn_sample_particles.argtypes = [c_d, C.c_int, C.c_ulonglong, c_d]
        L.orc_svn_particle_derivatives.restype = C.c_double
        L.orc_svn_particle_derivatives.argtypes = [C.c_void_p, c_d, c_d, c_d]
        L.orc_svn_step.argtypes = [C.c_void_p, C.c_int, c_d, c_d, c_d, c_d, c_d]
        L.orc_svn_stage2.argtypes = [C.c_void_p, C.c_int, c_d, c_d, c_d, c_d]
        L.orc_svn_align.restype = C.c_int
        L.orc_svn_align.argtypes = [C.c_void_p, c_d, C.c_int, c_d, c_d, c_d, c_i]
        for name in ("orc_eig3_sym", "orc_mat3_inverse", "orc_svd6_solve", "orc_eig6_sym", "orc_pose_expmap",
                     "orc_pose_logmap", "orc_rot_rpy"):
            getattr(L, name).restype = None
        L.orc_ldlt6_solve.restype = C.c_int
        c_u16 = C.POINTER(C.c_uint16)
        L.orc_lidar_create.restype = C.c_void_p
        L.orc_lidar_create.argtypes = [c_i, c_f, c_d, c_f, c_f]
        L.orc_lidar_destroy.argtypes = [C.c_void_p]
        L.orc_lidar_packet_bytes.restype = C.c_size_t
        L.orc_lidar_packet_bytes.argtypes = [C.c_void_p]
        L.orc_lidar_tables.argtypes = [C.c_void_p, c_f, c_f, c_f, c_f, c_f, c_f]
        L.orc_lidar_decode.restype = C.c_long
        L.orc_lidar_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t]
        L.orc_lidar_frame_info.argtypes = [C.c_void_p, C.c_long, c_d]
        L.orc_lidar_frame_get.argtypes = [C.c_void_p, C.c_long, c_f, c_f, c_f, c_u16, c_u16, c_d, c_f, c_u16, c_u16, c_u16]
        L.orc_lidar_clear_frames.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def _f(a):
    return a.ctypes.data_as(c_f)


def _d(a):
    return a.ctypes.data_as(c_d)


def _i(a):
    return a.ctypes.data_as(c_i)


def as_xyz(cloud):
    a = np.ascontiguousarray(cloud, dtype=np.float32)
    assert a.ndim == 2 and a.shape[1] >= 3
    return a


SEARCH = {"KDTREE": 0, "DIRECT26": 1, "DIRECT7": 2, "DIRECT1": 3}


def export_grid(gptr):
    L = lib()
    V = L.orc_grid_count(gptr)
    idx = np.zeros(V, np.int64); n = np.zeros(V, np.int32)
    mean = np.zeros((V, 3)); cov = np.zeros((V, 9)); icov = np.zeros((V, 9)); evecs = np.zeros((V, 9)); evals = np.zeros((V, 3))
    if V:
        L.orc_grid_export(gptr, idx.ctypes.data_as(c_ll), _i(n), _d(mean), _d(cov), _d(icov), _d(evecs), _d(evals))
    mn = np.zeros(3, np.int32); mx = np.zeros(3, np.int32); dv = np.zeros(3, np.int32)
    L.orc_grid_meta(gptr, _i(mn), _i(mx), _i(dv))
    return dict(idx=idx, n=n, mean=mean, cov=cov, icov=icov, evecs=evecs, evals=evals, min_b=mn, max_b=mx, div_b=dv)


class OracleGrid:
    def __init__(self, flavour=0, leaf=1.0, min_pts=6, eig_mult=0.01):
        self.h = lib().orc_grid_create(flavour, leaf, min_pts, eig_mult)

    def build(self, cloud, is_dense=True):
        a = as_xyz(cloud)
        return lib().orc_grid_build(self.h, _f(a), a.shape[0], a.shape[1], int(is_dense))

    def export(self):
        return export_grid(self.h)

    def neighbours(self, p, n_probe=7):
        out = np.zeros((7, 3))
        n = lib().orc_grid_neighbours(self.h, float(p[0]), float(p[1]), float(p[2]), n_probe, _d(out))
        return out[:n]

    def __del__(self):
        try:
            lib().orc_grid_destroy(self.h)
        except Exception:
            pass


def assemble_window(clouds, poses):
    """Sliding-window target: every keyframe cloud moved by its double 4x4 pose (pcl::transformPointCloud<.,double>),
    concatenated in order.  Returns packed (sum n, 4) float32 with w = 1."""
    arrs = [as_xyz(c) for c in clouds]
    stride = arrs[0].shape[1]
    assert all(a.shape[1] == stride for a in arrs)
    ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
    ns = (C.c_size_t * len(arrs))(*[a.shape[0] for a in arrs])
    P = np.ascontiguousarray(poses, np.float64).reshape(len(arrs), 16)
    out = np.zeros((sum(a.shape[0] for a in arrs), 4), np.float32)
    lib().orc_assemble_window(len(arrs), ptrs, ns, stride, _d(P), _f(out))
    return out


def voxelgrid_filter(cloud, leaf):
    """pcl::VoxelGrid down-sampling restatement: (m, 4) float32 centroids in ascending leaf index."""
    a = as_xyz(cloud)
    out = np.zeros((a.shape[0], 4), np.float32)
    m = lib().orc_voxelgrid_filter(_f(a), a.shape[0], a.shape[1], float(leaf), _f(out))
    if m < 0:
        raise OverflowError("leaf size too small")
    return out[:m].copy()


class OracleNdt:
    """pclomp::NormalDistributionsTransform restatement."""

    def __init__(self, resolution=1.0, step_size=0.1, trans_epsilon=0.1, max_iterations=35, search="DIRECT7",
                 num_threads=1, reg_scale=0.0, use_line_search=False, outlier_ratio=0.55, min_pts=6, eig_mult=0.01):
        self.h = lib().orc_ndt_create()
        lib().orc_ndt_set_params(self.h, resolution, step_size, trans_epsilon, max_iterations, SEARCH[search], num_threads,
                                 reg_scale, int(use_line_search), outlier_ratio, min_pts, eig_mult)

    def gauss(self):
        d = np.zeros(3); lib().orc_ndt_gauss(self.h, _d(d)); return d

    def set_target(self, cloud, is_dense=True):
        a = as_xyz(cloud)
        return lib().orc_ndt_set_target(self.h, _f(a), a.shape[0], a.shape[1], int(is_dense))

    def set_source(self, cloud):
        a = as_xyz(cloud)
        lib().orc_ndt_set_source(self.h, _f(a), a.shape[0], a.shape[1])

    def set_regularization_pose(self, T):
        if T is None:
            lib().orc_ndt_set_regularization_pose(self.h, None)
        else:
            t = np.ascontiguousarray(T, np.float32).reshape(16)
            lib().orc_ndt_set_regularization_pose(self.h, _f(t))

    def grid(self):
        return export_grid(lib().orc_ndt_grid(self.h))

    def compute_derivatives(self, p, compute_hessian=True, T=None):
        p = np.ascontiguousarray(p, np.float64)
        g = np.zeros(6); H = np.zeros((6, 6)); nv = np.zeros(1)
        if T is None:
            s = lib().orc_ndt_compute_derivatives(self.h, _d(p), int(compute_hessian), _d(g), _d(H), _d(nv))
        else:
            t = np.ascontiguousarray(T, np.float32).reshape(16)
            s = lib().orc_ndt_compute_derivatives_T(self.h, _f(t), _d(p), int(compute_hessian), _d(g), _d(H), _d(nv))
        return s, g, H, nv[0]

    def calculate_scores(self, T):
        t = np.ascontiguousarray(T, np.float32).reshape(16); tp = np.zeros(1); nv = np.zeros(1)
        lib().orc_ndt_calculate_scores(self.h, _f(t), _d(tp), _d(nv))
        return tp[0], nv[0]

    def calculate_score(self, T):
        t = np.ascontiguousarray(T, np.float32).reshape(16)
        nv = C.c_size_t(); ne = C.c_size_t()
        lib().orc_ndt_calculate_score(self.h, _f(t), 0, None, None, C.byref(nv), 0, None, C.byref(ne))
        idx = np.zeros(nv.value, np.int64); val = np.zeros(nv.value); emp = np.zeros((ne.value, 3), np.int32)
        s = lib().orc_ndt_calculate_score(self.h, _f(t), nv.value, idx.ctypes.data_as(c_ll), _d(val), C.byref(nv), ne.value, _i(emp), C.byref(ne))
        return s, dict(zip(idx.tolist(), val.tolist())), emp

    def write_files(self, ellipsoid_path, voxel_path):
        return lib().orc_ndt_write_files(self.h, str(ellipsoid_path).encode(), str(voxel_path).encode())

    def align(self, guess=None, max_trace=128):
        guess = np.eye(4, dtype=np.float32) if guess is None else np.ascontiguousarray(guess, np.float32)
        T = np.zeros((4, 4), np.float32); H = np.zeros((6, 6)); tp = np.zeros(1); nv = np.zeros(1)
        it = np.zeros(1, np.int32); cv = np.zeros(1, np.int32); nt = np.zeros(1, np.int32); na = np.zeros(1, np.int32)
        trace = np.zeros((max_trace, 50)); tarr = np.zeros((max_trace, 16), np.float32)
        lib().orc_ndt_align(self.h, _f(guess.reshape(16)), _f(T), _d(H), _d(tp), _d(nv), _i(it), _i(cv), _d(trace), max_trace, _i(nt),
                            _f(tarr), max_trace, _i(na))
        return dict(T=T, hessian=H, trans_probability=tp[0], nvtl=nv[0], iterations=int(it[0]), converged=bool(cv[0]),
                    trace=trace[: nt[0]].copy(), transformation_array=tarr[: na[0]].reshape(-1, 4, 4).copy())

    def __del__(self):
        try:
            lib().orc_ndt_destroy(self.h)
        except Exception:
            pass


class OracleMultigrid:
    """pclomp::MultiGridNormalDistributionsTransform restatement (oracle/multigrid_ndt.hpp)."""

    def __init__(self, resolution=1.0, step_size=0.1, trans_epsilon=0.1, max_iterations=35, num_threads=1, reg_scale=0.0,
                 use_line_search=False, outlier_ratio=0.55, min_pts=6, eig_mult=0.01):
        self.h = lib().orc_mg_create()
        lib().orc_mg_set_params(self.h, resolution, step_size, trans_epsilon, max_iterations, num_threads, reg_scale, int(use_line_search),
                                outlier_ratio, min_pts, eig_mult)

    def add_target(self, cloud, grid_id, is_dense=True):
        a = as_xyz(cloud)
        return lib().orc_mg_add_target(self.h, _f(a), a.shape[0], a.shape[1], int(is_dense), str(grid_id).encode())

    def remove_target(self, grid_id):
        lib().orc_mg_remove_target(self.h, str(grid_id).encode())

    def create_kdtree(self):
        lib().orc_mg_create_kdtree(self.h)

    def set_input_target(self, cloud):          # multigrid_ndt_omp.h:143-148
        rc = self.add_target(cloud, "default"); self.create_kdtree(); return rc

    def leaves(self):
        n = lib().orc_mg_leaf_count(self.h)
        npts = np.zeros(n, np.int32); mean = np.zeros((n, 3)); icov = np.zeros((n, 9)); cen = np.zeros((n, 3), np.float32)
        lib().orc_mg_export(self.h, n, _i(npts), _d(mean), _d(icov), _f(cen))
        return dict(n=npts, mean=mean, icov=icov, centroid=cen)

    def radius_search(self, p, cap=128):
        means = np.zeros((cap, 3)); dist = np.zeros(cap, np.float32)
        k = lib().orc_mg_radius_search(self.h, float(p[0]), float(p[1]), float(p[2]), cap, _d(means), _f(dist))
        return means[:k].copy(), dist[:k].copy()

    def set_source(self, cloud):
        a = as_xyz(cloud)
        lib().orc_mg_set_source(self.h, _f(a), a.shape[0], a.shape[1])

    def set_regularization_pose(self, T):
        if T is None:
            lib().orc_mg_set_regularization_pose(self.h, None)
        else:
            t = np.ascontiguousarray(T, np.float32).reshape(16)
            lib().orc_mg_set_regularization_pose(self.h, _f(t))

    def compute_derivatives(self, p, compute_hessian=True):
        p = np.ascontiguousarray(p, np.float64)
        g = np.zeros(6); H = np.zeros((6, 6)); nv = np.zeros(1)
        s = lib().orc_mg_compute_derivatives(self.h, _d(p), int(compute_hessian), _d(g), _d(H), _d(nv))
        return s, g, H, nv[0]

    def calculate_scores(self, T):
        t = np.ascontiguousarray(T, np.float32).reshape(16); tp = np.zeros(1); nv = np.zeros(1)
        lib().orc_mg_calculate_scores(self.h, _f(t), _d(tp), _d(nv))
        return tp[0], nv[0]

    def align(self, guess=None, max_trace=128):
        guess = np.eye(4, dtype=np.float32) if guess is None else np.ascontiguousarray(guess, np.float32)
        T = np.zeros((4, 4), np.float32); H = np.zeros((6, 6)); tp = np.zeros(1); nv = np.zeros(1)
        it = np.zeros(1, np.int32); cv = np.zeros(1, np.int32); nt = np.zeros(1, np.int32); na = np.zeros(1, np.int32)
        trace = np.zeros((max_trace, 50)); tarr = np.zeros((max_trace, 16), np.float32)
        rc = lib().orc_mg_align(self.h, _f(guess.reshape(16)), _f(T), _d(H), _d(tp), _d(nv), _i(it), _i(cv), _d(trace), max_trace, _i(nt),
                                _f(tarr), max_trace, _i(na))
        if rc != 0:
            raise RuntimeError("kd-tree is stale: call create_kdtree() after add_target / remove_target")
        return dict(T=T, hessian=H, trans_probability=tp[0], nvtl=nv[0], iterations=int(it[0]), converged=bool(cv[0]),
                    trace=trace[: nt[0]].copy(), transformation_array=tarr[: na[0]].reshape(-1, 4, 4).copy())

    def __del__(self):
        try:
            lib().orc_mg_destroy(self.h)
        except Exception:
            pass


class OracleSvn:
    """svn_ndt::SvnNormalDistributionsTransform restatement."""

    def __init__(self, resolution=1.0, outlier_ratio=0.55, search="DIRECT7", num_threads=1, K=30, max_iter=50, kernel_h=1.0,
                 step_size=1.0, stop_thresh=1e-4, use_gn=True, min_pts=6, eig_mult=0.01):
        self.h = lib().orc_svn_create()
        self.K = K
        lib().orc_svn_set_params(self.h, resolution, outlier_ratio, SEARCH[search], num_threads, K, max_iter, kernel_h, step_size,
                                 stop_thresh, int(use_gn), min_pts, eig_mult)

    def gauss(self):
        d = np.zeros(3); lib().orc_svn_gauss(self.h, _d(d)); return d

    def set_target(self, cloud):
        a = as_xyz(cloud)
        return lib().orc_svn_set_target(self.h, _f(a), a.shape[0], a.shape[1])

    def set_source(self, cloud):
        a = as_xyz(cloud)
        lib().orc_svn_set_source(self.h, _f(a), a.shape[0], a.shape[1])

    def grid(self):
        return export_grid(lib().orc_svn_grid(self.h))

    @staticmethod
    def sample_particles(prior, K, seed):
        prior = np.ascontiguousarray(prior, np.float64).reshape(16)
        out = np.zeros((K, 4, 4))
        lib().orc_svn_sample_particles(_d(prior), K, seed, _d(out))
        return out

    def particle_derivatives(self, pose):
        pose = np.ascontiguousarray(pose, np.float64).reshape(16)
        g = np.zeros(6); H = np.zeros((6, 6))
        s = lib().orc_svn_particle_derivatives(self.h, _d(pose), _d(g), _d(H))
        return s, g, H

    def step(self, particles):
        P = np.ascontiguousarray(particles, np.float64).copy(); K = P.shape[0]
        s = np.zeros(K); g = np.zeros((K, 6)); H = np.zeros((K, 6, 6)); u = np.zeros((K, 6))
        lib().orc_svn_step(self.h, K, _d(P), _d(s), _d(g), _d(H), _d(u))
        return P, s, g, H, u

    def stage2(self, particles, grads, hess):
        P = np.ascontiguousarray(particles, np.float64); K = P.shape[0]
        g = np.ascontiguousarray(grads, np.float64); H = np.ascontiguousarray(hess, np.float64); u = np.zeros((K, 6))
        lib().orc_svn_stage2(self.h, K, _d(P), _d(g), _d(H), _d(u))
        return u

    def align(self, prior, particles):
        prior = np.ascontiguousarray(prior, np.float64).reshape(16)
        P = np.ascontiguousarray(particles, np.float64).copy(); K = P.shape[0]
        fp = np.zeros((4, 4)); cov = np.zeros((6, 6)); cv = np.zeros(1, np.int32)
        it = lib().orc_svn_align(self.h, _d(prior), K, _d(P), _d(fp), _d(cov), _i(cv))
        return dict(final_pose=fp, final_covariance=cov, converged=bool(cv[0]), iterations=it, particles=P)

    def __del__(self):
        try:
            lib().orc_svn_destroy(self.h)
        except Exception:
            pass


class OracleLidar:
    """oracle/ouster_decode.hpp: LidarCallback::Initialize + DecodePacketRng19 restated (src/lidarcallback.cpp:191-341, 631-878).
    cfg: any object with the fields of ndtb_lidar_config (the ctypes struct of the product binding is accepted)."""

    def __init__(self, cfg, az_deg, alt_deg):
        self.L = lib()
        self.cfg = cfg
        ci = np.array([cfg.columns_per_frame, cfg.pixels_per_column, cfg.columns_per_packet, cfg.channel_stride, cfg.reflectivity_threshold,
                       int(getattr(cfg, "udp_profile_legacy", 0))], np.int32)
        cf = np.array([cfg.lidar_origin_to_beam_origin_mm, cfg.range_min, cfg.range_max, cfg.z_min, cfg.z_max, *cfg.box_center, *cfg.box_dimensions], np.float32)
        cd = np.array([*cfg.body_to_lidar_rotation, *cfg.body_to_lidar_translation], np.float64)
        az = np.ascontiguousarray(az_deg, np.float32); alt = np.ascontiguousarray(alt_deg, np.float32)
        self.h = self.L.orc_lidar_create(_i(ci), _f(cf), _d(cd), _f(az), _f(alt))
        if not self.h:
            raise ValueError("the reference's LidarCallback constructor would throw for this configuration")
        self.expected_size = self.L.orc_lidar_packet_bytes(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_lidar_destroy(self.h)
            self.h = None

    def tables(self):
        W, Hh = self.cfg.columns_per_frame, self.cfg.pixels_per_column
        t = [np.empty(W * Hh, np.float32) for _ in range(3)] + [np.empty(W, np.float32) for _ in range(3)]
        self.L.orc_lidar_tables(self.h, *[_f(a) for a in t])
        return t

    def decode(self, packets):
        """DecodePacket for every row of `packets` (uint8, (n, size)); returns the frames completed by these packets."""
        a = np.ascontiguousarray(packets, np.uint8)
        self.L.orc_lidar_clear_frames(self.h)
        n = self.L.orc_lidar_decode(self.h, a.ctypes.data, a.shape[0], a.shape[1], a.shape[1])
        frames = []
        u16 = C.POINTER(C.c_uint16)
        for k in range(n):
            info = np.zeros(5)
            self.L.orc_lidar_frame_info(self.h, k, _d(info))
            m = int(info[4])
            f = {"x": np.empty(m, np.float32), "y": np.empty(m, np.float32), "z": np.empty(m, np.float32), "c_id": np.empty(m, np.uint16),
                 "m_id": np.empty(m, np.uint16), "timestamp_points": np.empty(m, np.float64), "relative_timestamp": np.empty(m, np.float32),
                 "reflectivity": np.empty(m, np.uint16), "signal": np.empty(m, np.uint16), "nir": np.empty(m, np.uint16)}
            self.L.orc_lidar_frame_get(self.h, k, _f(f["x"]), _f(f["y"]), _f(f["z"]), f["c_id"].ctypes.data_as(u16), f["m_id"].ctypes.data_as(u16),
                                       _d(f["timestamp_points"]), _f(f["relative_timestamp"]), f["reflectivity"].ctypes.data_as(u16),
                                       f["signal"].ctypes.data_as(u16), f["nir"].ctypes.data_as(u16))
            f.update(frame_id=int(info[0]), timestamp=info[1], timestamp_end=info[2], interframe_timedelta=info[3], numberpoints=m)
            frames.append(f)
        return frames
