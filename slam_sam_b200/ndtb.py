"""ctypes binding of libndtb.so -- the C-ABI CUDA library (include/ndtb.h).

There is no CPU fallback: importing works anywhere (so the symbol table can be checked without a
GPU), but creating a handle without a CUDA device raises NdtbError, and a missing library raises
ImportError-like RuntimeError loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NDTB_LIB", os.path.join(_HERE, "libndtb.so"))  # NDTB_LIB: development builds of the same CUDA library

NDTB_OK = 0
NDTB_ERR_INVALID, NDTB_ERR_CUDA, NDTB_ERR_EMPTY, NDTB_ERR_OVERFLOW, NDTB_ERR_NOT_READY = -1, -2, -3, -4, -5
ERR ={-1: "NDTB_ERR_INVALID", -2: "NDTB_ERR_CUDA", -3: "NDTB_ERR_EMPTY", -4: "NDTB_ERR_OVERFLOW", -5: "NDTB_ERR_NOT_READY"}
KDTREE, DIRECT26, DIRECT7, DIRECT1 = 0, 1, 2, 3
FLAVOUR_PCLOMP, FLAVOUR_SVN = 0, 1

c_f = C.POINTER(C.c_float)
c_d = C.POINTER(C.c_double)
c_i = C.POINTER(C.c_int)


class Params(C.Structure):
    _fields_ = [("resolution", C.c_float), ("step_size", C.c_double), ("trans_epsilon", C.c_double), ("max_iterations", C.c_int),
                ("search_method", C.c_int), ("outlier_ratio", C.c_double), ("min_points_per_voxel", C.c_int),
                ("min_covar_eigvalue_mult", C.c_double), ("regularization_scale_factor", C.c_float), ("use_line_search", C.c_int),
                ("grid_flavour", C.c_int), ("is_dense", C.c_int)]


class Result(C.Structure):
    _fields_ = [("pose", C.c_float * 16), ("hessian", C.c_double * 36), ("transform_probability", C.c_double),
                ("nearest_voxel_transformation_likelihood", C.c_double), ("iteration_num", C.c_int), ("converged", C.c_int),
                ("transformation_array_len", C.c_int), ("evaluations", C.c_int)]


class SvnParams(C.Structure):
    _fields_ = [("particle_count", C.c_int), ("max_iterations", C.c_int), ("kernel_bandwidth", C.c_double), ("step_size", C.c_double),
                ("stop_threshold", C.c_double), ("use_gauss_newton", C.c_int)]


class SvnResult(C.Structure):
    _fields_ = [("final_pose", C.c_double * 16), ("final_covariance", C.c_double * 36), ("converged", C.c_int), ("iterations", C.c_int)]


# every symbol include/ndtb.h declares (tests check the library exports all of them)
SYMBOLS = ["ndtb_create", "ndtb_destroy", "ndtb_last_error", "ndtb_version", "ndtb_default_params", "ndtb_set_params", "ndtb_get_params",
           "ndtb_set_target", "ndtb_set_source", "ndtb_set_target_device", "ndtb_set_source_device", "ndtb_set_regularization_pose",
           "ndtb_compute_derivatives", "ndtb_align", "ndtb_get_trace", "ndtb_calculate_scores", "ndtb_get_voxel_count", "ndtb_get_voxels",
           "ndtb_get_last_hits", "ndtb_svn_default_params", "ndtb_svn_set_params", "ndtb_svn_particle_derivatives", "ndtb_svn_stage2",
           "ndtb_svn_step", "ndtb_svn_align", "ndtb_svn_stage1_local", "ndtb_svn_finish_step", "ndtb_align_batch", "ndtb_get_timings",
           "ndtb_get_stream", "ndtb_get_launch_count", "ndtb_get_session_stats", "ndtb_set_profiling", "ndtb_set_option", "ndtb_time_eval_kernel", "ndtb_newton_direction", "ndtb_hessian_to_covariance", "ndtb_keyframe_put", "ndtb_keyframe_drop", "ndtb_keyframe_count",
           "ndtb_set_target_window", "ndtb_get_target_points", "ndtb_voxelgrid_filter", "ndtb_svn_mean_pose", "ndtb_svn_pose_distance",
           "ndtb_lidar_default_config", "ndtb_lidar_create", "ndtb_lidar_destroy", "ndtb_lidar_last_error", "ndtb_lidar_packet_bytes",
           "ndtb_lidar_get_tables", "ndtb_lidar_push_packets", "ndtb_lidar_pop_frame", "ndtb_lidar_read_frame", "ndtb_lidar_frame_points",
           "ndtb_lidar_sync", "ndtb_lidar_get_stats",
           "ndtb_mg_add_target", "ndtb_mg_remove_target", "ndtb_mg_create_kdtree", "ndtb_mg_set_input_target", "ndtb_mg_get_info",
           "ndtb_mg_get_grid_ids", "ndtb_mg_get_leaves", "ndtb_mg_compute_derivatives", "ndtb_mg_align", "ndtb_mg_calculate_scores",
           "ndtb_svn_get_timings", "ndtb_comm_window_handle", "ndtb_comm_connect", "ndtb_comm_init", "ndtb_comm_close", "ndtb_comm_info",
           "ndtb_svn_align_sharded", "ndtb_calculate_score", "ndtb_write_ndt_files", "ndtb_compute_derivatives_static", "ndtb_stage_target", "ndtb_stage_source",
           "ndtb_commit_target", "ndtb_commit_source", "ndtb_staged_count", "ndtb_stage_sync", "ndtb_test_newton_step"]

_LIB = None


class LidarConfig(C.Structure):
    """ndtb_lidar_config (include/ndtb.h)."""
    _fields_ = [("columns_per_frame", C.c_int), ("pixels_per_column", C.c_int), ("columns_per_packet", C.c_int), ("channel_stride", C.c_int),
                ("reflectivity_threshold", C.c_int), ("lidar_origin_to_beam_origin_mm", C.c_float), ("range_min", C.c_float),
                ("range_max", C.c_float), ("z_min", C.c_float), ("z_max", C.c_float), ("box_center", C.c_float * 3),
                ("box_dimensions", C.c_float * 3), ("body_to_lidar_rotation", C.c_double * 9), ("body_to_lidar_translation", C.c_double * 3),
                ("udp_profile_legacy", C.c_int)]


class LidarFrameInfo(C.Structure):
    """ndtb_lidar_frame_info (include/ndtb.h)."""
    _fields_ = [("timestamp", C.c_double), ("timestamp_end", C.c_double), ("interframe_timedelta", C.c_double), ("numberpoints", C.c_uint32),
                ("frame_id", C.c_uint16), ("packets", C.c_uint16)]


class NdtbError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__(f"{ERR.get(code, code)}: {msg}")
        self.code = code


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        H = C.c_void_p
        L.ndtb_create.argtypes = [C.c_int, C.POINTER(H)]
        L.ndtb_destroy.argtypes = [H]
        L.ndtb_last_error.restype = C.c_char_p
        L.ndtb_last_error.argtypes = [H]
        L.ndtb_version.restype = C.c_char_p
        L.ndtb_default_params.argtypes = [C.c_int, C.POINTER(Params)]
        L.ndtb_set_params.argtypes = [H, C.POINTER(Params)]
        L.ndtb_get_params.argtypes = [H, C.POINTER(Params)]
        L.ndtb_set_target.argtypes = [H, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_set_source.argtypes = [H, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_set_target_device.argtypes = [H, C.c_void_p, C.c_size_t]
        L.ndtb_set_source_device.argtypes = [H, C.c_void_p, C.c_size_t]
        L.ndtb_stage_target.argtypes = [H, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_stage_source.argtypes = [H, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_commit_target.argtypes = [H]
        L.ndtb_commit_source.argtypes = [H]
        L.ndtb_stage_sync.argtypes = [H]
        L.ndtb_test_newton_step.argtypes = [C.c_int, C.c_int, c_d, c_d, c_d, c_d, C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                            c_d, c_f, c_d]
        L.ndtb_staged_count.argtypes = [H, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ndtb_set_regularization_pose.argtypes = [H, c_f]
        L.ndtb_compute_derivatives.argtypes = [H, c_d, c_f, C.c_int, c_d, c_d, c_d, c_d]
        L.ndtb_compute_derivatives_static.argtypes = [H, C.c_void_p, C.c_size_t, c_d, c_f, C.c_int, c_d, c_d, c_d, c_d]
        L.ndtb_align.argtypes = [H, c_f, C.POINTER(Result), c_f, C.c_int, c_f]
        L.ndtb_get_trace.argtypes = [H, c_d, C.c_int, c_i]
        L.ndtb_calculate_scores.argtypes = [H, c_f, c_d, c_d]
        L.ndtb_calculate_score.argtypes = [H, c_f, c_d, C.c_size_t, C.POINTER(C.c_int64), c_d, C.POINTER(C.c_size_t), C.c_size_t,
                                           C.POINTER(C.c_int32), C.POINTER(C.c_size_t)]
        L.ndtb_write_ndt_files.argtypes = [H, C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_get_voxel_count.argtypes = [H, C.POINTER(C.c_size_t), c_i, c_i, c_i]
        L.ndtb_get_voxels.argtypes = [H, C.POINTER(C.c_int64), C.POINTER(C.c_int32), c_d, c_d, c_d, c_d, c_d]
        L.ndtb_get_last_hits.argtypes = [H, c_d, c_d]
        L.ndtb_svn_default_params.argtypes = [C.POINTER(SvnParams)]
        L.ndtb_svn_set_params.argtypes = [H, C.POINTER(SvnParams)]
        L.ndtb_svn_particle_derivatives.argtypes = [H, C.c_int, c_d, c_d, c_d, c_d]
        L.ndtb_svn_stage2.argtypes = [H, C.c_int, c_d, c_d, c_d, c_d]
        L.ndtb_svn_step.argtypes = [H, C.c_int, c_d, c_d, c_d, c_d, c_d]
        L.ndtb_svn_align.argtypes = [H, c_d, c_d, C.c_uint64, C.POINTER(SvnResult), c_d]
        L.ndtb_svn_align_sharded.argtypes = [H, c_d, c_d, C.c_uint64, C.POINTER(SvnResult), c_d]
        L.ndtb_svn_get_timings.argtypes = [H, c_f, c_f, c_f, c_i]
        L.ndtb_comm_window_handle.argtypes = [H, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.ndtb_comm_connect.argtypes = [H, C.c_int, C.c_int, C.c_void_p]
        L.ndtb_comm_init.argtypes = [C.POINTER(H), C.c_int]
        L.ndtb_comm_close.argtypes = [H]
        L.ndtb_comm_info.argtypes = [H, c_i, c_i]
        L.ndtb_svn_stage1_local.argtypes = [H, C.c_int, c_d, C.c_int, C.c_int, c_d]
        L.ndtb_svn_finish_step.argtypes = [H, C.c_int, c_d, c_d, c_d]
        L.ndtb_align_batch.argtypes = [H, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_size_t, c_f, C.POINTER(Result)]
        L.ndtb_get_timings.argtypes = [H, c_f, c_f, c_f, c_i]
        L.ndtb_get_stream.argtypes = [H, C.POINTER(C.c_void_p)]
        L.ndtb_get_launch_count.argtypes = [H, C.POINTER(C.c_uint64)]
        L.ndtb_get_session_stats.argtypes = [H, C.POINTER(C.c_ulonglong)]
        L.ndtb_set_profiling.argtypes = [H, C.c_int]
        L.ndtb_set_option.argtypes = [H, C.c_char_p, C.c_double]
        L.ndtb_time_eval_kernel.argtypes = [H, C.POINTER(C.c_double), C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float)]
        L.ndtb_hessian_to_covariance.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.ndtb_keyframe_put.argtypes = [H, C.c_int64, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_keyframe_drop.argtypes = [H, C.c_int64]
        L.ndtb_keyframe_count.argtypes = [H, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
        L.ndtb_set_target_window.argtypes = [H, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_double)]
        L.ndtb_get_target_points.argtypes = [H, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.ndtb_voxelgrid_filter.argtypes = [H, C.c_void_p, C.c_size_t, C.c_size_t, C.c_float, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.ndtb_newton_direction.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.ndtb_svn_mean_pose.argtypes = [C.c_int, c_d, c_d, c_d]
        L.ndtb_svn_pose_distance.argtypes = [c_d, c_d, c_d]
        L.ndtb_mg_add_target.argtypes = [H, C.c_char_p, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_mg_remove_target.argtypes = [H, C.c_char_p]
        L.ndtb_mg_create_kdtree.argtypes = [H]
        L.ndtb_mg_set_input_target.argtypes = [H, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ndtb_mg_get_info.argtypes = [H, c_i, c_i, C.POINTER(C.c_size_t), c_i]
        L.ndtb_mg_get_grid_ids.argtypes = [H, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.ndtb_mg_get_leaves.argtypes = [H, C.c_size_t, C.POINTER(C.c_int32), c_d, c_d, c_f, C.POINTER(C.c_size_t)]
        L.ndtb_mg_compute_derivatives.argtypes = [H, c_d, c_f, C.c_int, c_d, c_d, c_d, c_d]
        L.ndtb_mg_align.argtypes = [H, c_f, C.POINTER(Result), c_f, C.c_int, c_f]
        L.ndtb_mg_calculate_scores.argtypes = [H, c_f, c_d, c_d]
        c_u16 = C.POINTER(C.c_uint16)
        L.ndtb_lidar_default_config.argtypes = [C.POINTER(LidarConfig)]
        L.ndtb_lidar_create.argtypes = [C.c_int, C.POINTER(LidarConfig), c_f, c_f, C.POINTER(H)]
        L.ndtb_lidar_destroy.argtypes = [H]
        L.ndtb_lidar_last_error.restype = C.c_char_p
        L.ndtb_lidar_last_error.argtypes = [H]
        L.ndtb_lidar_packet_bytes.argtypes = [H, C.POINTER(C.c_size_t)]
        L.ndtb_lidar_get_tables.argtypes = [H, c_f, c_f, c_f, c_f, c_f, c_f]
        L.ndtb_lidar_push_packets.argtypes = [H, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, c_i]
        L.ndtb_lidar_pop_frame.argtypes = [H, C.POINTER(LidarFrameInfo), c_i]
        L.ndtb_lidar_read_frame.argtypes = [H, C.c_size_t, c_f, c_f, c_f, c_u16, c_u16, c_d, c_f, c_u16, c_u16, c_u16]
        L.ndtb_lidar_frame_points.argtypes = [H, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        L.ndtb_lidar_sync.argtypes = [H]
        L.ndtb_lidar_get_stats.argtypes = [H, C.POINTER(C.c_uint64), c_f]
        _LIB = L
    return _LIB


def _ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def host_cloud(cloud):
    """float32 (N, >=3) C-contiguous view; returns (array, n, stride_bytes)."""
    a = np.ascontiguousarray(cloud, dtype=np.float32)
    if a.ndim != 2 or a.shape[1] < 3:
        raise ValueError("cloud must be (N, >=3) float32")
    return a, a.shape[0], a.shape[1] * 4


def test_newton_step(which, g, H, p, step_max, step_min, eps, max_iterations, first, nr_iterations, a_t_prev, device=0):
    """ndtb_test_newton_step: (status, nr_iterations, a_t, state81 float32, next pose)."""
    g = np.ascontiguousarray(g, np.float64); H = np.ascontiguousarray(H, np.float64).reshape(36); p = np.ascontiguousarray(p, np.float64)
    prm = np.array([step_max, step_min, eps, 0.0])
    st, nr = C.c_int(0), C.c_int(0)
    a_t = np.zeros(1); state = np.zeros(81, np.float32); pn = np.zeros(6)
    rc = lib().ndtb_test_newton_step(which, device, _ptr(g, c_d), _ptr(H, c_d), _ptr(p, c_d), _ptr(prm, c_d), max_iterations, int(first), nr_iterations,
                                     float(a_t_prev), C.byref(st), C.byref(nr), _ptr(a_t, c_d), _ptr(state, c_f), _ptr(pn, c_d))
    if rc != NDTB_OK:
        raise NdtbError(rc, "ndtb_test_newton_step")
    return st.value, nr.value, a_t[0], state, pn


test_newton_step.__test__ = False   # not a pytest test


def svn_mean_pose(poses, prior):
    P = np.ascontiguousarray(poses, np.float64); pr = np.ascontiguousarray(prior, np.float64).reshape(16); out = np.zeros(16)
    rc = lib().ndtb_svn_mean_pose(P.shape[0], _ptr(P, c_d), _ptr(pr, c_d), _ptr(out, c_d))
    if rc != NDTB_OK:
        raise NdtbError(rc, "ndtb_svn_mean_pose")
    return out.reshape(4, 4)


def svn_pose_distance(a, b):
    a = np.ascontiguousarray(a, np.float64).reshape(16); b = np.ascontiguousarray(b, np.float64).reshape(16); n = np.zeros(1)
    rc = lib().ndtb_svn_pose_distance(_ptr(a, c_d), _ptr(b, c_d), _ptr(n, c_d))
    if rc != NDTB_OK:
        raise NdtbError(rc, "ndtb_svn_pose_distance")
    return float(n[0])


def hessian_to_covariance(H):
    """-(H + 1e-6 I)^-1 in GTSAM block order (run/pipeline_ligo_tc.cpp:536-542)."""
    H = np.ascontiguousarray(H, np.float64).reshape(36); out = np.zeros(36)
    rc = lib().ndtb_hessian_to_covariance(_ptr(H, c_d), _ptr(out, c_d))
    if rc != NDTB_OK:
        raise NdtbError(rc, "ndtb_hessian_to_covariance")
    return out.reshape(6, 6)


def newton_direction(H, b, method=1):
    """x = H^-1 b as ndtb_align computes it (host code).  Returns (x, used_direct)."""
    H = np.ascontiguousarray(H, np.float64).reshape(36); b = np.ascontiguousarray(b, np.float64).reshape(6)
    x = np.zeros(6); used = C.c_int(0)
    rc = lib().ndtb_newton_direction(_ptr(H, c_d), _ptr(b, c_d), int(method), _ptr(x, c_d), C.byref(used))
    if rc != NDTB_OK:
        raise NdtbError(rc, "ndtb_newton_direction")
    return x, bool(used.value)


def comm_init(handles):
    """ndtb_comm_init: one process, one handle per GPU -- the record windows are reached through peer access."""
    arr = (C.c_void_p * len(handles))(*[h.h for h in handles])
    rc = lib().ndtb_comm_init(arr, len(handles))
    if rc != NDTB_OK:
        raise NdtbError(rc, lib().ndtb_last_error(handles[0].h).decode())


class Handle:
    """Thin RAII wrapper over ndtb_handle."""

    def __init__(self, device=0):
        self.L = lib()
        self.h = C.c_void_p()
        rc = self.L.ndtb_create(device, C.byref(self.h))
        if rc != NDTB_OK:
            raise NdtbError(rc, "ndtb_create failed (no CUDA device? there is no CPU fallback)")
        self._keep = []

    def close(self):
        if getattr(self, "h", None):
            self.L.ndtb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc):
        if rc != NDTB_OK:
            raise NdtbError(rc, self.L.ndtb_last_error(self.h).decode())
        return rc

    # parameters
    def get_params(self):
        p = Params()
        self.check(self.L.ndtb_get_params(self.h, C.byref(p)))
        return p

    def set_params(self, **kw):
        p = self.get_params()
        for k, v in kw.items():
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
        return self.L.ndtb_set_params(self.h, C.byref(p))

    def set_svn_params(self, **kw):
        p = SvnParams()
        self.L.ndtb_svn_default_params(C.byref(p))
        for k, v in getattr(self, "_svn_kw", {}).items():
            setattr(p, k, v)
        for k, v in kw.items():
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
        self._svn_kw = {**getattr(self, "_svn_kw", {}), **kw}
        self.check(self.L.ndtb_svn_set_params(self.h, C.byref(p)))

    # clouds
    def set_target(self, cloud):
        a, n, s = host_cloud(cloud)
        return self.L.ndtb_set_target(self.h, a.ctypes.data, n, s)

    def set_source(self, cloud):
        a, n, s = host_cloud(cloud)
        return self.L.ndtb_set_source(self.h, a.ctypes.data, n, s)

    def stage_target(self, cloud):
        """Starts the upload of the NEXT target (keep `cloud` alive until commit_target); see ndtb_stage_target."""
        a, n, s = host_cloud(cloud)
        self._staged = getattr(self, "_staged", []) + [a]
        return self.L.ndtb_stage_target(self.h, a.ctypes.data, n, s)

    def stage_source(self, cloud):
        a, n, s = host_cloud(cloud)
        self._staged = getattr(self, "_staged", []) + [a]
        return self.L.ndtb_stage_source(self.h, a.ctypes.data, n, s)

    def commit_target(self):
        return self.L.ndtb_commit_target(self.h)

    def commit_source(self):
        return self.L.ndtb_commit_source(self.h)

    def staged_count(self):
        t, s = C.c_int(0), C.c_int(0)
        self.check(self.L.ndtb_staged_count(self.h, C.byref(t), C.byref(s)))
        if t.value == 0 and s.value == 0:
            self._staged = []
        return t.value, s.value

    def set_target_device(self, ptr, n):
        return self.L.ndtb_set_target_device(self.h, C.c_void_p(ptr), n)

    def set_source_device(self, ptr, n):
        return self.L.ndtb_set_source_device(self.h, C.c_void_p(ptr), n)

    def set_regularization_pose(self, T):
        t = None if T is None else np.ascontiguousarray(T, np.float32).reshape(16)
        self.check(self.L.ndtb_set_regularization_pose(self.h, _ptr(t, c_f)))

    # pclomp
    def compute_derivatives(self, p, T=None, compute_hessian=True):
        p = np.ascontiguousarray(p, np.float64)
        t = None if T is None else np.ascontiguousarray(T, np.float32).reshape(16)
        s = np.zeros(1); g = np.zeros(6); H = np.zeros((6, 6)); nv = np.zeros(1)
        self.check(self.L.ndtb_compute_derivatives(self.h, _ptr(p, c_d), _ptr(t, c_f), int(compute_hessian), _ptr(s, c_d), _ptr(g, c_d),
                                                   _ptr(H, c_d), _ptr(nv, c_d)))
        return s[0], g, H, nv[0]

    def compute_derivatives_static(self, p, T=None, compute_hessian=True, d_source=None, n_source=0):
        """ndtb_compute_derivatives_static: the same evaluation over the library's static scratch, no allocation in the call.
        `d_source` is a device address of n_source packed float4 points; None -> the handle's resident source."""
        p = np.ascontiguousarray(p, np.float64)
        t = None if T is None else np.ascontiguousarray(T, np.float32).reshape(16)
        s = np.zeros(1); g = np.zeros(6); H = np.zeros((6, 6)); nv = np.zeros(1)
        self.check(self.L.ndtb_compute_derivatives_static(self.h, C.c_void_p(d_source) if d_source else None, int(n_source), _ptr(p, c_d),
                                                          _ptr(t, c_f), int(compute_hessian), _ptr(s, c_d), _ptr(g, c_d), _ptr(H, c_d),
                                                          _ptr(nv, c_d)))
        return s[0], g, H, nv[0]

    def align(self, guess=None, capacity=64, want_output=False):
        guess = np.eye(4, dtype=np.float32) if guess is None else np.ascontiguousarray(guess, np.float32)
        r = Result()
        arr = np.zeros((capacity, 16), np.float32)
        rc = self.L.ndtb_align(self.h, _ptr(guess.reshape(16), c_f), C.byref(r), _ptr(arr, c_f), capacity, None)
        self.check(rc)
        return dict(T=np.array(r.pose, np.float32).reshape(4, 4), hessian=np.array(r.hessian).reshape(6, 6),
                    trans_probability=r.transform_probability, nvtl=r.nearest_voxel_transformation_likelihood,
                    iterations=r.iteration_num, converged=bool(r.converged), evaluations=r.evaluations,
                    transformation_array=arr[: r.transformation_array_len].reshape(-1, 4, 4).copy())

    # ---- multigrid NDT (pclomp::MultiGridNormalDistributionsTransform) ----
    def mg_add_target(self, cloud, grid_id):
        a, n, st = host_cloud(cloud)
        return self.L.ndtb_mg_add_target(self.h, str(grid_id).encode(), a.ctypes.data, n, st)

    def mg_remove_target(self, grid_id):
        return self.L.ndtb_mg_remove_target(self.h, str(grid_id).encode())

    def mg_create_kdtree(self):
        return self.L.ndtb_mg_create_kdtree(self.h)

    def mg_set_input_target(self, cloud):
        a, n, st = host_cloud(cloud)
        return self.L.ndtb_mg_set_input_target(self.h, a.ctypes.data, n, st)

    def mg_info(self):
        a = np.zeros(1, np.int32); b = np.zeros(1, np.int32); c = C.c_size_t(0); d = np.zeros(1, np.int32)
        self.check(self.L.ndtb_mg_get_info(self.h, _ptr(a, c_i), _ptr(b, c_i), C.byref(c), _ptr(d, c_i)))
        return dict(grids=int(a[0]), grids_in_tree=int(b[0]), leaves=int(c.value), current=bool(d[0]))

    def mg_grid_ids(self):
        need = C.c_size_t(0)
        self.check(self.L.ndtb_mg_get_grid_ids(self.h, None, 0, C.byref(need)))
        buf = C.create_string_buffer(need.value)
        self.check(self.L.ndtb_mg_get_grid_ids(self.h, buf, need.value, None))
        txt = buf.value.decode()
        return txt.split("\n") if txt else []

    def mg_leaves(self):
        cnt = C.c_size_t(0)
        self.check(self.L.ndtb_mg_get_leaves(self.h, 0, None, None, None, None, C.byref(cnt)))
        k = cnt.value
        n = np.zeros(k, np.int32); mean = np.zeros((k, 3)); icov = np.zeros((k, 9)); cen = np.zeros((k, 3), np.float32)
        self.check(self.L.ndtb_mg_get_leaves(self.h, k, _ptr(n, C.POINTER(C.c_int32)), _ptr(mean, c_d), _ptr(icov, c_d), _ptr(cen, c_f), None))
        return dict(n=n, mean=mean, icov=icov, centroid=cen)

    def mg_compute_derivatives(self, p, T=None, compute_hessian=True):
        p = np.ascontiguousarray(p, np.float64)
        t = None if T is None else np.ascontiguousarray(T, np.float32).reshape(16)
        s = np.zeros(1); g = np.zeros(6); H = np.zeros((6, 6)); nv = np.zeros(1)
        self.check(self.L.ndtb_mg_compute_derivatives(self.h, _ptr(p, c_d), _ptr(t, c_f), int(compute_hessian), _ptr(s, c_d), _ptr(g, c_d),
                                                      _ptr(H, c_d), _ptr(nv, c_d)))
        return s[0], g, H, nv[0]

    def mg_calculate_scores(self, T):
        t = np.ascontiguousarray(T, np.float32).reshape(16); tp = np.zeros(1); nv = np.zeros(1)
        self.check(self.L.ndtb_mg_calculate_scores(self.h, _ptr(t, c_f), _ptr(tp, c_d), _ptr(nv, c_d)))
        return tp[0], nv[0]

    def mg_align(self, guess=None, capacity=64):
        guess = np.eye(4, dtype=np.float32) if guess is None else np.ascontiguousarray(guess, np.float32)
        r = Result()
        arr = np.zeros((capacity, 16), np.float32)
        self.check(self.L.ndtb_mg_align(self.h, _ptr(guess.reshape(16), c_f), C.byref(r), _ptr(arr, c_f), capacity, None))
        return dict(T=np.array(r.pose, np.float32).reshape(4, 4), hessian=np.array(r.hessian).reshape(6, 6),
                    trans_probability=r.transform_probability, nvtl=r.nearest_voxel_transformation_likelihood,
                    iterations=r.iteration_num, converged=bool(r.converged), evaluations=r.evaluations,
                    transformation_array=arr[: r.transformation_array_len].reshape(-1, 4, 4).copy())

    def align_batch(self, scans, guesses):
        """ndtb_align_batch: n independent scans (each float32 (n_i, 4)) against the current target; one result per scan."""
        arrs = [np.ascontiguousarray(sc, np.float32) for sc in scans]
        assert all(a.ndim == 2 and a.shape[1] == 4 for a in arrs)
        ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        ns = (C.c_size_t * len(arrs))(*[a.shape[0] for a in arrs])
        g = np.ascontiguousarray(guesses, np.float32).reshape(len(arrs), 16)
        res = (Result * len(arrs))()
        self.check(self.L.ndtb_align_batch(self.h, len(arrs), ptrs, ns, 16, _ptr(g, c_f), res))
        return [dict(T=np.array(r.pose, np.float32).reshape(4, 4), hessian=np.array(r.hessian).reshape(6, 6),
                     trans_probability=r.transform_probability, nvtl=r.nearest_voxel_transformation_likelihood,
                     iterations=r.iteration_num, converged=bool(r.converged), evaluations=r.evaluations) for r in res]

    def trace(self, capacity=128):
        t = np.zeros((capacity, 50)); n = np.zeros(1, np.int32)
        self.check(self.L.ndtb_get_trace(self.h, _ptr(t, c_d), capacity, _ptr(n, c_i)))
        return t[: min(n[0], capacity)].copy()

    def calculate_scores(self, T):
        t = np.ascontiguousarray(T, np.float32).reshape(16)
        tp = np.zeros(1); nv = np.zeros(1)
        self.check(self.L.ndtb_calculate_scores(self.h, _ptr(t, c_f), _ptr(tp, c_d), _ptr(nv, c_d)))
        return tp[0], nv[0]

    def calculate_score(self, T):
        """calculateScore (ndt_omp_impl.hpp:1083-1167): (score, {voxel index: mean score}, sorted empty voxel ids (m, 3))."""
        t = np.ascontiguousarray(T, np.float32).reshape(16)
        sc = np.zeros(1); nv = C.c_size_t(); ne = C.c_size_t()
        self.check(self.L.ndtb_calculate_score(self.h, _ptr(t, c_f), _ptr(sc, c_d), 0, None, None, C.byref(nv), 0, None, C.byref(ne)))
        idx = np.zeros(nv.value, np.int64); val = np.zeros(nv.value); emp = np.zeros((ne.value, 3), np.int32)
        self.check(self.L.ndtb_calculate_score(self.h, _ptr(t, c_f), _ptr(sc, c_d), nv.value, idx.ctypes.data_as(C.POINTER(C.c_int64)), _ptr(val, c_d),
                                               C.byref(nv), ne.value, emp.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(ne)))
        return sc[0], dict(zip(idx.tolist(), val.tolist())), emp

    def write_ndt_files(self, ellipsoid_path, voxel_path, points_path=None, points=None):
        """writeNdtDataToFiles (include/pipeline.hpp:224-274): ndt_ellipsoids.txt / ndt_voxels.txt (/ the map points)."""
        if points is not None and points_path is not None:
            a, n, st = host_cloud(points)
            return self.check(self.L.ndtb_write_ndt_files(self.h, str(ellipsoid_path).encode(), str(voxel_path).encode(), str(points_path).encode(), a.ctypes.data, n, st))
        return self.check(self.L.ndtb_write_ndt_files(self.h, str(ellipsoid_path).encode(), str(voxel_path).encode(), None, None, 0, 0))

    # voxel map
    def voxels(self):
        cnt = C.c_size_t()
        mn = np.zeros(3, np.int32); mx = np.zeros(3, np.int32); dv = np.zeros(3, np.int32)
        self.check(self.L.ndtb_get_voxel_count(self.h, C.byref(cnt), _ptr(mn, c_i), _ptr(mx, c_i), _ptr(dv, c_i)))
        V = cnt.value
        idx = np.zeros(V, np.int64); n = np.zeros(V, np.int32)
        mean = np.zeros((V, 3)); cov = np.zeros((V, 9)); icov = np.zeros((V, 9)); evecs = np.zeros((V, 9)); evals = np.zeros((V, 3))
        if V:
            self.check(self.L.ndtb_get_voxels(self.h, idx.ctypes.data_as(C.POINTER(C.c_int64)), n.ctypes.data_as(C.POINTER(C.c_int32)),
                                              _ptr(mean, c_d), _ptr(cov, c_d), _ptr(icov, c_d), _ptr(evecs, c_d), _ptr(evals, c_d)))
        return dict(idx=idx, n=n, mean=mean, cov=cov, icov=icov, evecs=evecs, evals=evals, min_b=mn, max_b=mx, div_b=dv)

    def last_hits(self):
        a = np.zeros(1); b = np.zeros(1)
        self.L.ndtb_get_last_hits(self.h, _ptr(a, c_d), _ptr(b, c_d))
        return a[0], b[0]

    def timings(self):
        b = np.zeros(1, np.float32); a = np.zeros(1, np.float32); e = np.zeros(1, np.float32); n = np.zeros(1, np.int32)
        self.L.ndtb_get_timings(self.h, _ptr(b, c_f), _ptr(a, c_f), _ptr(e, c_f), _ptr(n, c_i))
        return dict(build_ms=float(b[0]), align_ms=float(a[0]), eval_ms_total=float(e[0]), eval_launches=int(n[0]))

    def launch_count(self):
        n = C.c_uint64()
        self.L.ndtb_get_launch_count(self.h, C.byref(n))
        return n.value

    def session_stats(self):
        """ndtb_get_session_stats: device Newton steps confirmed by the host / handed to the host / comparisons that failed."""
        st = (C.c_ulonglong * 4)()
        self.check(self.L.ndtb_get_session_stats(self.h, st))
        return dict(accepted=int(st[0]), declined=int(st[1]), mismatches=int(st[2]), libm_port_ok=bool(st[3]))

    def set_profiling(self, on):
        self.check(self.L.ndtb_set_profiling(self.h, int(bool(on))))

    def keyframe_put(self, kid, cloud):
        """Upload a keyframe cloud once (pointsArchive entry); float32 (n, 3|4|8) array."""
        a, n, st = host_cloud(cloud)
        return self.L.ndtb_keyframe_put(self.h, int(kid), a.ctypes.data, n, st)

    def keyframe_drop(self, kid):
        return self.L.ndtb_keyframe_drop(self.h, int(kid))

    def keyframe_count(self):
        c = C.c_size_t(); t = C.c_size_t()
        self.L.ndtb_keyframe_count(self.h, C.byref(c), C.byref(t))
        return c.value, t.value

    def set_target_window(self, ids, poses):
        """Transform the listed keyframes by their poses (n x 4 x 4 float64), concatenate, rebuild the voxel map."""
        ids = np.ascontiguousarray(ids, np.int64)
        P = np.ascontiguousarray(poses, np.float64).reshape(len(ids), 16)
        return self.L.ndtb_set_target_window(self.h, len(ids), ids.ctypes.data_as(C.POINTER(C.c_int64)), _ptr(P, c_d))

    def voxelgrid_filter(self, cloud, leaf, dest="host"):
        """pcl::VoxelGrid down-sampling on the device.  dest: "host" -> returns the (m, 4) centroids, "source" / "target" ->
        they become that cloud (returns m)."""
        a, n, st = host_cloud(cloud)
        m = C.c_size_t()
        if dest == "host":
            out = np.zeros((n, 4), np.float32)
            self.check(self.L.ndtb_voxelgrid_filter(self.h, a.ctypes.data, n, st, float(leaf), 0, out.ctypes.data, n, C.byref(m)))
            return out[: m.value].copy()
        self.check(self.L.ndtb_voxelgrid_filter(self.h, a.ctypes.data, n, st, float(leaf), 1 if dest == "source" else 2, None, 0, C.byref(m)))
        return m.value

    def target_points(self):
        n = C.c_size_t()
        self.check(self.L.ndtb_get_target_points(self.h, None, 0, C.byref(n)))
        out = np.zeros((n.value, 4), np.float32)
        if n.value:
            self.check(self.L.ndtb_get_target_points(self.h, out.ctypes.data, n.value, C.byref(n)))
        return out

    def time_eval_kernel(self, p, T=None, reps=50):
        """Mean duration (us) of one evaluation kernel over `reps` back-to-back launches between one CUDA event pair."""
        p = np.ascontiguousarray(p, np.float64)
        t = None if T is None else np.ascontiguousarray(T, np.float32).reshape(16)
        us = np.zeros(1, np.float32)
        self.check(self.L.ndtb_time_eval_kernel(self.h, _ptr(p, c_d), _ptr(t, c_f), int(reps), _ptr(us, c_f)))
        return float(us[0])

    def set_option(self, name, value):
        """ndtb_set_option: "direct_solver", "sessions", "session_idle_timeout_ms", "session_free_sms", "device_newton", "batch_interleave"."""
        self.check(self.L.ndtb_set_option(self.h, name.encode(), float(value)))

    def stream_ptr(self):
        p = C.c_void_p()
        self.L.ndtb_get_stream(self.h, C.byref(p))
        return p.value

    # svn
    def svn_particle_derivatives(self, poses):
        P = np.ascontiguousarray(poses, np.float64); K = P.shape[0]
        s = np.zeros(K); g = np.zeros((K, 6)); H = np.zeros((K, 6, 6))
        self.check(self.L.ndtb_svn_particle_derivatives(self.h, K, _ptr(P, c_d), _ptr(s, c_d), _ptr(g, c_d), _ptr(H, c_d)))
        return s, g, H

    def svn_stage2(self, poses, grads, hess):
        P = np.ascontiguousarray(poses, np.float64); K = P.shape[0]
        g = np.ascontiguousarray(grads, np.float64); H = np.ascontiguousarray(hess, np.float64); u = np.zeros((K, 6))
        self.check(self.L.ndtb_svn_stage2(self.h, K, _ptr(P, c_d), _ptr(g, c_d), _ptr(H, c_d), _ptr(u, c_d)))
        return u

    def svn_step(self, poses):
        P = np.ascontiguousarray(poses, np.float64).copy(); K = P.shape[0]
        s = np.zeros(K); g = np.zeros((K, 6)); H = np.zeros((K, 6, 6)); u = np.zeros((K, 6))
        self.check(self.L.ndtb_svn_step(self.h, K, _ptr(P, c_d), _ptr(s, c_d), _ptr(g, c_d), _ptr(H, c_d), _ptr(u, c_d)))
        return P, s, g, H, u

    def svn_stage1_local(self, poses, rank, world):
        P = np.ascontiguousarray(poses, np.float64); K = P.shape[0]
        rec = np.zeros((K, 43))
        self.check(self.L.ndtb_svn_stage1_local(self.h, K, _ptr(P, c_d), rank, world, _ptr(rec, c_d)))
        return rec

    def svn_finish_step(self, poses, rec):
        P = np.ascontiguousarray(poses, np.float64).copy(); K = P.shape[0]
        rec = np.ascontiguousarray(rec, np.float64); u = np.zeros((K, 6))
        self.check(self.L.ndtb_svn_finish_step(self.h, K, _ptr(P, c_d), _ptr(rec, c_d), _ptr(u, c_d)))
        return P, u

    def svn_timings(self):
        e = np.zeros(1, np.float32); s = np.zeros(1, np.float32); w = np.zeros(1, np.float32); n = np.zeros(1, np.int32)
        self.L.ndtb_svn_get_timings(self.h, _ptr(e, c_f), _ptr(s, c_f), _ptr(w, c_f), _ptr(n, c_i))
        return dict(eval_ms=float(e[0]), stein_ms=float(s[0]), exchange_wait_ms=float(w[0]), iterations=int(n[0]))

    # multi-GPU SVN communicator (one process per GPU): window handle -> all-gather -> connect -> barrier
    def comm_window_handle(self):
        nb = C.c_size_t()
        self.check(self.L.ndtb_comm_window_handle(self.h, None, 0, C.byref(nb)))
        buf = np.zeros(nb.value, np.uint8)
        self.check(self.L.ndtb_comm_window_handle(self.h, buf.ctypes.data, nb.value, None))
        return buf

    def comm_connect(self, rank, world, all_handles):
        a = np.ascontiguousarray(all_handles, np.uint8)
        self.check(self.L.ndtb_comm_connect(self.h, int(rank), int(world), a.ctypes.data))

    def comm_close(self):
        self.check(self.L.ndtb_comm_close(self.h))

    def comm_info(self):
        r = np.zeros(1, np.int32); w = np.zeros(1, np.int32)
        self.L.ndtb_comm_info(self.h, _ptr(r, c_i), _ptr(w, c_i))
        return int(r[0]), int(w[0])

    def svn_align(self, prior, particles=None, seed=0):
        prior = np.ascontiguousarray(prior, np.float64).reshape(16)
        K = self._svn_kw.get("particle_count", 30) if hasattr(self, "_svn_kw") else 30
        init = None if particles is None else np.ascontiguousarray(particles, np.float64)
        out = np.zeros((K, 4, 4)); r = SvnResult()
        self.check(self.L.ndtb_svn_align(self.h, _ptr(prior, c_d), _ptr(init, c_d), seed, C.byref(r), _ptr(out, c_d)))
        return dict(final_pose=np.array(r.final_pose).reshape(4, 4), final_covariance=np.array(r.final_covariance).reshape(6, 6),
                    converged=bool(r.converged), iterations=r.iterations, particles=out)
