"""Small end-to-end run of every kernel family (compute-sanitizer memcheck target): map build (1-tile and multi-tile sort paths), align with and
without a session, KDTREE / DIRECT26, line search, svn align, staged uploads, window, voxelgrid, multigrid, lidar (RNG19 + LEGACY)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from slam_sam_b200 import ndtb, synth

scene = synth.Scene()
A = synth.trajectory_pose(40.0)
B = A @ synth.pose_matrix((0.5, 0.1, 0.02), np.deg2rad((0.5, 0.3, 2.0)))
tgt = synth.ouster_scan(scene, A, beams=32, cols=512, seed=1)
src = synth.ouster_scan(scene, B, beams=32, cols=512, seed=2)
big = synth.surface_map(scene, 1_300_000, seed=7)          # > 148 tiles: the multi-tile path of the sort kernel
for search in (ndtb.DIRECT7, ndtb.DIRECT1, ndtb.KDTREE, ndtb.DIRECT26):
    h = ndtb.Handle(0)
    h.set_params(resolution=1.0, search_method=search)
    assert h.set_target(tgt) == 0 and h.set_source(src) == 0
    r = h.align(); print("align", search, r["iterations"], r["converged"])
    h.set_option("sessions", 0); h.align()
    if search in (ndtb.DIRECT7, ndtb.DIRECT1):
        print("static", h.compute_derivatives_static(np.zeros(6))[0])
h = ndtb.Handle(0); h.set_params(resolution=1.0)
assert h.set_target(big) == 0 and h.set_source(src) == 0
print("big map leaves", len(h.voxels()["idx"]), h.align()["iterations"])
assert h.stage_target(tgt) == 0 and h.stage_source(src) == 0 and h.commit_target() == 0 and h.commit_source() == 0
print("staged", h.align()["iterations"])
h2 = ndtb.Handle(0); h2.set_params(resolution=1.0, use_line_search=1, step_size=0.5)
h2.set_target(tgt); h2.set_source(src); print("line search", h2.align()["iterations"])
hs = ndtb.Handle(0)
hs.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=1.0, min_points_per_voxel=3, search_method=ndtb.DIRECT7)
hs.set_svn_params(particle_count=12, max_iterations=6, kernel_bandwidth=1.0, step_size=1.0, stop_threshold=1e-4, use_gauss_newton=1)
hs.set_target(tgt); hs.set_source(src)
rng = np.random.Generator(np.random.PCG64(7))
P0 = np.stack([synth.pose_matrix(rng.normal(0, 0.05, 3), rng.normal(0, 0.01, 3)) for _ in range(12)])
print("svn", hs.svn_align(np.eye(4), P0)["iterations"])
print("voxelgrid", h.voxelgrid_filter(tgt, 0.5).shape if hasattr(h, "voxelgrid_filter") else "n/a")
import test_lidar_decode as T
from slam_sam_b200.lidar import LidarCallback
for legacy in (0, 1):
    cfg, az, alt = T.make_config(stride=2, udp_profile_legacy=legacy)
    pk = T.make_legacy_stream(cfg, [64, 40, 64], seed=3) if legacy else T.make_stream(cfg, 2, seed=3)
    print("lidar", legacy, len(LidarCallback(cfg, az, alt).DecodePackets(pk)))
print("ok")
