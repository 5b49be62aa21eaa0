"""Development aid: SVN-NDT iteration time, K particles x 131072-pt scan (configs[2])."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from slam_sam_b200 import ndtb, synth
K = int(sys.argv[1]) if len(sys.argv) > 1 else 100
scene = synth.Scene()
A = synth.trajectory_pose(40.0); B = A @ synth.pose_matrix((0.5, 0.1, 0.02), np.deg2rad((0.5, 0.3, 2.0)))
tgt = synth.ouster_scan(scene, A, seed=1); src = synth.ouster_scan(scene, B, seed=2)
h = ndtb.Handle(0)
h.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=1.0, min_points_per_voxel=3)
h.set_svn_params(particle_count=K, max_iterations=50, kernel_bandwidth=1.0, step_size=1.0, stop_threshold=1e-4, use_gauss_newton=1)
assert h.set_target(tgt) == 0 and h.set_source(src) == 0
rng = np.random.default_rng(7)
prior = np.eye(4)
P = np.stack([prior @ synth.pose_matrix(rng.normal(0, 0.05, 3), rng.normal(0, 0.01, 3)) for _ in range(K)])
for it in range(3):
    t = time.perf_counter(); P, s, g, H, u = h.svn_step(P); dt = time.perf_counter() - t
    print("iter", it, "wall ms", dt * 1e3, h.timings(), "mean |u|", np.abs(u).mean())
t = time.perf_counter(); r = h.svn_align(prior, particles=P, seed=0); dt = time.perf_counter() - t
print("svn_align wall ms", dt * 1e3, "iterations", r["iterations"], "converged", r["converged"], h.timings())
print("rel pose error vs truth", np.abs(r["final_pose"] - np.linalg.inv(A) @ B).max())
