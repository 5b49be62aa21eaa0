import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from slam_sam_b200 import ndtb, synth
tgt, scans, guesses, truths = bench.make_workload(0)
h = ndtb.Handle(0)
h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
n_kf = 5
kf_poses = np.stack([synth.trajectory_pose(30.0 + 2.0 * j) for j in range(n_kf)])
bounds = np.linspace(0, len(tgt), n_kf + 1).astype(int)
for j in range(n_kf):
    Tinv = np.linalg.inv(kf_poses[j])
    part = tgt[bounds[j]:bounds[j + 1], :3].astype(np.float64)
    body = np.zeros((len(part), 4), np.float32)
    body[:, :3] = (part @ Tinv[:3, :3].T + Tinv[:3, 3]).astype(np.float32)
    assert h.keyframe_put(j, body) == 0
print("poses det", [np.linalg.det(P[:3, :3]) for P in kf_poses])
assert h.set_target_window(np.arange(n_kf), kf_poses) == 0
w = h.target_points()
print("window vs map max abs diff", np.abs(w[:, :3] - tgt[:, :3]).max(), "n", len(w), len(tgt))
vw = h.voxels()
h2 = ndtb.Handle(0); h2.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
h2.set_target(tgt); v = h2.voxels()
print("voxels", len(vw["idx"]), len(v["idx"]), "same idx set", np.array_equal(vw["idx"], v["idx"]))
for k in range(len(scans)):
    h.set_source(scans[k]); h2.set_source(scans[k])
    r1 = h.align(guesses[k], capacity=0); r2 = h2.align(guesses[k], capacity=0)
    print(k, "window: it", r1["iterations"], "err", np.abs(r1["T"][:3, 3] - truths[k][:3, 3]).max(), "| map: it", r2["iterations"], "err", np.abs(r2["T"][:3, 3] - truths[k][:3, 3]).max())
