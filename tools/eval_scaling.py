"""Development aid: eval-kernel time vs source size (separates the fixed launch/tail cost from the per-point cost)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from slam_sam_b200 import ndtb, synth
scene = synth.Scene()
tgt = synth.surface_map(scene, 1_000_000, seed=7)
pose = synth.trajectory_pose(37.0)
scan = synth.ouster_scan(scene, pose, seed=100)
h = ndtb.Handle(0)
h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
assert h.set_target(tgt) == 0
p6 = np.array([pose[0, 3], pose[1, 3], pose[2, 3], 0, 0, 0.0])
for mult in (0.125, 0.25, 0.5, 1, 2, 4, 8):
    if mult < 1:
        src = scan[: int(len(scan) * mult)]
    else:
        src = np.concatenate([scan] * int(mult))
    assert h.set_source(src) == 0
    for _ in range(3):
        h.compute_derivatives(p6, T=pose.astype(np.float32))
    t0 = h.timings()["eval_ms_total"]; reps = 20
    for _ in range(reps):
        s, g, H, nv = h.compute_derivatives(p6, T=pose.astype(np.float32))
    t1 = h.timings()["eval_ms_total"]
    print(f"n={len(src):8d} eval_us={(t1 - t0) / reps * 1e3:8.2f} hits={h.last_hits()[0]:.3f} score={s:.6e}")
