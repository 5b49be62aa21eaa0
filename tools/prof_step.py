"""Profiling target: ONE bench step (1M-pt map build + 131072-pt DIRECT7 align) after one warm-up step.
Run plain first, then under ncu (B200_PROFILING.md recipe)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from slam_sam_b200 import ndtb, synth

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
scene = synth.Scene()
tgt = synth.surface_map(scene, 1_000_000, seed=7)
pose = synth.trajectory_pose(37.0)
scan = synth.ouster_scan(scene, pose, seed=100)
guess = (pose @ synth.pose_matrix((0.15, -0.1, 0.02), np.deg2rad((0.4, -0.3, 0.8)))).astype(np.float32)
h = ndtb.Handle(0)
h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
import os as _os
h.set_profiling(_os.environ.get("NDTB_EVENTS", "1") == "1")
for rep in range(reps):
    assert h.set_target(tgt) == 0 and h.set_source(scan) == 0
    r = h.align(guess, capacity=0)
    print(rep, h.timings(), r["iterations"], r["converged"], "hits", h.last_hits(), "launches", h.launch_count(),
          "terr", float(np.abs(r["T"][:3, 3] - pose[:3, 3]).max()))
