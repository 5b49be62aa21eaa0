"""Profiling target: the eval kernel alone on an 8x tiled scan (steady state, no launch/tail effects)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from slam_sam_b200 import ndtb, synth
mult = int(sys.argv[1]) if len(sys.argv) > 1 else 8
scene = synth.Scene()
tgt = synth.surface_map(scene, 1_000_000, seed=7)
pose = synth.trajectory_pose(37.0)
scan = synth.ouster_scan(scene, pose, seed=100)
h = ndtb.Handle(0)
h.set_params(resolution=1.0, search_method=ndtb.DIRECT7)
assert h.set_target(tgt) == 0 and h.set_source(np.concatenate([scan] * mult)) == 0
p6 = np.array([pose[0, 3], pose[1, 3], pose[2, 3], 0, 0, 0.0])
for _ in range(4):
    s, g, H, nv = h.compute_derivatives(p6, T=pose.astype(np.float32))
print(s, h.timings())
