"""Prints a hash of the per-evaluation traces of a few aligns (bit-level regression check between two builds:
NDTB_LIB=... python tools/trace_hash.py)."""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from slam_sam_b200 import ndtb, synth
scene = synth.Scene()
tgt = synth.surface_map(scene, 300_000, seed=7)
h = ndtb.Handle(0)
hs = hashlib.sha256()
for search in (ndtb.DIRECT7, ndtb.DIRECT1):
    h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=search)
    assert h.set_target(tgt) == 0
    for k in range(3):
        pose = synth.trajectory_pose(20.0 + 7.0 * k)
        scan = synth.ouster_scan(scene, pose, beams=64, cols=512, seed=100 + k)
        guess = (pose @ synth.pose_matrix((0.15, -0.1, 0.02), np.deg2rad((0.4, -0.3, 0.8)))).astype(np.float32)
        h.set_source(scan)
        r = h.align(guess, capacity=0)
        hs.update(h.trace().tobytes()); hs.update(r["T"].tobytes())
        s, g, H, nv = h.compute_derivatives(np.array([0.1, 0.2, 0.0, 0.01, -0.02, 0.3]), compute_hessian=False)
        hs.update(np.float64(s).tobytes()); hs.update(g.tobytes())
print("trace hash", hs.hexdigest()[:16], "evals", r["evaluations"])
