"""Development aid: build + align on the synthetic Ouster pair, twice, printing timings."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from slam_sam_b200 import synth, ndtb
scene = synth.Scene()
A = synth.trajectory_pose(40.0); B = A @ synth.pose_matrix((0.5, 0.1, 0.02), np.deg2rad((0.5, 0.3, 2.0)))
tgt = synth.ouster_scan(scene, A, seed=1); src = synth.ouster_scan(scene, B, seed=2)
h = ndtb.Handle(0)
h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01)
for rep in range(3):
    t = time.time(); rc = h.set_target(tgt); t1 = time.time() - t
    h.set_source(src)
    t = time.time(); r = h.align(); t2 = time.time() - t
    print(rep, "set_target wall", t1, "align wall", t2, h.timings(), r['iterations'])
v = h.voxels(); print("V", len(v['n']), "max n", v['n'].max(), "n>1000:", (v['n'] > 1000).sum(), "valid", (v['n'] >= 6).sum())
