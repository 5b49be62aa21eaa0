"""Wall time of ndtb_set_target for the strides a caller can hand over: packed float4 (pcl::PointXYZ, 16 B), pcl::PointXYZI (32 B), packed xyz (12 B); pinned and pageable."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from slam_sam_b200 import ndtb, synth
scene = synth.Scene()
tgt = synth.surface_map(scene, bench.N_MAP, seed=7)
h = ndtb.Handle(0); h.set_params(resolution=1.0)
for width in (4, 8, 3):
    c = np.zeros((tgt.shape[0], width), np.float32); c[:, :3] = tgt[:, :3]
    for pinned in (True, False):
        t = torch.from_numpy(c).pin_memory() if pinned else torch.from_numpy(c)
        ts = []
        for i in range(6):
            t0 = time.perf_counter()
            rc = h.L.ndtb_set_target(h.h, t.data_ptr(), c.shape[0], width * 4)
            ts.append(time.perf_counter() - t0)
            assert rc == 0
        print("stride", width * 4, "pinned" if pinned else "pageable", "set_target wall ms", round(min(ts[1:]) * 1e3, 3), "build_ms", round(h.timings()["build_ms"], 3), flush=True)
