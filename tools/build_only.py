"""Builds the bench map a few times (nothing else): the command ncu profiles the build kernels on."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from slam_sam_b200 import ndtb, synth
scene = synth.Scene()
tgt = synth.surface_map(scene, bench.N_MAP, seed=7)
h = ndtb.Handle(0)
h.set_params(resolution=1.0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
for i in range(n):
    t0 = time.perf_counter()
    assert h.set_target(tgt) == 0
    dt = time.perf_counter() - t0
    print(i, "set_target wall ms", dt * 1e3, "build_ms", h.timings()["build_ms"], flush=True)
print("V", len(h.voxels()["idx"]))
