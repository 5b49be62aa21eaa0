"""Replay-batch throughput (BASELINE configs[3] in miniature): N scans against one 1M-point map, serial vs interleaved."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from slam_sam_b200 import ndtb, synth
scene = synth.Scene()
tgt = synth.surface_map(scene, 1_000_000, seed=7)
scans, guesses = [], []
for k in range(12):
    pose = synth.trajectory_pose(20.0 + 9.0 * k)
    scans.append(synth.ouster_scan(scene, pose, seed=300 + k))
    guesses.append((pose @ synth.pose_matrix((0.15, -0.1, 0.02), np.deg2rad((0.4, -0.3, 0.8)))).astype(np.float32))
import torch
pinned = [torch.from_numpy(sc).pin_memory() for sc in scans]          # page-locked host buffers: uploads are truly asynchronous
scans = [t.numpy() for t in pinned] * 4; guesses = np.stack(guesses * 4)
h = ndtb.Handle(0)
h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
assert h.set_target(tgt) == 0
for mode in (0, 1, 0, 1):
    h.set_option("batch_interleave", mode)
    t0 = time.perf_counter()
    r = h.align_batch(scans, guesses)
    dt = time.perf_counter() - t0
    ev = sum(x["evaluations"] for x in r)
    print(f"interleave={mode}: {len(scans) / dt:8.1f} scans/s, {1e6 * dt / ev:6.2f} us per evaluation, {ev / len(scans):.1f} evaluations per scan")
