"""A/B of the Newton-loop plumbing on the bench workload: resident session kernel vs one launch per evaluation,
pivoted elimination vs Jacobi SVD for the Newton direction.  Prints wall time per align and the largest relative
difference of the per-evaluation trace against the (launch, svd) baseline."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from slam_sam_b200 import ndtb, synth

scene = synth.Scene()
tgt = synth.surface_map(scene, 1_000_000, seed=7)
h = ndtb.Handle(0)
h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
assert h.set_target(tgt) == 0
cases = []
for k in range(6):
    pose = synth.trajectory_pose(20.0 + 7.0 * k)
    scan = synth.ouster_scan(scene, pose, seed=100 + k)
    guess = (pose @ synth.pose_matrix((0.15, -0.1, 0.02), np.deg2rad((0.4, -0.3, 0.8)))).astype(np.float32)
    cases.append((scan, guess))
base = None
for sessions, direct in ((0, 0), (0, 1), (1, 0), (1, 1)):
    h.set_option("sessions", sessions); h.set_option("direct_solver", direct)
    traces, evals, t_al, t_dev = [], 0, 0.0, 0.0
    for rep in range(4):
        for scan, guess in cases:
            h.set_source(scan)
            t0 = time.perf_counter()
            r = h.align(guess, capacity=0)
            t1 = time.perf_counter()
            if rep:
                t_al += t1 - t0; evals += r["evaluations"]; t_dev += h.timings()["eval_ms_total"]
            if rep == 3:
                traces.append(h.trace())
    if base is None:
        base = traces
    worst = 0.0
    for a, b in zip(base, traces):
        if a.shape != b.shape:
            worst = float("inf"); continue
        den = np.maximum(np.abs(a), 1e-300)
        sc = np.abs(a[:, 6:7])  # score column scales g/H comparisons poorly; use per-row max instead
        rowmax = np.abs(a).max(axis=1, keepdims=True)
        worst = max(worst, float((np.abs(a - b) / rowmax).max()))
    print(f"sessions={sessions} direct_solver={direct}: {1e6 * t_al / evals:7.2f} us wall per evaluation, {1e3 * t_al / (3 * len(cases)):.3f} ms per align, "
          f"{evals / (3 * len(cases)):.1f} evals/align, device eval {1e3 * t_dev / max(evals, 1):.2f} us, trace diff vs baseline {worst:.2e}")
