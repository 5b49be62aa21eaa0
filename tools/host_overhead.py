"""Wall clock of the C-ABI calls of one bench step against the CUDA-event times they report (where does the rest go?)."""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from slam_sam_b200 import ndtb, synth
scene = synth.Scene(); tgt = synth.surface_map(scene, 1_000_000, seed=7); pose = synth.trajectory_pose(37.0)
scan = synth.ouster_scan(scene, pose, seed=100)
guess = np.ascontiguousarray((pose @ synth.pose_matrix((0.15, -0.1, 0.02), np.deg2rad((0.4, -0.3, 0.8)))).astype(np.float32)).reshape(16)
h = ndtb.Handle(0); h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
d_t = torch.from_numpy(tgt).cuda(); d_s = torch.from_numpy(scan).cuda(); torch.cuda.synchronize()
res = ndtb.Result(); gp = guess.ctypes.data_as(ndtb.c_f)
T = {"set_target": [], "set_source": [], "align": [], "build_ms": [], "align_ms": [], "eval_ms": [], "evals": []}
for i in range(30):
    t0 = time.perf_counter_ns(); h.L.ndtb_set_target_device(h.h, ctypes.c_void_p(d_t.data_ptr()), 1_000_000)
    t1 = time.perf_counter_ns(); h.L.ndtb_set_source_device(h.h, ctypes.c_void_p(d_s.data_ptr()), scan.shape[0])
    t2 = time.perf_counter_ns(); h.L.ndtb_align(h.h, gp, ctypes.byref(res), None, 0, None)
    t3 = time.perf_counter_ns()
    tm = h.timings()
    if i >= 5:
        T["set_target"].append((t1 - t0) / 1e3); T["set_source"].append((t2 - t1) / 1e3); T["align"].append((t3 - t2) / 1e3)
        T["build_ms"].append(tm["build_ms"] * 1e3); T["align_ms"].append(tm["align_ms"] * 1e3); T["eval_ms"].append(tm["eval_ms_total"] * 1e3); T["evals"].append(res.evaluations)
for k, v in T.items():
    print(f"{k:12s} mean {np.mean(v):9.2f}  min {np.min(v):9.2f}  (us)")
