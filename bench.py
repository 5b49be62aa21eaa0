#!/usr/bin/env python
"""bench.py -- NDT align scans/s at 131 k points (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload map_align|svn|replay|multigrid]

Workload `map_align` (default; BASELINE.json configs[1]): one step = target voxel-map build of a
1 M-point synthetic map + full NDT align (DIRECT7, 1.0 m voxels) of one 131 072-point OS1-128-shaped
scan against it -- the reference rebuilds the target every keyframe (run/pipeline_ligo_tc.cpp:529-532).
  value : scans/s with map and scan already resident in HBM when the timed region starts
  e2e   : scans/s through the C ABI with HOST (pinned) buffers: H2D of map + scan and D2H of the
          result inside the timed region -- the number to hold against the reference arm
N > 1 (torchrun, one rank per GPU): independent replicas, scan streams sharded by rank, no data-path
collective ("weak" scaling); time = max over ranks.
`--impl reference` times the CPU restatement of the reference path (oracle/, all host threads); the
reference itself needs PCL/Eigen/GTSAM and cannot be built in this image.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_SCAN = 131072
N_MAP = 1_000_000
RESOLUTION = 1.0
N_SCANS_POOL = 6


def make_workload(rank):
    from slam_sam_b200 import synth

    scene = synth.Scene()
    tgt = synth.surface_map(scene, N_MAP, seed=7)
    scans, guesses, truths = [], [], []
    rng = np.random.Generator(np.random.PCG64(1000 + rank))
    for i in range(N_SCANS_POOL):
        s = 20.0 + 17.0 * i + 3.0 * rank
        pose = synth.trajectory_pose(s)
        scans.append(synth.ouster_scan(scene, pose, seed=100 * rank + i))
        # guess = truth perturbed by N(0, 0.2 m / 1 deg), clipped to stay in the basin
        dt = np.clip(rng.normal(0, 0.2, 3), -0.3, 0.3) * [1, 1, 0.2]
        dr = np.deg2rad(np.clip(rng.normal(0, 1.0, 3), -1.5, 1.5))
        guesses.append((pose @ synth.pose_matrix(dt, dr)).astype(np.float32))
        truths.append(pose)
    return tgt, scans, guesses, truths


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), samples=len(sm), reasons=sorted(reasons))
        return out


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def build_traffic():
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "build_kernels_traffic.json")))["dram_bytes_per_build"])
    except Exception:
        return None


def ncu_inst_executed():
    """Warp instructions one eval-kernel launch executes (ncu smsp__inst_executed.sum, profiles/eval_kernel_traffic.json)."""
    path = os.path.join(ROOT, "profiles", "eval_kernel_traffic.json")
    try:
        return float(json.load(open(path))["inst_executed_per_launch"])
    except Exception:
        return None


def ncu_traffic():
    path = os.path.join(ROOT, "profiles", "eval_kernel_traffic.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["dram_bytes_per_launch"])
        except Exception:
            return None
    return None


def run_cpu(tgt, scans, guesses, steps, warmup, threads):
    """The CPU restatement of the reference path, all host threads: build + align per step.  The per-step results (pose,
    iteration count, per-evaluation trace) are kept: bench's `parity` block holds the GPU path against them."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O

    O.build()
    o = O.OracleNdt(resolution=RESOLUTION, max_iterations=35, trans_epsilon=0.01, step_size=0.1, num_threads=threads)
    times, iters, build_t, results = [], [], [], {}
    for i in range(warmup + steps):
        k = i % len(scans)
        t0 = time.perf_counter()
        assert o.set_target(tgt) == 0
        t1 = time.perf_counter()
        o.set_source(scans[k])
        r = o.align(guesses[k])
        t2 = time.perf_counter()
        results[k] = r
        if i >= warmup:
            times.append(t2 - t0); build_t.append(t1 - t0); iters.append(r["iterations"])
    return dict(total_s=float(np.sum(times)), mean_s=float(np.mean(times)), build_s=float(np.mean(build_t)), iterations=float(np.mean(iters)),
                results=results)


def pose_delta(Ta, Tb):
    d = np.linalg.inv(np.asarray(Ta, float)) @ np.asarray(Tb, float)
    return float(np.linalg.norm(d[:3, 3])), float(np.arccos(np.clip((np.trace(d[:3, :3]) - 1) / 2, -1, 1)))


def parity_block(h, tgt, scans, guesses, cpu_results):
    """The bench line carries its own parity: the scans the CPU leg aligned are aligned again on the GPU (untimed) and every
    evaluation of every Newton loop is held against the oracle's.  Bars of BASELINE.json north_star; a miss fails the run."""
    assert h.set_target(tgt) == 0
    it_equal, dt_max, dr_max, tr_max, n_eval = True, 0.0, 0.0, 0.0, 0
    for k, ro in sorted(cpu_results.items()):
        assert h.set_source(scans[k]) == 0
        rg = h.align(guesses[k])
        tg, to = h.trace(), ro["trace"]
        it_equal = it_equal and rg["iterations"] == ro["iterations"] and rg["converged"] == ro["converged"] and tg.shape == to.shape
        if tg.shape == to.shape:
            for a, b in zip(tg, to):
                tr_max = max(tr_max, abs(a[6] - b[6]) / abs(b[6]), float(np.abs(a[7:13] - b[7:13]).max() / np.abs(b[7:13]).max()),
                             float(np.abs(a[13:49] - b[13:49]).max() / np.abs(b[13:49]).max()))
                n_eval += 1
        dt, dr = pose_delta(rg["T"], ro["T"])
        dt_max, dr_max = max(dt_max, dt), max(dr_max, dr)
    out = {"against": "oracle (CPU restatement of ndt_omp), same scans as cpu_baseline", "scans": len(cpu_results), "evaluations_compared": n_eval,
           "iterations_equal": bool(it_equal), "max_dt_m": dt_max, "max_drot_rad": dr_max, "trace_rel_inf": tr_max,
           "bars": {"trace_rel_inf": 1e-9, "max_dt_m": 1e-4, "max_drot_rad": 1e-5}}
    out["ok"] = bool(it_equal and tr_max <= 1e-9 and dt_max <= 1e-4 and dr_max <= 1e-5)
    return out


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from slam_sam_b200 import distributed as D
    from slam_sam_b200 import ndtb, synth

    tgt, scans, guesses, truths = make_workload(rank)
    h = ndtb.Handle(local_rank)  # no CPU fallback: raises if the CUDA library / device is missing
    h.set_params(resolution=RESOLUTION, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
    if os.environ.get("NDTB_BENCH_DEVICE_NEWTON") == "1":   # A/B aid: the resident kernel takes the Newton steps (opt-in)
        h.set_option("device_newton", 1)
    stream = torch.cuda.ExternalStream(h.stream_ptr(), device=local_rank)
    # device-resident copies (value) and pinned host copies (e2e)
    d_tgt = torch.from_numpy(tgt).cuda()
    d_scans = [torch.from_numpy(s).cuda() for s in scans]
    p_tgt = torch.from_numpy(tgt).pin_memory()
    p_scans = [torch.from_numpy(s).pin_memory() for s in scans]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    torch.cuda.synchronize()

    # sliding-window variant of the end-to-end step: the map window lives in HBM as 5 archived keyframe clouds (uploaded
    # once, outside the timed region, like pointsArchive of run/pipeline_ligo_tc.cpp:519-526); per step the window is
    # re-assembled on the device with the keyframes' poses, and only the new scan crosses PCIe
    n_kf = 5
    # identity poses: the assembled window is then bit-identical to the 1M-pt map of the other regions (same voxels, same
    # Newton trajectories -- a non-trivial pose moves points by a float ulp, which this un-line-searched Newton loop
    # amplifies into different iteration counts); the kernel does the same arithmetic for any pose
    kf_poses = np.stack([np.eye(4) for j in range(n_kf)])
    kf_ids = np.arange(n_kf, dtype=np.int64)
    bounds = np.linspace(0, N_MAP, n_kf + 1).astype(int)
    for j in range(n_kf):
        Tinv = np.linalg.inv(kf_poses[j])
        part = tgt[bounds[j]:bounds[j + 1], :3].astype(np.float64)
        body = np.zeros((len(part), 4), np.float32)
        body[:, :3] = (part @ Tinv[:3, :3].T + Tinv[:3, 3]).astype(np.float32)
        assert h.keyframe_put(int(kf_ids[j]), body) == 0

    # scan-producer variant: the scan arrives as the sensor sends it -- 64 RNG19 lidar packets of 24 832 bytes
    # (one 1024 x 128 revolution, pinned host memory) plus the first packet of the next revolution, which closes the frame
    # (src/lidarcallback.cpp:658-669); packets are decoded on the device and the frame becomes the NDT source without leaving HBM
    from slam_sam_b200 import lidar as lidar_mod, synth as synth_mod
    lcfg = ndtb.LidarConfig()
    ndtb.lib().ndtb_lidar_default_config(ctypes.byref(lcfg))
    lcfg.columns_per_frame, lcfg.pixels_per_column, lcfg.columns_per_packet, lcfg.channel_stride = 1024, 128, 16, 1
    lcfg.range_min, lcfg.range_max, lcfg.z_min, lcfg.z_max = 0.3, 250.0, -300.0, 300.0
    scene_pk = synth_mod.Scene()
    p_packets, lc = [], None
    for i in range(len(scans)):
        rm, az_deg, alt_deg, Cb2s, tb2s = synth_mod.ouster_range_image(scene_pk, truths[i], seed=100 * rank + i)
        if lc is None:
            lcfg.body_to_lidar_rotation[:] = list(Cb2s); lcfg.body_to_lidar_translation[:] = list(tb2s)
            lc = lidar_mod.LidarCallback(lcfg, az_deg, alt_deg, device=local_rank)
        ts_ns = np.uint64(1_762_300_000_000_000_000) + np.arange(1024, dtype=np.uint64) * np.uint64(97_656)
        zeros = np.zeros_like(rm)
        pk = lidar_mod.encode_rng19_packets(rm, zeros, zeros, zeros, ts_ns, np.zeros(1024), 16)
        p_packets.append(torch.from_numpy(np.concatenate([pk, pk[:1]])).pin_memory())  # 64 packets + the closing one
    n_pk = p_packets[0].shape[0] - 1
    pk_bytes = p_packets[0].shape[1]
    pk_state = {"fid": 1, "points": 0}

    # the timed region covers the C-ABI calls a C++ caller makes; arguments are marshalled once, results are unpacked after the
    # closing event (the Python wrappers' numpy / dict conversions are binding overhead, not part of the path)
    g16 = [np.ascontiguousarray(g, np.float32).reshape(16) for g in guesses]
    g16_ptr = [g.ctypes.data_as(ndtb.c_f) for g in g16]
    res_struct = ndtb.Result()
    kf_ids64 = np.ascontiguousarray(kf_ids, np.int64)
    kf_ids_ptr = kf_ids64.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))
    kf_P = np.ascontiguousarray(kf_poses, np.float64).reshape(n_kf, 16)
    kf_P_ptr = kf_P.ctypes.data_as(ndtb.c_d)

    pipe_state = {"primed": False}
    tgt_ptr = p_tgt.data_ptr()
    scan_ptr = [t.data_ptr() for t in p_scans]

    def unpack(r):
        return dict(T=np.array(r.pose, np.float32).reshape(4, 4), iterations=r.iteration_num, converged=bool(r.converged), evaluations=r.evaluations)

    def step(i, resident):
        k = i % len(scans)
        if resident == "packets":
            buf = p_packets[k].numpy()
            fid = pk_state["fid"]; pk_state["fid"] = (fid + 2) & 0xffff
            buf[:n_pk, 2] = fid & 0xff; buf[:n_pk, 3] = fid >> 8
            buf[n_pk, 2] = (fid + 1) & 0xff; buf[n_pk, 3] = ((fid + 1) >> 8) & 0xff
            lc.push(p_packets[k].data_ptr(), n_pk + 1, pk_bytes)           # DMA + decode kernel queued on the producer's stream
            rc = h.L.ndtb_set_target_window(h.h, n_kf, kf_ids_ptr, kf_P_ptr)  # meanwhile: window assembly + map build
            while True:
                f = lc.pop()
                assert f is not None
                if f.packets == n_pk:                                        # (the one-packet frame of the previous closer is skipped)
                    break
            ptr, npts = f.device_points()
            pk_state["points"] = npts
            rc |= h.set_source_device(ptr, npts)
        elif resident == "window":
            rc = h.L.ndtb_set_target_window(h.h, n_kf, kf_ids_ptr, kf_P_ptr)
            rc |= h.L.ndtb_set_source(h.h, p_scans[k].data_ptr(), N_SCAN, 16)
        elif resident is True:
            rc = h.set_target_device(d_tgt.data_ptr(), N_MAP)
            rc |= h.set_source_device(d_scans[k].data_ptr(), N_SCAN)
        elif resident == "serial":
            rc = h.L.ndtb_set_target(h.h, p_tgt.data_ptr(), N_MAP, 16)
            rc |= h.L.ndtb_set_source(h.h, p_scans[k].data_ptr(), N_SCAN, 16)
        else:
            # pipelined uploads (ndtb_stage_* / ndtb_commit_*): the copy of step i + 1's map and scan is started first and runs on
            # the copy stream while step i's map is built and aligned; the step ends once that copy has arrived, so every timed
            # step contains one full 18 MB host-to-device copy
            kn = (i + 1) % len(scans)
            if not pipe_state["primed"]:                                     # first step of a region: nothing was staged for it
                rc = h.L.ndtb_stage_target(h.h, tgt_ptr, N_MAP, 16) | h.L.ndtb_stage_source(h.h, scan_ptr[k], N_SCAN, 16)
                assert rc == 0, h.L.ndtb_last_error(h.h)
                pipe_state["primed"] = True
            rc = h.L.ndtb_stage_target(h.h, tgt_ptr, N_MAP, 16)
            rc |= h.L.ndtb_stage_source(h.h, scan_ptr[kn], N_SCAN, 16)
            rc |= h.L.ndtb_commit_target(h.h)
            rc |= h.L.ndtb_commit_source(h.h)
        assert rc == 0, h.L.ndtb_last_error(h.h)
        rc = h.L.ndtb_align(h.h, g16_ptr[k], ctypes.byref(res_struct), None, 0, None)
        assert rc == 0, h.L.ndtb_last_error(h.h)
        if resident is False:
            assert h.L.ndtb_stage_sync(h.h) == 0
        return res_struct

    def timed(resident, steps, warmup):
        for i in range(warmup):
            step(i, resident)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        launches0 = h.launch_count()
        ev, stats = [], dict(build_ms=0.0, align_ms=0.0, eval_ms=0.0, eval_launches=0, iters=0, hits=0.0, err=0.0)
        wall0 = time.perf_counter()
        for i in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()          # L2 flush between timed iterations, outside the event pair
            stream.synchronize()       # the flush has finished before the opening event: no host work of the step hides behind it
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            r = step(warmup + i, resident)
            e1.record(stream)
            ev.append((e0, e1))
            r = unpack(r)
            t = h.timings()
            stats["build_ms"] += t["build_ms"]; stats["align_ms"] += t["align_ms"]; stats["eval_ms"] += t["eval_ms_total"]
            stats["eval_launches"] += t["eval_launches"]; stats["iters"] += r["iterations"]; stats["hits"] += h.last_hits()[0]
            k = (warmup + i) % len(scans)
            stats["err"] = max(stats["err"], float(np.abs(r["T"][:3, 3] - truths[k][:3, 3]).max()))
            assert r["converged"], "alignment did not converge"
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        wall = time.perf_counter() - wall0
        total_ms = sum(a.elapsed_time(b) for a, b in ev)
        stats["launches"] = h.launch_count() - launches0
        stats["wall_s"] = wall
        return total_ms, stats

    clocks = ClockSampler(local_rank)
    clocks.start()
    # value / e2e regions: the library's default path (one resident session kernel per align, no per-kernel events);
    # a third region of the same K resident steps runs one launch per evaluation with a CUDA event pair around each and
    # yields the eval kernel's launch duration (roofline).
    h.set_profiling(False)
    total_ms, st = timed(True, args.steps, args.warmup)
    e2e_ms, st2 = timed(False, args.steps, args.warmup)
    while h.staged_count() != (0, 0):                                        # drain what the last pipelined step staged
        h.commit_target(); h.commit_source()
    pipe_state["primed"] = False
    ser_ms, st6 = timed("serial", args.steps, args.warmup)
    win_ms, st4 = timed("window", args.steps, args.warmup)
    pkt_ms, st5 = timed("packets", args.steps, args.warmup)
    decode_launches, decode_us = lc.stats()
    h.set_profiling(True)
    roof_ms, st3 = timed(True, args.steps, 1)
    # the same kernel timed as 50 back-to-back launches between ONE event pair (event overhead amortised), scan 0 at its guess
    h.set_source_device(d_scans[0].data_ptr(), N_SCAN)
    g0 = guesses[0].astype(np.float64)
    p0 = np.array([g0[0, 3], g0[1, 3], g0[2, 3], 0.0, 0.0, 0.0])
    b2b_us = h.time_eval_kernel(p0, guesses[0], reps=50)
    clk = clocks.stop()
    total_ms = D.max_over_ranks(total_ms, local_rank)
    e2e_ms = D.max_over_ranks(e2e_ms, local_rank)
    ser_ms = D.max_over_ranks(ser_ms, local_rank)
    win_ms = D.max_over_ranks(win_ms, local_rank)
    pkt_ms = D.max_over_ranks(pkt_ms, local_rank)
    launches = int(D.sum_over_ranks(st["launches"], local_rank))
    if world > 1:
        dist.barrier()
    if rank != 0:
        return None
    K = args.steps
    value = world * K / (total_ms * 1e-3)
    e2e = world * K / (e2e_ms * 1e-3)
    hbar = st["hits"] / K
    b_eval = N_SCAN * (16 + 8 * 7 + 64 * hbar)                 # algorithmic bytes per evaluation
    t_eval = st3["eval_ms"] / max(st3["eval_launches"], 1) * 1e-3  # CUDA event pairs on the library's stream, third timed region
    peak, peak_src = measured_peak()
    achieved = b_eval / t_eval / 1e9
    line = {
        "metric": "NDT align scans/sec @131k pts (1/2/4/8 B200) + % gather-BW roofline vs CPU OMP", "value": value, "unit": "scans/s",
        "n_gpus": world, "steps": K, "warmup": args.warmup, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 pair maths, f64 accumulate", "data": "synthetic",
        "config": {"workload": "configs[1]: voxel-map build of a 1M-pt synthetic map + DIRECT7 NDT align of one 131072-pt OS1-128-shaped scan, "
                               "1.0 m voxels, per step", "n_source": N_SCAN, "n_target": N_MAP, "resolution": RESOLUTION, "max_iterations": 35,
                   "trans_epsilon": 0.01, "l2": "256 MB flush between timed steps", "parallelism": f"replicas x{world} (no collective)",
                   "build_ms": st["build_ms"] / K, "align_ms": st["align_ms"] / K, "newton_iterations": st["iters"] / K,
                   "evaluations_per_align": st["eval_launches"] / K, "max_abs_translation_error_m": st["err"],
                   "eval_kernel_share_of_step": st3["eval_ms"] / roof_ms,
                   "timing": "value/e2e regions: every align runs its evaluations in ONE resident kernel (commands through a pinned mailbox), "
                             "roofline.session_eval_us is its device-side (globaltimer) time per evaluation, command seen -> results stored; "
                             "roofline.kernel_us / achieved / frac come from a third region of the same steps with one launch per evaluation "
                             "and a CUDA event pair around each (an empty launch between two events reads ~7 us on this box); "
                             "roofline.kernel_us_back_to_back is the same kernel as 50 back-to-back launches between one event pair"},
        "e2e": {"value": e2e, "unit": "scans/s", "h2d_bytes_per_step": 16 * (N_MAP + N_SCAN) + 64, "d2h_bytes_per_step": int(368 * st2["eval_launches"] / K + 64),
                "ms_per_step": e2e_ms / K, "max_abs_translation_error_m": st2["err"], "build_ms": st2["build_ms"] / K, "align_ms": st2["align_ms"] / K,
                "session_eval_us": st2["eval_ms"] / max(st2["eval_launches"], 1) * 1e3,
                "what": "host buffers in, result out, through the C ABI: ndtb_stage_target / ndtb_stage_source start the copy of step i + 1's map "
                        "and scan (18 MB, pinned host memory, copy stream) before ndtb_commit_target / ndtb_commit_source / ndtb_align of step i; "
                        "the step ends with ndtb_stage_sync, i.e. once that copy has arrived, so every timed step holds one full H2D copy, "
                        "one map build, one align and the result read-back"},
        "e2e_serial": {"value": world * K / (ser_ms * 1e-3), "unit": "scans/s", "h2d_bytes_per_step": 16 * (N_MAP + N_SCAN) + 64,
                       "d2h_bytes_per_step": int(368 * st6["eval_launches"] / K + 64), "ms_per_step": ser_ms / K,
                       "what": "the same step without pipelining: ndtb_set_target + ndtb_set_source (copy, then build) + ndtb_align, one after the other"},
        "e2e_window": {"value": world * K / (win_ms * 1e-3), "unit": "scans/s", "h2d_bytes_per_step": 16 * N_SCAN + 64 + 128 * n_kf,
                       "d2h_bytes_per_step": int(368 * st4["eval_launches"] / K + 64), "ms_per_step": win_ms / K,
                       "what": "same step through ndtb_set_target_window: the 1M-pt target is assembled on the device from 5 archived keyframe "
                               "clouds (resident, uploaded once) with their poses (identity here, so the map is bit-identical to the other regions), "
                               "then built; only the new scan is copied from pinned host memory",
                       "max_abs_translation_error_m": st4["err"]},
        "e2e_packets": {"value": world * K / (pkt_ms * 1e-3), "unit": "scans/s", "h2d_bytes_per_step": (n_pk + 1) * pk_bytes + 64 + 128 * n_kf,
                        "d2h_bytes_per_step": int(368 * st5["eval_launches"] / K + 64 + 32), "ms_per_step": pkt_ms / K,
                        "what": "the e2e_window step with the scan arriving as the sensor sends it: 64 RNG19_RFL8_SIG16_NIR16 lidar packets "
                                "(one 1024 x 128 revolution of the same synthetic scene, ray-cast, no padding) + the packet that closes the "
                                "frame, from pinned host memory; decoded on the device (lidar_decode_kernel, the reference's "
                                "LidarCallback::DecodePacketRng19) into the NDT source without leaving HBM",
                        "points_per_scan": int(pk_state["points"]), "newton_iterations": st5["iters"] / K,
                        "max_abs_translation_error_m": st5["err"],
                        "decode_kernel_us": decode_us, "decode_kernel_launches": int(decode_launches),
                        "decode_algorithmic_bytes": int(n_pk * pk_bytes + 12 * 1024 * 128 + 42 * pk_state["points"]),
                        "decode_gbs": (n_pk * pk_bytes + 12 * 1024 * 128 + 42 * pk_state["points"]) / max(decode_us, 1e-3) / 1e3},
        "gpu_launches": launches,
        "device_newton": dict(h.session_stats(), what="opt-in (NDTB_BENCH_DEVICE_NEWTON=1; off by default because the host step is faster, "
                              "DESIGN 4.4): Newton steps the resident kernel took itself and the host confirmed word for word (accepted) / "
                              "handed to the host (declined) / that the host had to repair (mismatches)"),
        "roofline": {"kernel": "eval_kernel_pclomp<7,hessian>", "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(), "peak_source": peak_src, "algorithmic_bytes_per_launch": b_eval,
                     "mean_hits_per_point": hbar, "kernel_us": t_eval * 1e6,
                     "session_eval_us": st["eval_ms"] / max(st["eval_launches"], 1) * 1e3,
                     "kernel_us_back_to_back": b2b_us, "frac_back_to_back": b_eval / (b2b_us * 1e-6) / 1e9 / peak},
        "clocks": clk,
    }
    # the kernel is L2-resident and issue/latency-bound: its instruction-issue floor beside the HBM roofline
    inst = ncu_inst_executed()
    if inst and clk and clk.get("sm_mhz"):
        floor_us = inst / (148 * 4) / (clk["sm_mhz"] * 1e6) * 1e6
        sess_us = line["roofline"]["session_eval_us"]
        line["roofline"]["issue"] = {"bound": "warp-instruction issue (4 schedulers x 148 SMs, 1 instruction per cycle each)",
                                     "inst_executed_per_launch": inst, "floor_us": floor_us, "frac_of_kernel_us": floor_us / (t_eval * 1e6),
                                     "frac_of_session_eval_us": floor_us / sess_us if sess_us else None,
                                     "source": "ncu smsp__inst_executed.sum (profiles/eval_kernel_traffic.json) / live SM clock"}
    # the voxel-map build is the other half of the step: its own roofline entry (B_build = 2*16*N_t + 64*V)
    l0 = h.launch_count()
    assert h.set_target_device(d_tgt.data_ptr(), N_MAP) == 0
    build_launches = h.launch_count() - l0
    n_leaves = int(len(h.voxels()["idx"]))
    b_build = 2 * 16 * N_MAP + 64 * n_leaves
    build_s = st["build_ms"] / K * 1e-3
    line["roofline_build"] = {"kernel": f"voxel-map build ({build_launches} launches)", "bound": "hbm", "achieved": b_build / build_s / 1e9, "peak": peak,
                              "unit": "GB/s", "frac": b_build / build_s / 1e9 / peak, "algorithmic_bytes": b_build, "leaves": n_leaves,
                              "build_us": build_s * 1e6, "traffic": build_traffic(),
                              "note": "CUDA events on the library's stream around the whole build (sort_segments_kernel + voxel_stats_kernel, one host "
                                      "wait at the end); latency-bound, not bandwidth-bound: DESIGN 4.3; traffic = ncu dram bytes of the two kernels "
                                      "(profiles/build_kernels_traffic.json)"}
    if not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        c = run_cpu(tgt, scans, guesses, steps=2, warmup=1, threads=threads)
        line["parity"] = parity_block(h, tgt, scans, guesses, c["results"])
        assert line["parity"]["ok"], f"parity bars missed: {line['parity']}"
        line["cpu_baseline"] = {"value": 1.0 / c["mean_s"], "unit": "scans/s", "cores": threads, "kind": "port",
                                "sample": f"2 steps (after 1 warm-up) of the same workload: 1M-pt map build ({c['build_s']*1e3:.0f} ms, serial like the "
                                          f"reference) + align ({c['iterations']:.0f} iterations) with {threads} OpenMP threads"}
        # the scan producer's CPU side: the oracle restatement of DecodePacketRng19 (one thread, like the reference's callback)
        # decoding the same 65 packets, timed beside the device decode of e2e_packets
        import oracle_lib as O
        od = O.OracleLidar(lcfg, az_deg, alt_deg)
        pk_np = p_packets[0].numpy()
        pk_np[:n_pk, 2] = 1; pk_np[:n_pk, 3] = 0; pk_np[n_pk, 2] = 2; pk_np[n_pk, 3] = 0
        od.decode(pk_np)
        t0 = time.perf_counter()
        reps = 5
        for r in range(reps):
            pk_np[:n_pk, 2] = (3 + 2 * r) & 0xff; pk_np[n_pk, 2] = (4 + 2 * r) & 0xff
            fr = od.L.orc_lidar_decode(od.h, pk_np.ctypes.data, pk_np.shape[0], pk_np.shape[1], pk_np.shape[1])
        line["e2e_packets"]["decode_cpu_port_us"] = (time.perf_counter() - t0) / reps * 1e6
        line["e2e_packets"]["decode_cpu_port_what"] = "oracle/ouster_decode.hpp (restatement of LidarCallback::DecodePacketRng19), 1 thread, same 65 packets"
    return line


def svn_initial_particles(K):
    """K particles around the identity prior, seed 7, sigmas of svn_ndt_impl.hpp:710 (translation 0.05 m, rotation 0.01 rad)."""
    from slam_sam_b200 import synth

    rng = np.random.Generator(np.random.PCG64(7))
    return np.stack([synth.pose_matrix(rng.normal(0, 0.05, 3), rng.normal(0, 0.01, 3)) for _ in range(K)])


def make_svn_workload(n_scans):
    """configs[2]: one target scan + `n_scans` source scans of the same place (131 072 points each)."""
    from slam_sam_b200 import synth

    scene = synth.Scene()
    A = synth.trajectory_pose(40.0)
    tgt = synth.ouster_scan(scene, A, seed=1)
    pool = []
    for i in range(n_scans):
        B = A @ synth.pose_matrix((0.5 - 0.1 * i, 0.1, 0.02), np.deg2rad((0.5, 0.3, 2.0 - 0.3 * i)))
        pool.append((synth.ouster_scan(scene, B, seed=2 + i), np.linalg.inv(A) @ B))
    return tgt, pool


def run_gpu_svn(args, rank, world, local_rank, sub=False):
    """configs[2]: SVN-NDT, K = 100 particles, 131072-pt scan pair; particles sharded over the ranks (strong scaling).  One step =
    one full align(source, prior).  `sub`: a short run whose summary becomes the "svn" object of the default bench line."""
    import torch
    import torch.distributed as dist

    from slam_sam_b200 import distributed as D
    from slam_sam_b200 import ndtb

    K = 100
    tgt, pool = make_svn_workload(2 if sub else 4)
    h = ndtb.Handle(local_rank)
    h.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=RESOLUTION, min_points_per_voxel=3, search_method=ndtb.DIRECT7)
    h.set_svn_params(particle_count=K, max_iterations=50, kernel_bandwidth=1.0, step_size=1.0, stop_threshold=1e-4, use_gauss_newton=1)
    assert h.set_target(tgt) == 0
    p_scans = [torch.from_numpy(s).pin_memory() for s, _ in pool]
    d_scans = [torch.from_numpy(s).cuda() for s, _ in pool]
    prior = np.eye(4)
    P0 = svn_initial_particles(K)
    stream = torch.cuda.ExternalStream(h.stream_ptr(), device=local_rank)
    comm = D.SvnComm(h, rank, world, local_rank) if world > 1 else None

    def align(resident, k):
        if resident:
            assert h.set_source_device(d_scans[k].data_ptr(), N_SCAN) == 0
        else:
            assert h.L.ndtb_set_source(h.h, p_scans[k].data_ptr(), N_SCAN, 16) == 0
        return h.svn_align(prior, P0) if comm is None else comm.align(prior, P0)

    def step(i, resident):
        k = i % len(pool)
        r = align(resident, k)
        return r, float(np.abs(r["final_pose"][:3, 3] - pool[k][1][:3, 3]).max())

    def timed(resident, steps, warmup):
        for i in range(warmup):
            step(i, resident)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        l0 = h.launch_count()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        iters = 0; err = 0.0; eval_ms = 0.0; stein_ms = 0.0; xchg_ms = 0.0
        e0.record(stream)
        for i in range(steps):
            r, er = step(warmup + i, resident)
            iters += r["iterations"]; err = max(err, er)
            t = h.svn_timings()
            eval_ms += t["eval_ms"]; stein_ms += t["stein_ms"]; xchg_ms += t["exchange_wait_ms"]
        e1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        return e0.elapsed_time(e1), dict(iters=iters, err=err, launches=h.launch_count() - l0, eval_ms=eval_ms, stein_ms=stein_ms, xchg_ms=xchg_ms,
                                         hits=h.last_hits()[0])

    clocks = None
    if not sub:
        clocks = ClockSampler(local_rank); clocks.start()
    steps = 3 if sub else max(1, min(args.steps, 10))
    total_ms, st = timed(True, steps, 1 if sub else min(args.warmup, 3))
    e2e_ms, st2 = timed(False, steps, 1)
    clk = clocks.stop() if clocks else None
    total_ms = D.max_over_ranks(total_ms, local_rank); e2e_ms = D.max_over_ranks(e2e_ms, local_rank)
    launches = int(D.sum_over_ranks(st["launches"], local_rank))
    # the sharded result must be the single-GPU result, bit for bit (same records, same Stein update on every rank)
    equal = None
    if world > 1:
        r_sh = align(True, 0)
        if rank == 0:
            h1 = ndtb.Handle(local_rank)
            h1.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=RESOLUTION, min_points_per_voxel=3, search_method=ndtb.DIRECT7)
            h1.set_svn_params(particle_count=K, max_iterations=50, kernel_bandwidth=1.0, step_size=1.0, stop_threshold=1e-4, use_gauss_newton=1)
            assert h1.set_target(tgt) == 0 and h1.set_source_device(d_scans[0].data_ptr(), N_SCAN) == 0
            r_1 = h1.svn_align(prior, P0)
            equal = bool(r_sh["iterations"] == r_1["iterations"] and np.array_equal(r_sh["final_pose"], r_1["final_pose"])
                         and np.array_equal(r_sh["particles"], r_1["particles"]) and np.array_equal(r_sh["final_covariance"], r_1["final_covariance"]))
            h1.close()
        dist.barrier()
    if comm is not None:
        comm.close()
    h.close()
    if rank != 0:
        return None
    assert equal is not False, "sharded SVN align differs from the single-GPU align"
    k_local = len(range(0, K, world))
    hb = st["hits"]
    b64 = k_local * N_SCAN * (16 + 8 * 7 + 64 * hb)   # K * B_eval with the 64-byte record
    b96 = k_local * N_SCAN * (16 + 8 * 7 + 96 * hb)   # what the svn kernel reads: FP64 mean (24 B) + FP64 inverse covariance (72 B)
    t_eval = st["eval_ms"] / max(st["iters"], 1) * 1e-3
    peak, peak_src = measured_peak()
    roof = {"kernel": "eval_kernel<svn,DIRECT7,GN>", "bound": "hbm", "achieved": b64 / t_eval / 1e9, "peak": peak, "unit": "GB/s",
            "frac": b64 / t_eval / 1e9 / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": b64,
            "frac_with_96_byte_fp64_records": b96 / t_eval / 1e9 / peak, "mean_hits_per_point": hb, "kernel_us": t_eval * 1e6,
            "note": "frac uses SURVEY 8(d)'s formula (64-byte record per hit); the svn kernel actually gathers 96 bytes of FP64 mean + inverse covariance per hit"}
    summary = {"value": steps / (total_ms * 1e-3), "unit": "aligns/s", "ms_per_step": total_ms / steps, "steps": steps, "scaling": "strong",
               "particles": K, "particles_on_rank0": k_local, "iterations": st["iters"] / steps,
               "eval_us_per_iter": st["eval_ms"] / max(st["iters"], 1) * 1e3, "stein_us_per_iter": st["stein_ms"] / max(st["iters"], 1) * 1e3,
               "allreduce_us_per_iter": st["xchg_ms"] / max(st["iters"], 1) * 1e3,
               "exchange": "none (one GPU)" if world == 1 else "records peer-written over NVLink by the eval kernel's last block; allreduce_us_per_iter = time the Stein kernel waited for peer records",
               "e2e_value": steps / (e2e_ms * 1e-3), "max_abs_translation_error_m": st["err"], "sharded_equals_single_gpu": equal,
               "roofline_frac": roof["frac"], "gpu_launches": launches}
    if sub:
        return summary
    return {"metric": "NDT align scans/sec @131k pts (1/2/4/8 B200) + % gather-BW roofline vs CPU OMP", "value": steps / (total_ms * 1e-3), "unit": "scans/s",
            "n_gpus": world, "steps": steps, "warmup": min(args.warmup, 3), "ms_per_step": total_ms / steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64 Mahalanobis/exp, f32 products, f64 accumulate", "data": "synthetic",
            "config": {"workload": "configs[2]: SVN-NDT align, K=100 particles, 131072-pt OS1-128-shaped scan pair, 1.0 m voxels, Gauss-Newton Hessian, per step",
                       "particles": K, "parallelism": f"particles sharded x{world}, per-particle records exchanged once per iteration", "svn_iterations_per_align": st["iters"] / steps,
                       "max_abs_translation_error_m": st["err"], "l2": "inputs of one align (K x 131072 point evaluations per iteration) exceed nothing: L2-resident by design"},
            "e2e": {"value": steps / (e2e_ms * 1e-3), "unit": "scans/s", "h2d_bytes_per_step": 16 * N_SCAN + 16 * 8 * K, "d2h_bytes_per_step": 16 * 8 * K + 53 * 8,
                    "ms_per_step": e2e_ms / steps},
            "gpu_launches": launches, "svn": summary, "roofline": roof, "clocks": clk}


def run_gpu_replay(args, rank, world, local_rank, sub=False):
    """configs[3]: replay batch.  Every rank holds the (replicated, built once) 1M-pt map and aligns its own shard of the scans
    through ndtb_align_batch from pinned host buffers; no data-path collective (weak scaling).  One step = one batch of 32 scans."""
    import torch
    import torch.distributed as dist

    from slam_sam_b200 import distributed as D
    from slam_sam_b200 import ndtb

    tgt, scans, guesses, truths = make_workload(rank)
    B = 32
    h = ndtb.Handle(local_rank)
    h.set_params(resolution=RESOLUTION, max_iterations=35, trans_epsilon=0.01, step_size=0.1, search_method=ndtb.DIRECT7)
    assert h.set_target(tgt) == 0
    pinned = [torch.from_numpy(s).pin_memory() for s in scans]
    batch = [pinned[i % len(pinned)].numpy() for i in range(B)]
    gb = np.stack([guesses[i % len(guesses)] for i in range(B)])
    stream = torch.cuda.ExternalStream(h.stream_ptr(), device=local_rank)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def timed(steps, warmup, interleave):
        h.set_option("batch_interleave", interleave)
        for _ in range(warmup):
            h.align_batch(batch, gb)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        l0 = h.launch_count()
        ev, evals, err = [], 0, 0.0
        for _ in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            res = h.align_batch(batch, gb)
            e1.record(stream)
            ev.append((e0, e1))
            evals += sum(r["evaluations"] for r in res)
            err = max(err, max(float(np.abs(r["T"][:3, 3] - truths[i % len(truths)][:3, 3]).max()) for i, r in enumerate(res)))
            assert all(r["converged"] for r in res)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        return sum(a.elapsed_time(b) for a, b in ev), dict(evals=evals, err=err, launches=h.launch_count() - l0)

    clocks = None
    if not sub:
        clocks = ClockSampler(local_rank); clocks.start()
    steps = 3 if sub else max(1, min(args.steps, 10))
    ms, st = timed(steps, 3, 1)
    ms_serial, st_s = timed(steps, 1, 0)
    clk = clocks.stop() if clocks else None
    ms = D.max_over_ranks(ms, local_rank); ms_serial = D.max_over_ranks(ms_serial, local_rank)
    launches = int(D.sum_over_ranks(st["launches"], local_rank))
    if world > 1:
        dist.barrier()
    hbar = h.last_hits()[0]
    h.close()
    if rank != 0:
        return None
    if sub:
        return {"value": world * steps * B / (ms * 1e-3), "unit": "scans/s", "ms_per_step": ms / steps, "steps": steps, "scaling": "weak",
                "scans_per_step_per_rank": B, "evaluations_per_scan": st["evals"] / (steps * B), "serial_scans_per_s": world * steps * B / (ms_serial * 1e-3),
                "wall_us_per_evaluation": ms * 1e3 / st["evals"], "max_abs_translation_error_m": st["err"], "gpu_launches": launches,
                "what": "configs[3]: 32 scans per step and rank through ndtb_align_batch from pinned host buffers against the replicated 1M-pt map "
                        "(built once), scans sharded by rank, no collective"}
    b_eval = N_SCAN * (16 + 8 * 7 + 64 * hbar)
    t_eval = ms * 1e-3 / st["evals"]        # wall time per evaluation with three alignments in flight: the GPU is never idle
    peak, peak_src = measured_peak()
    v = world * steps * B / (ms * 1e-3)
    return {"metric": "NDT align scans/sec @131k pts (1/2/4/8 B200) + % gather-BW roofline vs CPU OMP", "value": v, "unit": "scans/s",
            "n_gpus": world, "steps": steps, "warmup": 3, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 pair maths, f64 accumulate", "data": "synthetic",
            "config": {"workload": "configs[3]: replay batch, 32 synthetic 131072-pt OS1-128-shaped scans per step and rank against a replicated 1M-pt map "
                                   "(built once), DIRECT7, 1.0 m voxels, through ndtb_align_batch from pinned host buffers",
                       "scans_per_step_per_rank": B, "parallelism": f"scan shards x{world} (no collective)", "l2": "256 MB flush between timed steps",
                       "evaluations_per_scan": st["evals"] / (steps * B), "max_abs_translation_error_m": st["err"],
                       "serial_scans_per_s": world * steps * B / (ms_serial * 1e-3),
                       "note": "value == e2e: the batch entry point takes host buffers; three alignments are kept in flight on one stream "
                               "(serial_scans_per_s is the same batch with batch_interleave = 0)"},
            "e2e": {"value": v, "unit": "scans/s", "h2d_bytes_per_step": 16 * N_SCAN * B, "d2h_bytes_per_step": int(736 * st["evals"] / steps), "ms_per_step": ms / steps},
            "gpu_launches": launches,
            "roofline": {"kernel": "eval_kernel_pclomp<7,hessian>", "bound": "hbm", "achieved": b_eval / t_eval / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": b_eval / t_eval / 1e9 / peak, "traffic": ncu_traffic(), "peak_source": peak_src, "algorithmic_bytes_per_launch": b_eval,
                         "mean_hits_per_point": hbar, "kernel_us": t_eval * 1e6,
                         "note": "kernel_us here is WALL time per evaluation of the whole batch (launch gaps, uploads and host work included)"},
            "clocks": clk}


def run_reference(args, rank):
    if rank != 0:
        return None
    tgt, scans, guesses, truths = make_workload(0)
    threads = os.cpu_count() or 1
    steps, warmup = max(1, args.steps), max(0, args.warmup)   # ~0.4 s of CPU work per step: 20 + 5 steps end within ~15 s
    c = run_cpu(tgt, scans, guesses, steps=steps, warmup=warmup, threads=threads)
    v = 1.0 / c["mean_s"]
    sample = f"{steps} steps of the same workload (1M-pt map build + 131072-pt align), {threads} OpenMP threads, oracle port of ndt_omp"
    return {"impl": "reference", "metric": "NDT align scans/sec @131k pts (1/2/4/8 B200) + % gather-BW roofline vs CPU OMP", "value": v, "unit": "scans/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": c["mean_s"] * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 pair maths, f64 accumulate", "data": "synthetic",
            "config": {"workload": "configs[1]: voxel-map build of a 1M-pt synthetic map + DIRECT7 NDT align of one 131072-pt OS1-128-shaped scan, "
                                   "1.0 m voxels, per step", "build_ms": c["build_s"] * 1e3, "newton_iterations": c["iterations"]},
            "cpu_baseline": {"value": v, "unit": "scans/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "scans/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def run_gpu_multigrid(args, rank, world, local_rank):
    """Multigrid NDT (pclomp::MultiGridNormalDistributionsTransform) as a dynamic-map localiser drives it: the 1M-pt map is held
    as 16 tiles (4 x 4 by x / y quantiles); one step = one tile replaced (addTarget from pinned host memory: upload + voxel-map
    build of that tile), createVoxelKdtree, and one align of a 131072-pt scan.  Replicas only across GPUs (no collective)."""
    import torch
    import torch.distributed as dist

    from slam_sam_b200 import distributed as D
    from slam_sam_b200 import ndtb

    tgt, scans, guesses, truths = make_workload(rank)
    qx = np.quantile(tgt[:, 0], np.linspace(0, 1, 5)); qy = np.quantile(tgt[:, 1], np.linspace(0, 1, 5))
    qx[-1] += 1; qy[-1] += 1
    tiles = {}
    for i in range(4):
        for j in range(4):
            m = (tgt[:, 0] >= qx[i]) & (tgt[:, 0] < qx[i + 1]) & (tgt[:, 1] >= qy[j]) & (tgt[:, 1] < qy[j + 1])
            tiles["tile_%d_%d" % (i, j)] = torch.from_numpy(np.ascontiguousarray(tgt[m])).pin_memory()
    keys = sorted(tiles)
    h = ndtb.Handle(local_rank)
    h.set_params(resolution=RESOLUTION, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
    for k in keys:
        assert h.mg_add_target(tiles[k].numpy(), k) == 0
    assert h.mg_create_kdtree() == 0
    pinned = [torch.from_numpy(sc).pin_memory() for sc in scans]
    stream = torch.cuda.ExternalStream(h.stream_ptr(), device=local_rank)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def step(i):
        k = keys[i % len(keys)]
        assert h.mg_add_target(tiles[k].numpy(), k) == 0          # tile swap: H2D + build of one tile
        assert h.mg_create_kdtree() == 0
        assert h.set_source(pinned[i % len(pinned)].numpy()) == 0
        return h.mg_align(guesses[i % len(guesses)])

    for i in range(max(args.warmup, 3)):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local_rank); clocks.start()
    l0 = h.launch_count()
    ev, evals, err, iters = [], 0, 0.0, 0
    for i in range(args.steps):
        with torch.cuda.stream(stream):
            flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        r = step(i)
        e1.record(stream)
        ev.append((e0, e1))
        assert r["converged"]
        evals += r["evaluations"]; iters += r["iterations"]
        err = max(err, float(np.abs(r["T"][:3, 3] - truths[i % len(truths)][:3, 3]).max()))
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    launches = h.launch_count() - l0
    # the evaluation kernel alone: event pairs around single evaluations at the converged pose (kernel + the one-block reduction)
    p = np.zeros(6)
    kt = []
    for _ in range(20):
        with torch.cuda.stream(stream):
            flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream); h.mg_compute_derivatives(p, T=r["T"]); e1.record(stream)
        torch.cuda.synchronize()
        kt.append(e0.elapsed_time(e1))
    clk = clocks.stop()
    ms = D.max_over_ranks(ms, local_rank)
    launches = int(D.sum_over_ranks(launches, local_rank))
    if world > 1:
        dist.barrier()
    if rank != 0:
        return None
    hbar = h.last_hits()[0]
    info = h.mg_info()
    # algorithmic bytes of one evaluation: the point (16 B), 27 hash slots (8 B) per overlapping grid (~1), and per hit the
    # centroid (16 B) + FP64 mean and inverse covariance (24 + 72 B)
    b_eval = N_SCAN * (16 + 8 * 27 + (16 + 96) * hbar)
    t_eval = float(np.median(kt)) * 1e-3
    peak, peak_src = measured_peak()
    tile_pts = int(np.mean([len(t) for t in tiles.values()]))
    v = world * args.steps / (ms * 1e-3)
    return {"metric": "NDT align scans/sec @131k pts (1/2/4/8 B200) + % gather-BW roofline vs CPU OMP", "value": v, "unit": "scans/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "multigrid NDT (pclomp::MultiGridNormalDistributionsTransform): 1M-pt synthetic map held as 16 tiles; per step one "
                                   "tile replaced from pinned host memory (addTarget), createVoxelKdtree, align of one 131072-pt scan, 1.0 m voxels",
                       "tiles": len(keys), "points_per_tile": tile_pts, "leaves_in_tree": info["leaves"], "parallelism": f"replicas x{world} (no collective)",
                       "l2": "256 MB flush between timed steps", "newton_iterations": iters / args.steps, "evaluations_per_align": evals / args.steps,
                       "max_abs_translation_error_m": err, "note": "value == e2e: tile and scan come from pinned host buffers every step"},
            "e2e": {"value": v, "unit": "scans/s", "h2d_bytes_per_step": 16 * (N_SCAN + tile_pts), "d2h_bytes_per_step": int(368 * evals / args.steps),
                    "ms_per_step": ms / args.steps},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "mg_eval_kernel<hessian>", "bound": "hbm", "achieved": b_eval / t_eval / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": b_eval / t_eval / 1e9 / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": b_eval,
                         "mean_hits_per_point": hbar, "kernel_us": t_eval * 1e6,
                         "note": "all-FP64 pair maths, 250 registers -> 8 warps per SM: latency-bound (ncu: issue-active 24 %, FP64 pipe 15 %, 3.5 MB of DRAM traffic per launch, profiles/r01_mg_eval_kernel_ncu_full.txt), not HBM bound; the HBM figure is reported for comparability with the pclomp kernel"},
            "clocks": clk}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub-workloads", action="store_true", help="skip the replay / svn sub-objects of the default line")
    ap.add_argument("--workload", default="map_align", choices=["map_align", "svn", "replay", "multigrid"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        line = run_reference(args, rank)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    if args.workload == "svn":
        line = run_gpu_svn(args, rank, world, local_rank)
    elif args.workload == "replay":
        line = run_gpu_replay(args, rank, world, local_rank)
    elif args.workload == "multigrid":
        line = run_gpu_multigrid(args, rank, world, local_rank)
    else:
        # the headline line is configs[1] (replicas); the two workloads north_star shards across GPUs ride along as sub-objects,
        # measured with a bounded step count in the same process group: configs[3] replay (scans sharded, weak scaling) and
        # configs[2] SVN (particles sharded, strong scaling, records exchanged by the kernels over NVLink)
        line = run_gpu(args, rank, world, local_rank)
        if not args.no_sub_workloads:
            torch.cuda.empty_cache()
            replay = run_gpu_replay(args, rank, world, local_rank, sub=True)
            torch.cuda.empty_cache()
            svn = run_gpu_svn(args, rank, world, local_rank, sub=True)
            if line is not None:
                line["replay"] = replay
                line["svn"] = svn
                line["gpu_launches"] += replay["gpu_launches"] + svn["gpu_launches"]
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
