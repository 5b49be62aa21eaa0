"""Sliding-window target assembly (the step before setInputTarget, run/pipeline_ligo_tc.cpp:519-526):
CPU -- the oracle restatement of pcl::transformPointCloud<PointT,double> + concatenation against an independent
numpy float64 evaluation; GPU -- ndtb_keyframe_put / ndtb_set_target_window bit-exact against the oracle, and the
voxel map / alignment identical to uploading the host-assembled window through ndtb_set_target."""
import numpy as np
import pytest

import oracle_lib as O
from slam_sam_b200 import ndtb, synth


def make_window(seed=0, n_kf=5, beams=16, cols=256):
    """n_kf keyframe scans in their own body frames + the poses that put them into the map frame."""
    scene = synth.Scene()
    rng = np.random.default_rng(seed)
    clouds, poses = [], []
    for k in range(n_kf):
        Tb2m = synth.trajectory_pose(30.0 + 1.5 * k)
        scan_map = synth.ouster_scan(scene, Tb2m, beams=beams, cols=cols, seed=10 + k)   # map-frame points of the scan
        # express in the body frame in double, round to float: what an archived keyframe cloud holds
        Tm2b = np.linalg.inv(Tb2m)
        body = (scan_map[:, :3].astype(np.float64) @ Tm2b[:3, :3].T + Tm2b[:3, 3]).astype(np.float32)
        body = body[: len(body) - int(rng.integers(0, 50))]                               # ragged sizes
        clouds.append(np.ascontiguousarray(body))
        poses.append(Tb2m)
    return clouds, np.stack(poses)


def numpy_window(clouds, poses):
    out = []
    for c, T in zip(clouds, poses):
        x, y, z = (c[:, i].astype(np.float64) for i in range(3))
        rows = [((T[r, 0] * x + T[r, 1] * y) + T[r, 2] * z) + T[r, 3] for r in range(3)]
        out.append(np.stack(rows + [np.ones_like(x)], axis=1).astype(np.float32))
    return np.concatenate(out)


def test_oracle_window_matches_numpy_float64(oracle):
    clouds, poses = make_window()
    a = O.assemble_window(clouds, poses)
    b = numpy_window(clouds, poses)
    assert a.shape == b.shape and np.array_equal(a, b)
    # identity poses reproduce the clouds; a single keyframe works; order is the listed order
    ident = np.stack([np.eye(4)] * len(clouds))
    assert np.array_equal(O.assemble_window(clouds, ident)[:, :3], np.concatenate(clouds)[:, :3])
    assert np.array_equal(O.assemble_window(clouds[::-1], poses[::-1])[: len(clouds[-1])], numpy_window(clouds[-1:], poses[-1:]))


@pytest.mark.gpu
def test_gpu_window_bit_exact_and_same_map(oracle):
    clouds, poses = make_window(seed=1)
    h = ndtb.Handle(0)
    h.set_params(resolution=1.0)
    for k, c in enumerate(clouds):
        assert h.keyframe_put(100 + k, c) == 0
    assert h.keyframe_count() == (len(clouds), sum(len(c) for c in clouds))
    ids = [100 + k for k in range(len(clouds))]
    assert h.set_target_window(ids, poses) == 0
    want = O.assemble_window(clouds, poses)
    got = h.target_points()
    assert got.shape == want.shape and np.array_equal(got, want)          # bit-exact, order preserved
    v_win = h.voxels()
    h2 = ndtb.Handle(0); h2.set_params(resolution=1.0)
    assert h2.set_target(want) == 0
    v_host = h2.voxels()
    for k in ("idx", "n", "mean", "icov"):
        assert np.array_equal(v_win[k], v_host[k])
    # a different window (subset, different order, updated poses) reuses the archive without new uploads
    poses2 = poses.copy(); poses2[:, 0, 3] += 0.25
    sub = [ids[3], ids[1], ids[4]]
    assert h.set_target_window(sub, poses2[[3, 1, 4]]) == 0
    assert np.array_equal(h.target_points(), O.assemble_window([clouds[3], clouds[1], clouds[4]], poses2[[3, 1, 4]]))
    # alignment against the device-assembled window == alignment against the uploaded one
    scene = synth.Scene()
    Ts = synth.trajectory_pose(30.0 + 1.5 * len(clouds))
    scan = synth.ouster_scan(scene, Ts, beams=16, cols=256, seed=99)
    body = (scan[:, :3].astype(np.float64) @ np.linalg.inv(Ts)[:3, :3].T + np.linalg.inv(Ts)[:3, 3]).astype(np.float32)
    guess = (Ts @ synth.pose_matrix((0.1, -0.05, 0.0), np.deg2rad((0.2, 0.1, 0.5)))).astype(np.float32)
    assert h.set_target_window(ids, poses) == 0
    h.set_source(body); h2.set_source(body)
    r1, r2 = h.align(guess), h2.align(guess)
    assert np.array_equal(h.trace(), h2.trace()) and np.array_equal(r1["T"], r2["T"])
    assert np.abs(r1["T"][:3, 3] - Ts[:3, 3]).max() < 0.3       # same basin (a 16-beam scan against a 5-keyframe map)


@pytest.mark.gpu
def test_gpu_keyframe_management():
    clouds, poses = make_window(seed=2, n_kf=3)
    h = ndtb.Handle(0)
    assert h.set_target_window([1], poses[:1]) == ndtb.NDTB_ERR_INVALID          # unknown id
    assert h.keyframe_put(1, clouds[0]) == 0 and h.keyframe_put(2, clouds[1]) == 0
    assert h.keyframe_put(1, clouds[2]) == 0                                      # replace in place
    assert h.keyframe_count() == (2, len(clouds[2]) + len(clouds[1]))
    assert h.keyframe_drop(2) == 0 and h.keyframe_drop(2) == ndtb.NDTB_ERR_INVALID
    assert h.keyframe_put(3, np.zeros((0, 3), np.float32)) == ndtb.NDTB_ERR_EMPTY
    assert h.set_target_window([], np.zeros((0, 4, 4))) == ndtb.NDTB_ERR_EMPTY
    # 20 keyframes: more than one launch of the assembly kernel (16 segments each)
    many, mp = make_window(seed=3, n_kf=20, beams=8, cols=128)
    for k, c in enumerate(many):
        assert h.keyframe_put(1000 + k, c) == 0
    assert h.set_target_window([1000 + k for k in range(20)], mp) == 0
    assert np.array_equal(h.target_points(), O.assemble_window(many, mp))
