"""BASELINE.json configs[4] (large map stress: 20 M-point map at 0.5 m voxels, multi-resolution 2.0 / 1.0 / 0.5 m
align) and configs[3] (replay batch) as parity cases: the 20 M-point voxel map bit-exact against the CPU oracle,
size-independent properties, and the coarse-to-fine alignment against the oracle's on the same inputs."""
import numpy as np
import pytest

import oracle_lib as O
from slam_sam_b200 import distributed as D
from slam_sam_b200 import ndtb, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def big():
    scene = synth.Scene()
    tgt = synth.surface_map(scene, 20_000_000, seed=21)
    pose = synth.trajectory_pose(45.0)
    scan = synth.ouster_scan(scene, pose, seed=5)
    return dict(scene=scene, target=tgt, pose=pose, scan=scan)


def test_20m_point_map_bit_exact(oracle, big):
    h = ndtb.Handle(0)
    h.set_params(resolution=0.5)
    assert h.set_target(big["target"]) == 0
    v = h.voxels()
    assert len(v["idx"]) > 300_000                                          # ~370 k occupied leaves (surfaces only)
    assert np.all(np.diff(v["idx"]) > 0)
    assert int(v["n"].clip(0).sum()) <= len(big["target"])
    o = O.OracleGrid(0, 0.5, 6)
    assert o.build(big["target"]) == 0
    w = o.export()
    assert np.array_equal(v["min_b"], w["min_b"]) and np.array_equal(v["div_b"], w["div_b"])
    assert np.array_equal(v["idx"], w["idx"]) and np.array_equal(v["n"], w["n"])
    valid = w["n"] >= 1
    assert np.array_equal(v["mean"][valid], w["mean"][valid]) and np.array_equal(v["icov"], w["icov"])
    # build time at this size, second build (the first one allocates gigabytes of scratch): 20x the 1 M-point build is ~4 ms
    assert h.set_target(big["target"]) == 0
    assert h.timings()["build_ms"] < 20.0
    assert np.array_equal(h.voxels()["idx"], w["idx"])


def test_multi_resolution_align_on_20m_map(oracle, big):
    """Coarse-to-fine: align at 2.0 m, then 1.0 m from that pose, then 0.5 m (setResolution rebuilds the map,
    ndt_omp.h:135-142).  Same chain on the oracle; every stage must agree."""
    guess = (big["pose"] @ synth.pose_matrix((0.4, -0.3, 0.05), np.deg2rad((0.5, -0.4, 1.5)))).astype(np.float32)
    # the oracle works on a 4 M-point crop around the scan to keep its serial map build short; the GPU gets the same crop
    c = big["pose"][:3, 3]
    t = big["target"]
    crop = np.ascontiguousarray(t[(np.abs(t[:, 0] - c[0]) < 60) & (np.abs(t[:, 1] - c[1]) < 60)])
    assert len(crop) > 2_000_000
    h = ndtb.Handle(0)
    h.set_params(resolution=2.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
    assert h.set_target(crop) == 0 and h.set_source(big["scan"]) == 0
    o = O.OracleNdt(resolution=2.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1, num_threads=16)
    o.set_target(crop); o.set_source(big["scan"])
    Tg, To = guess, guess
    for res in (2.0, 1.0, 0.5):
        h.set_params(resolution=res, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
        r = h.align(Tg)
        oo = O.OracleNdt(resolution=res, max_iterations=35, trans_epsilon=0.01, step_size=0.1, num_threads=16)
        oo.set_target(crop); oo.set_source(big["scan"])
        ro = oo.align(To)
        assert (r["iterations"], r["converged"]) == (ro["iterations"], ro["converged"])
        d = np.linalg.inv(r["T"].astype(float)) @ ro["T"].astype(float)
        assert np.linalg.norm(d[:3, 3]) <= 1e-4 and np.arccos(np.clip((np.trace(d[:3, :3]) - 1) / 2, -1, 1)) <= 1e-5
        Tg, To = r["T"], ro["T"]
    assert np.abs(Tg[:3, 3] - big["pose"][:3, 3]).max() < 0.05


def test_replay_batch_sharding_matches_serial(big):
    """configs[3] in miniature: 12 scans against one map; rank r of `world` aligns scans r, r + world, ... and the
    union over ranks equals the serial batch bit for bit (no data-path collective)."""
    scene = big["scene"]
    tgt = np.ascontiguousarray(big["target"][:1_000_000])
    scans, guesses = [], []
    for k in range(12):
        pose = synth.trajectory_pose(20.0 + 6.0 * k)
        scans.append(synth.ouster_scan(scene, pose, beams=32, cols=512, seed=40 + k))
        guesses.append((pose @ synth.pose_matrix((0.1, -0.1, 0.0), np.deg2rad((0.2, 0.1, 0.6)))).astype(np.float32))
    h = ndtb.Handle(0)
    h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
    assert h.set_target(tgt) == 0
    serial = h.align_batch(scans, np.stack(guesses))
    world = 4
    merged = [None] * len(scans)
    for rank in range(world):
        mine = D.local_indices(len(scans), rank, world)
        part = h.align_batch([scans[i] for i in mine], np.stack([guesses[i] for i in mine]))
        for i, r in zip(mine, part):
            merged[i] = r
    for a, b in zip(serial, merged):
        assert np.array_equal(a["T"], b["T"]) and a["iterations"] == b["iterations"] and a["converged"] == b["converged"]
    # the batch keeps two alignments in flight; every one of them must equal the stand-alone ndtb_align bit for bit
    h.set_option("batch_interleave", 0)
    one_by_one = h.align_batch(scans, np.stack(guesses))
    for i, (a, b) in enumerate(zip(serial, one_by_one)):
        h.set_source(scans[i])
        c = h.align(guesses[i], capacity=0)
        for r in (b, c):
            assert np.array_equal(a["T"], r["T"]) and np.array_equal(a["hessian"], r["hessian"])
            assert (a["iterations"], a["converged"], a["evaluations"]) == (r["iterations"], r["converged"], r["evaluations"])
            assert a["trans_probability"] == r["trans_probability"] and a["nvtl"] == r["nvtl"]
    # ragged batch: different scan sizes, identity guess, a batch of one
    h.set_option("batch_interleave", 1)
    ragged = [np.ascontiguousarray(sc[: len(sc) - 101 * k]) for k, sc in enumerate(scans[:5])]
    gg = np.stack(guesses[:5]); gg[2] = np.eye(4, dtype=np.float32)
    rb = h.align_batch(ragged, gg)
    for i, a in enumerate(rb):
        h.set_source(ragged[i])
        c = h.align(gg[i], capacity=0)
        assert np.array_equal(a["T"], c["T"]) and a["iterations"] == c["iterations"]
    assert np.array_equal(h.align_batch(ragged[:1], gg[:1])[0]["T"], rb[0]["T"])
