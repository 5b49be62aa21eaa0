"""pcl::VoxelGrid down-sampling (the preparation step of extern/ndt_omp/apps/align.cpp:65-74 and of the map pipelines):
CPU -- the oracle restatement against an independent numpy evaluation and against the committed real-data fixture's
shape; GPU -- ndtb_voxelgrid_filter bit-exact against the oracle, including the straight-to-source / straight-to-target
variants."""
import os

import numpy as np
import pytest

import oracle_lib as O
from slam_sam_b200 import ndtb, synth


def numpy_voxelgrid(p, leaf):
    p = np.ascontiguousarray(p[:, :3], np.float32)
    p = p[np.isfinite(p).all(axis=1)]
    inv = np.float32(1.0) / np.float32(leaf)
    min_b = np.floor(p.min(axis=0) * inv).astype(np.int32)
    div_b = np.floor(p.max(axis=0) * inv).astype(np.int32) - min_b + 1
    mul = np.array([1, div_b[0], div_b[0] * div_b[1]], np.int64)
    idx = (np.floor(p * inv) - min_b.astype(np.float32)).astype(np.int32).astype(np.int64) @ mul
    order = np.argsort(idx, kind="stable")
    idx_s, ps = idx[order], p[order]
    starts = np.flatnonzero(np.r_[True, idx_s[1:] != idx_s[:-1]])
    ends = np.r_[starts[1:], len(idx_s)]
    out = np.ones((len(starts), 4), np.float32)
    for k, (a, b) in enumerate(zip(starts, ends)):
        s = np.zeros(3, np.float32)
        for q in ps[a:b]:
            s = s + q
        out[k, :3] = s / np.float32(b - a)
    return out


@pytest.fixture(scope="module")
def scan():
    scene = synth.Scene()
    return synth.ouster_scan(scene, synth.trajectory_pose(33.0), beams=32, cols=512, seed=4)


@pytest.mark.parametrize("leaf", [0.1, 0.5, 2.0])
def test_oracle_voxelgrid_matches_numpy(oracle, scan, leaf):
    a = O.voxelgrid_filter(scan, leaf)
    b = numpy_voxelgrid(scan, leaf)
    assert a.shape == b.shape and np.array_equal(a, b)
    assert len(a) < len(scan)


def test_oracle_voxelgrid_edge_cases(oracle, scan):
    one = O.voxelgrid_filter(scan[:1], 0.5)
    assert one.shape == (1, 4) and np.array_equal(one[0, :3], scan[0, :3])
    dirty = scan[:1000].copy(); dirty[::7, 1] = np.nan; dirty[3, 0] = np.inf
    assert np.array_equal(O.voxelgrid_filter(dirty, 0.5), numpy_voxelgrid(dirty, 0.5))
    with pytest.raises(OverflowError):
        O.voxelgrid_filter(scan, 1e-4)                      # dx*dy*dz > INT32_MAX (voxel_grid.hpp: "Leaf size is too small")


@pytest.mark.gpu
@pytest.mark.parametrize("leaf", [0.1, 0.5, 2.0])
def test_gpu_voxelgrid_bit_exact(oracle, scan, leaf):
    h = ndtb.Handle(0)
    got = h.voxelgrid_filter(scan, leaf)
    want = O.voxelgrid_filter(scan, leaf)
    assert got.shape == want.shape and np.array_equal(got, want)


@pytest.mark.gpu
def test_gpu_voxelgrid_into_source_and_target(oracle, scan):
    scene = synth.Scene()
    tgt = synth.surface_map(scene, 300_000, seed=3)
    h = ndtb.Handle(0); h.set_params(resolution=1.0, is_dense=0)
    dirty = tgt.copy(); dirty[::1000, 2] = np.nan           # non-finite points are skipped (is_dense = false)
    m_t = h.voxelgrid_filter(dirty, 0.25, dest="target")
    want_t = O.voxelgrid_filter(dirty, 0.25)
    assert m_t == len(want_t) and np.array_equal(h.target_points(), want_t)
    m_s = h.voxelgrid_filter(scan, 0.2, dest="source")
    want_s = O.voxelgrid_filter(scan, 0.2)
    assert m_s == len(want_s)
    # the down-sampled clouds are what the registration now uses: same result as uploading the oracle's clouds
    h2 = ndtb.Handle(0); h2.set_params(resolution=1.0)
    assert h2.set_target(want_t) == 0 and h2.set_source(want_s) == 0
    guess = synth.trajectory_pose(33.0).astype(np.float32)
    r1, r2 = h.align(guess), h2.align(guess)
    assert np.array_equal(h.trace(), h2.trace()) and np.array_equal(r1["T"], r2["T"])
    with pytest.raises(ndtb.NdtbError):
        h.voxelgrid_filter(scan, 1e-4)                      # NDTB_ERR_OVERFLOW
