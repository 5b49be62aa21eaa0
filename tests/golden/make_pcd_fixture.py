"""Builds tests/golden/ndt_omp_pair.npz from the scan pair the reference ships (extern/ndt_omp/data/251370668.pcd,
251371071.pcd) exactly as extern/ndt_omp/apps/align.cpp:49-74 prepares it: load both PCD files, downsample both with
pcl::VoxelGrid at a 0.1 m leaf.  Run in the build container (needs /root/reference); the .npz travels to the GPU box.

The published outputs this fixture pins are the `fitness:` lines of extern/ndt_omp/README.md:19-49 for this very pair
(DIRECT7 0.214205, DIRECT1 0.208511): the ONLY numeric results of the path that the reference publishes.

pcl::VoxelGrid<PointXYZ>::applyFilter is PCL code (>= 1.14, filters/impl/voxel_grid.hpp, not vendored), restated here:
  min_b = floor(min_p * inv_leaf), div_b = max_b - min_b + 1, idx = sum_k int(floor(p_k * inv_leaf_k) - float(min_b_k)) * divb_mul_k,
  points sorted by idx, one output point per occupied leaf in ascending idx = float centroid (CentroidPoint: float32 sums / n).
std::sort's order among equal keys is implementation-defined; a stable order is used here (float-sum rounding only)."""
import os
import sys

import numpy as np

REF = "/root/reference/extern/ndt_omp/data"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ndt_omp_pair.npz")


def load_pcd(path):
    raw = open(path, "rb").read()
    head_end = raw.index(b"DATA binary\n") + len(b"DATA binary\n")
    header = raw[:head_end].decode("ascii", "replace")
    fields = [l for l in header.splitlines() if l.startswith("FIELDS")][0].split()[1:]
    n = int([l for l in header.splitlines() if l.startswith("POINTS")][0].split()[1])
    assert fields == ["x", "y", "z", "intensity"]
    a = np.frombuffer(raw, dtype=np.float32, count=4 * n, offset=head_end).reshape(n, 4)
    return np.ascontiguousarray(a[:, :3])


def voxel_grid_downsample(p, leaf=np.float32(0.1)):
    p = p[np.isfinite(p).all(axis=1)]
    inv = np.float32(1.0) / np.float32(leaf)
    mn, mx = p.min(axis=0), p.max(axis=0)
    min_b = np.floor(mn * inv).astype(np.int32)
    max_b = np.floor(mx * inv).astype(np.int32)
    div_b = max_b - min_b + 1
    mul = np.array([1, div_b[0], div_b[0] * div_b[1]], np.int64)
    ijk = (np.floor(p * inv) - min_b.astype(np.float32)).astype(np.int32)
    idx = ijk.astype(np.int64) @ mul
    order = np.argsort(idx, kind="stable")
    idx_s, ps = idx[order], p[order]
    starts = np.flatnonzero(np.r_[True, idx_s[1:] != idx_s[:-1]])
    ends = np.r_[starts[1:], len(idx_s)]
    out = np.zeros((len(starts), 3), np.float32)
    for k, (a, b) in enumerate(zip(starts, ends)):
        s = np.zeros(3, np.float32)
        for q in ps[a:b]:
            s = s + q                      # float32 accumulation in sorted order
        out[k] = s / np.float32(b - a)
    return out


if __name__ == "__main__":
    tgt = voxel_grid_downsample(load_pcd(os.path.join(REF, "251370668.pcd")))
    src = voxel_grid_downsample(load_pcd(os.path.join(REF, "251371071.pcd")))
    np.savez_compressed(OUT, target=tgt, source=src, readme_fitness_direct7=0.214205, readme_fitness_direct1=0.208511)
    print("target", tgt.shape, "source", src.shape, "->", OUT, os.path.getsize(OUT), "bytes")
