"""Writes tests/golden/lidar_golden.npz: a small RNG19 packet stream and the frames the ORACLE restatement
(oracle/ouster_decode.hpp) decodes from it.  The reference (src/lidarcallback.cpp) cannot be built here (nlohmann_json and
Eigen are absent), so this file freezes OUR restatement, not output of the reference: parity unpinned.
Run from the repo root:  python tests/golden/make_lidar_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib  # noqa: E402
from test_lidar_decode import ARRAYS, SCALARS, make_config, make_stream  # noqa: E402

stride = 2
cfg, az, alt = make_config(stride=stride, rotated=True)
pk = make_stream(cfg, 3, seed=2024)
frames = oracle_lib.OracleLidar(cfg, az, alt).decode(pk)
out = {"packets": pk, "stride": stride, "n_frames": len(frames)}
for i, f in enumerate(frames):
    for k in ARRAYS:
        out[f"f{i}_{k}"] = f[k]
    out[f"f{i}_scalars"] = np.array([f[k] for k in SCALARS], np.float64)
np.savez_compressed(os.path.join(HERE, "lidar_golden.npz"), **out)
print("frames", len(frames), [f["numberpoints"] for f in frames])
