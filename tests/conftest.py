import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib

    oracle_lib.build()
    oracle_lib.lib()
    return oracle_lib


@pytest.fixture(scope="session")
def two_plane():
    from slam_sam_b200 import synth

    return synth.two_plane_scene()


@pytest.fixture(scope="session")
def street():
    """Small street-scene scan pair (32 beams x 512 columns) shared by CPU and GPU tests."""
    import numpy as np

    from slam_sam_b200 import synth

    scene = synth.Scene()
    A = synth.trajectory_pose(40.0)
    B = A @ synth.pose_matrix((0.5, 0.1, 0.02), np.deg2rad((0.5, 0.3, 2.0)))
    tgt = synth.ouster_scan(scene, A, beams=32, cols=512, seed=1)
    src = synth.ouster_scan(scene, B, beams=32, cols=512, seed=2)
    return dict(scene=scene, A=A, B=B, target=tgt, source=src, rel=np.linalg.inv(A) @ B)


def rel_inf(a, b):
    import numpy as np

    a = np.asarray(a, float)
    b = np.asarray(b, float)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
