"""Multi-GPU SVN (needs >= 2 CUDA devices; skipped on a one-GPU box): the device-resident sharded align -- records peer-written
over NVLink by the derivative kernel's last blocks, no collective call -- must equal the single-GPU align bit for bit."""
import threading

import numpy as np
import pytest

from slam_sam_b200 import ndtb, synth

pytestmark = pytest.mark.gpu


def n_devices():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def make_handle(dev, tgt, src, K):
    h = ndtb.Handle(dev)
    h.set_params(grid_flavour=ndtb.FLAVOUR_SVN, resolution=1.0, min_points_per_voxel=3, search_method=ndtb.DIRECT7)
    h.set_svn_params(particle_count=K, max_iterations=40, kernel_bandwidth=1.0, step_size=1.0, stop_threshold=1e-4, use_gauss_newton=1)
    assert h.set_target(tgt) == 0 and h.set_source(src) == 0
    return h


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_align_equals_single_gpu(street, world):
    if n_devices() < world:
        pytest.skip(f"needs {world} GPUs")
    K = 20
    rng = np.random.Generator(np.random.PCG64(7))
    P0 = np.stack([synth.pose_matrix(rng.normal(0, 0.05, 3), rng.normal(0, 0.01, 3)) for _ in range(K)])
    prior = np.eye(4)
    single = make_handle(0, street["target"], street["source"], K)
    r1 = single.svn_align(prior, P0)
    hs = [make_handle(d, street["target"], street["source"], K) for d in range(world)]
    ndtb.comm_init(hs)
    assert hs[1].comm_info() == (1, world)
    for rep in range(2):      # twice: stamps and window halves carry over from one collective align to the next
        out = [None] * world
        err = []

        def run(r):
            try:
                out[r] = hs[r].svn_align(prior, P0)
            except Exception as e:   # noqa: BLE001
                err.append(e)

        th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
        [t.start() for t in th]
        [t.join() for t in th]
        assert not err, err
        for r in range(world):
            assert out[r]["iterations"] == r1["iterations"] and out[r]["converged"] == r1["converged"]
            assert np.array_equal(out[r]["particles"], r1["particles"])
            assert np.array_equal(out[r]["final_pose"], r1["final_pose"])
            assert np.array_equal(out[r]["final_covariance"], r1["final_covariance"])
    for h in hs:
        h.comm_close()
    r2 = hs[0].svn_align(prior, P0)     # back to single-GPU operation
    assert np.array_equal(r2["particles"], r1["particles"])


def test_comm_example_cpp(street, tmp_path):
    """tests/cpp/comm_example.cpp: one C++ process, one registrar per GPU, ndtb_comm_init + threaded collective align -- every
    rank prints the same line as a one-GPU run of the same program."""
    import os
    import shutil
    import subprocess

    if n_devices() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.join(root, "slam_sam_b200")
    exe = os.path.join(root, "tests", "cpp", "comm_example")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    r = subprocess.run([gxx, "-std=c++17", "-O2", "-Wall", "-pthread", "-I" + os.path.join(root, "include"), "-DNDTB_FACADE_NO_EIGEN",
                        os.path.join(root, "tests", "cpp", "comm_example.cpp"), "-L" + lib_dir, "-lndtb", "-Wl,-rpath," + lib_dir, "-o", exe],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    t = tmp_path / "t.bin"; s = tmp_path / "s.bin"
    np.ascontiguousarray(street["target"], np.float32).tofile(t); np.ascontiguousarray(street["source"], np.float32).tofile(s)
    one = subprocess.run([exe, "1", str(t), str(s)], capture_output=True, text=True)
    assert one.returncode == 0, one.stderr
    ref = one.stdout.split()[1:]
    two = subprocess.run([exe, "2", str(t), str(s)], capture_output=True, text=True, timeout=120)
    assert two.returncode == 0, two.stderr
    lines = two.stdout.strip().splitlines()
    assert len(lines) == 2 and all(line.split()[1:] == ref for line in lines)
