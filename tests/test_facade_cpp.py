"""include/ndtb_pclomp.hpp: the C++ facade with the reference's class / method names compiles and links against
libndtb.so (CPU), and on a GPU box reproduces the Python binding's alignment (same C ABI underneath)."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "facade_example")


def build_example():
    lib_dir = os.path.join(ROOT, "slam_sam_b200")
    if not os.path.exists(os.path.join(lib_dir, "libndtb.so")):
        pytest.skip("libndtb.so not built")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    cmd = [gxx, "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "include"), "-DNDTB_FACADE_NO_EIGEN",
           os.path.join(ROOT, "tests", "cpp", "facade_example.cpp"), "-L" + lib_dir, "-lndtb", "-Wl,-rpath," + lib_dir, "-o", EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return EXE


def test_facade_compiles_and_links():
    exe = build_example()
    assert os.path.exists(exe)
    # without arguments it only prints its usage (no CUDA call is made)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


@pytest.mark.gpu
def test_facade_matches_python_binding(two_plane, tmp_path):
    from slam_sam_b200 import registration

    exe = build_example()
    t = tmp_path / "target.bin"; s = tmp_path / "source.bin"
    np.ascontiguousarray(two_plane["target"], np.float32).tofile(t)
    np.ascontiguousarray(two_plane["source"], np.float32).tofile(s)
    guess = two_plane["guess"].astype(np.float32)
    r = subprocess.run([exe, str(t), str(s)] + [repr(float(v)) for v in guess.reshape(16)] + ["svn"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    f = lines[0].split()
    ndt = registration.NormalDistributionsTransform()
    ndt.setResolution(1.0); ndt.setMaximumIterations(50); ndt.setTransformationEpsilon(1e-4); ndt.setStepSize(0.1)
    ndt.setInputTarget(two_plane["target"]); ndt.setInputSource(two_plane["source"])
    ndt.computeTransformation(guess)
    assert int(f[0]) == int(ndt.hasConverged()) == 1
    assert int(f[1]) == ndt.getFinalNumIteration()
    T = np.array([float(v) for v in f[2:18]]).reshape(4, 4)
    assert np.abs(T - ndt.getFinalTransformation()).max() <= 1e-6
    assert int(f[18]) == len(two_plane["source"]) and int(f[19]) == len(ndt.getTargetCells()["idx"])
    assert lines[1].startswith("svn ") and int(lines[1].split()[2]) == 5
    w = lines[2].split()   # device-assembled window of one identity-pose keyframe == the uploaded target
    assert w[0] == "win" and w[1:3] == f[0:2] and w[3:19] == f[2:18]
    # multigrid facade: two tiles split at x = 0, a third added and removed again
    m = lines[3].split()
    assert m[0] == "mg" and int(m[1]) == 1
    mg = registration.MultiGridNormalDistributionsTransform()
    mg.setResolution(1.0); mg.setMaximumIterations(50); mg.setTransformationEpsilon(1e-4); mg.setStepSize(0.1)
    tgt = np.ascontiguousarray(two_plane["target"], np.float32)
    mg.addTarget(np.ascontiguousarray(tgt[tgt[:, 0] < 0]), "west"); mg.addTarget(np.ascontiguousarray(tgt[tgt[:, 0] >= 0]), "east")
    mg.createVoxelKdtree(); mg.setInputSource(two_plane["source"])
    mg.align(guess)
    assert int(m[2]) == mg.getFinalNumIteration()
    assert np.abs(np.array([float(v) for v in m[3:19]]).reshape(4, 4) - mg.getFinalTransformation()).max() <= 1e-6
    assert int(m[19]) == 2 and int(m[20]) == len(mg.getVoxelPCD())
    assert abs(float(m[21]) - mg.calculateNearestVoxelTransformationLikelihood(mg.getFinalTransformation())) <= 1e-6


ADAPTER_EXE = os.path.join(ROOT, "tests", "cpp", "adapter_example")


def build_adapter_example():
    """include/ndtb_pcl_adapter.hpp against tests/cpp/mock_pcl (a stand-in for the PCL declarations the adapter uses)."""
    lib_dir = os.path.join(ROOT, "slam_sam_b200")
    if not os.path.exists(os.path.join(lib_dir, "libndtb.so")):
        pytest.skip("libndtb.so not built")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    cmd = [gxx, "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "tests", "cpp", "mock_pcl"),
           "-DNDTB_FACADE_NO_EIGEN", os.path.join(ROOT, "tests", "cpp", "adapter_example.cpp"), "-L" + lib_dir, "-lndtb",
           "-Wl,-rpath," + lib_dir, "-o", ADAPTER_EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return ADAPTER_EXE


def test_pcl_adapter_compiles_and_links():
    """The pcl::Registration-derived drop-in: RegisterCallback::registration keeps its base-pointer type (registercallback.hpp:35)."""
    exe = build_adapter_example()
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


@pytest.mark.gpu
def test_pcl_adapter_through_the_base_pointer(two_plane, tmp_path):
    """setInputTarget / setInputSource / align / getFinalTransformation through pcl::Registration::Ptr (run/pipeline.cpp:557-566),
    PointXYZI clouds (32-byte stride): same alignment as the Python binding, `output` = transformed source with intensity kept,
    and align() returns promptly although `output` is requested (the resident kernel is told to leave before the copy)."""
    import time

    from slam_sam_b200 import registration

    exe = build_adapter_example()
    t = tmp_path / "target.bin"; s = tmp_path / "source.bin"
    np.ascontiguousarray(two_plane["target"], np.float32).tofile(t)
    np.ascontiguousarray(two_plane["source"], np.float32).tofile(s)
    guess = two_plane["guess"].astype(np.float32)
    r = subprocess.run([exe, str(t), str(s)] + [repr(float(v)) for v in guess.reshape(16)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    f = r.stdout.strip().split()
    ndt = registration.NormalDistributionsTransform()
    ndt.setResolution(1.0); ndt.setMaximumIterations(50); ndt.setTransformationEpsilon(1e-4); ndt.setStepSize(0.1)
    ndt.setInputTarget(two_plane["target"]); ndt.setInputSource(two_plane["source"])
    ndt.computeTransformation(guess)
    assert int(f[0]) == 1 and int(f[1]) == ndt.getFinalNumIteration()
    T = np.array([float(v) for v in f[2:18]]).reshape(4, 4)
    assert np.abs(T - ndt.getFinalTransformation()).max() <= 1e-6
    assert int(f[18]) == len(two_plane["source"])
    p0 = T[:3, :3] @ two_plane["source"][0, :3].astype(np.float64) + T[:3, 3]
    assert np.abs(np.array([float(v) for v in f[19:22]]) - p0).max() <= 1e-4 and float(f[22]) == 42.0
    # with out_xyz set the D2H copy used to queue behind the resident session kernel until its 50 ms idle time-out
    h = ndt._h
    n = len(two_plane["source"])
    out = np.zeros((n, 3), np.float32)
    import ctypes

    from slam_sam_b200 import ndtb

    res = ndtb.Result()
    g = np.ascontiguousarray(guess.reshape(16))
    for _ in range(2):
        t0 = time.perf_counter()
        rc = h.L.ndtb_align(h.h, g.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), ctypes.byref(res), None, 0, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
        dt = time.perf_counter() - t0
        assert rc == 0
    assert dt < 0.025, f"align with out_xyz took {dt * 1e3:.1f} ms: queued behind the resident kernel?"
    assert np.abs(out[0] - p0).max() <= 1e-4
    # every point, bit for bit: T[4r] x + (T[4r+1] y + (T[4r+2] z + T[4r+3])) in float, two roundings per term (the device kernel
    # restates the host loop it replaced)
    Tf = np.array(res.pose, np.float32).reshape(4, 4)
    s3 = np.ascontiguousarray(two_plane["source"][:, :3], np.float32)
    want = np.stack([Tf[r, 0] * s3[:, 0] + (Tf[r, 1] * s3[:, 1] + (Tf[r, 2] * s3[:, 2] + Tf[r, 3])) for r in range(3)], axis=1).astype(np.float32)
    assert np.array_equal(out, want)


LIDAR_EXE = os.path.join(ROOT, "tests", "cpp", "lidar_example")


def build_lidar_example():
    lib_dir = os.path.join(ROOT, "slam_sam_b200")
    if not os.path.exists(os.path.join(lib_dir, "libndtb.so")):
        pytest.skip("libndtb.so not built")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else shutil.which("g++")
    cmd = [gxx, "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "lidar_example.cpp"),
           "-L" + lib_dir, "-lndtb", "-Wl,-rpath," + lib_dir, "-o", LIDAR_EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return LIDAR_EXE


def test_lidar_facade_compiles_and_links():
    """include/ndtb_lidar.hpp (LidarCallback / LidarFrame with the reference's names) builds against libndtb.so."""
    exe = build_lidar_example()
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


@pytest.mark.gpu
def test_lidar_facade_matches_oracle(oracle, tmp_path):
    from test_lidar_decode import make_config, make_stream

    exe = build_lidar_example()
    cfg, az, alt = make_config(cols=128, pixels=64, stride=2)
    pk = make_stream(cfg, 3, seed=77)
    path = tmp_path / "packets.bin"
    pk.tofile(path)
    r = subprocess.run([exe, str(path), "128", "64", "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    want = oracle.OracleLidar(cfg, az, alt).decode(pk)
    lines = r.stdout.strip().splitlines()
    assert len(lines) == len(want) == 3
    for line, w in zip(lines, want):
        f = line.split()
        assert int(f[0]) == w["frame_id"] and int(f[1]) == w["numberpoints"]
        assert float(f[2]) == w["timestamp"] and float(f[3]) == w["timestamp_end"] and float(f[4]) == w["interframe_timedelta"]
        assert float(f[5]) == float(np.sum(w["x"].astype(np.float64))) or abs(float(f[5]) - np.sum(w["x"].astype(np.float64))) < 1e-6
        assert int(f[6]) == int(w["c_id"].astype(np.int64).sum()) and int(f[7]) == int(w["signal"].astype(np.int64).sum())
