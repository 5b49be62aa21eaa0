"""The C-ABI library loads and exports every symbol include/ndtb.h declares (CPU: no compute calls)."""
import ctypes as C
import os
import re

import pytest

from slam_sam_b200 import ndtb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "ndtb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ndtb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert header_symbols() == sorted(ndtb.SYMBOLS)


def test_library_exports_every_symbol():
    assert os.path.exists(ndtb.LIB_PATH), "libndtb.so missing: run __graft_entry__.build()"
    lib = C.CDLL(ndtb.LIB_PATH)
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_struct_layouts_match_header():
    # sizes the C compiler gives the header's structs (x86-64 SysV)
    assert C.sizeof(ndtb.Params) == 72
    assert C.sizeof(ndtb.Result) == 64 + 288 + 16 + 16
    assert C.sizeof(ndtb.SvnParams) == 40
    assert C.sizeof(ndtb.SvnResult) == 128 + 288 + 8


def test_defaults_mirror_reference_constructors():
    L = ndtb.lib()
    p = ndtb.Params()
    assert L.ndtb_default_params(ndtb.FLAVOUR_PCLOMP, C.byref(p)) == 0
    # ndt_omp_impl.hpp:52-81, voxel_grid_covariance_omp.h:193-208
    assert (p.resolution, p.step_size, p.trans_epsilon, p.max_iterations, p.search_method) == (1.0, 0.1, 0.1, 35, ndtb.DIRECT7)
    assert (p.outlier_ratio, p.min_points_per_voxel, p.min_covar_eigvalue_mult, p.use_line_search) == (0.55, 6, 0.01, 0)
    s = ndtb.SvnParams()
    assert L.ndtb_svn_default_params(C.byref(s)) == 0
    # svn_ndt_impl.hpp:61-77
    assert (s.particle_count, s.max_iterations, s.kernel_bandwidth, s.step_size, s.stop_threshold, s.use_gauss_newton) == (30, 50, 1.0, 1.0, 1e-4, 1)
    assert L.ndtb_version().startswith(b"ndtb")


def test_null_handles_are_rejected():
    L = ndtb.lib()
    assert L.ndtb_destroy(None) == -1
    assert L.ndtb_set_params(None, None) == -1
    assert L.ndtb_get_voxel_count(None, None, None, None, None) == -1


def test_no_cpu_fallback_without_gpu():
    """Without a CUDA device the product must fail loudly, not fall back to the oracle."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(ndtb.NdtbError):
        ndtb.Handle(0)
    src = open(os.path.join(ROOT, "slam_sam_b200", "ndtb.py")).read() + open(os.path.join(ROOT, "slam_sam_b200", "registration.py")).read()
    assert "oracle" not in src.replace("oracle)", "").lower() or "import oracle" not in src
    for f in os.listdir(os.path.join(ROOT, "slam_sam_b200", "csrc")):
        if f.endswith((".cu", ".cuh")):
            assert "oracle/" not in open(os.path.join(ROOT, "slam_sam_b200", "csrc", f)).read(), f


def test_sincosf_restatement_matches_the_running_libm():
    """The device Newton step of the resident align kernel restates glibc's sincosf (newton_math.cuh, nm_sincosf_glibc); the
    library checks the restatement against the libm it runs with (24 000 arguments) and only then lets the device take steps.
    Host code, no GPU: on this image (glibc 2.39, FMA build selected) it must hold."""
    import ctypes as C

    from slam_sam_b200 import ndtb
    st = (C.c_ulonglong * 4)()
    assert ndtb.lib().ndtb_get_session_stats(None, st) == 0
    assert list(st) == [0, 0, 0, 1]
