"""ndtb_compute_derivatives_static (include/ndtb.h): the derivative evaluation over the library's static __device__ scratch, for
callers that cannot allocate.  No counterpart in the reference; the check is the existing entry point, ndtb_compute_derivatives.

Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_inf
from slam_sam_b200 import ndtb, synth

pytestmark = pytest.mark.gpu
TOL = 1e-9          # the bar of the other GPU parity tests (test_gpu_baseline_configs.py)
STATIC_MAX_POINTS = 4194304


def free_device_bytes():
    import torch

    return torch.cuda.mem_get_info(0)[0]


def assert_same(a, b):
    (sa, ga, Ha, na), (sb, gb, Hb, nb) = a, b
    assert abs(sa - sb) <= TOL * abs(sb)
    assert rel_inf(ga, gb) <= TOL and rel_inf(Ha, Hb) <= TOL
    assert abs(na - nb) <= TOL * abs(nb)


@pytest.mark.parametrize("search", ["DIRECT7", "DIRECT1"])
def test_static_scratch_evaluation_equals_the_allocating_one(street, search):
    import torch

    h = ndtb.Handle(0)
    h.set_params(search_method=getattr(ndtb, search), resolution=1.0)
    assert h.set_target(street["target"]) == 0 and h.set_source(street["source"]) == 0
    poses = [np.zeros(6), np.array([0.4, 0.1, 0.02, 0.005, 0.004, 0.03]), np.array([-0.2, 0.3, 0.0, 0.0, -0.01, -0.02])]
    want = [h.compute_derivatives(p) for p in poses]
    assert abs(want[0][0]) > 1.0                                         # a real evaluation, not zeros
    # (1) the handle's resident source, two calls in a row (the second runs over the first one's leftovers)
    for p, w in zip(poses[:2], want[:2]):
        assert_same(h.compute_derivatives_static(p), w)
    # (2) the caller's own device buffer; a different, shorter cloud first so that stale partial sums would show
    src = np.ascontiguousarray(street["source"], np.float32)
    d_src = torch.from_numpy(src).cuda()
    torch.cuda.synchronize()
    short = h.compute_derivatives_static(poses[1], d_source=d_src.data_ptr(), n_source=1000)
    assert short[0] != want[1][0]
    before = free_device_bytes()
    for p, w in zip(poses, want):
        assert_same(h.compute_derivatives_static(p, d_source=d_src.data_ptr(), n_source=src.shape[0]), w)
    assert free_device_bytes() == before                                 # nothing was allocated (or freed) by the calls
    # without the Hessian: score and gradient only, like the existing entry point
    a = h.compute_derivatives_static(poses[1], compute_hessian=False)
    b = h.compute_derivatives(poses[1], compute_hessian=False)
    assert abs(a[0] - b[0]) <= TOL * abs(b[0]) and rel_inf(a[1], b[1]) <= TOL
    # and the allocating path still agrees afterwards (the static call leaves the handle's own buffers alone)
    assert_same(h.compute_derivatives(poses[2]), want[2])


def test_static_scratch_limits_are_reported(street):
    h = ndtb.Handle(0)
    h.set_params(search_method=ndtb.DIRECT7, resolution=1.0)
    with pytest.raises(ndtb.NdtbError) as e:                             # no target yet
        h.compute_derivatives_static(np.zeros(6))
    assert e.value.code == ndtb.NDTB_ERR_NOT_READY
    assert h.set_target(street["target"]) == 0
    with pytest.raises(ndtb.NdtbError) as e:                             # no source either way
        h.compute_derivatives_static(np.zeros(6))
    assert e.value.code == ndtb.NDTB_ERR_NOT_READY
    with pytest.raises(ndtb.NdtbError) as e:                             # beyond the stated maximum: rejected before any launch
        h.compute_derivatives_static(np.zeros(6), d_source=0x1000, n_source=STATIC_MAX_POINTS + 1)
    assert e.value.code == ndtb.NDTB_ERR_INVALID and "NDTB_STATIC_MAX_POINTS" in str(e.value)
    h.set_params(search_method=ndtb.KDTREE, resolution=1.0)
    assert h.set_source(street["source"]) == 0
    with pytest.raises(ndtb.NdtbError) as e:                             # served search methods only
        h.compute_derivatives_static(np.zeros(6))
    assert e.value.code == ndtb.NDTB_ERR_INVALID
