"""The Newton loop's plumbing on the GPU: ndtb_align serves every evaluation from one resident kernel (an
"alignment session") by default.  It must give bit-identical traces to one launch per evaluation, survive its own
idle time-out, and the pivoted-elimination Newton direction must stay inside the parity bar of the Jacobi SVD."""
import os

import numpy as np
import pytest

from conftest import rel_inf
from slam_sam_b200 import ndtb

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ndt_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


def make_handle(gold, search=ndtb.DIRECT7):
    h = ndtb.Handle(0)
    h.set_params(resolution=2.0, search_method=search, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
    h.set_target(gold["target"]); h.set_source(gold["source"])
    return h


@pytest.mark.parametrize("search", ["DIRECT7", "DIRECT1"])
def test_session_is_bit_identical_to_one_launch_per_evaluation(gold, search):
    h = make_handle(gold, getattr(ndtb, search))
    n0 = h.launch_count()
    r_s = h.align(); tr_s = h.trace()
    session_launches = h.launch_count() - n0
    assert session_launches == 1                      # one resident kernel for the whole Newton loop
    assert h.timings()["eval_ms_total"] > 0           # device-side (globaltimer) evaluation time is reported
    h.set_option("sessions", 0)
    n0 = h.launch_count()
    r_l = h.align(); tr_l = h.trace()
    assert h.launch_count() - n0 == r_l["evaluations"]
    h.set_option("sessions", 1); h.set_profiling(True)  # per-kernel events imply one launch per evaluation
    n0 = h.launch_count()
    r_p = h.align(); tr_p = h.trace()
    assert h.launch_count() - n0 == r_p["evaluations"]
    assert np.array_equal(tr_s, tr_l) and np.array_equal(tr_s, tr_p)
    assert np.array_equal(r_s["T"], r_l["T"]) and r_s["iterations"] == r_l["iterations"] == r_p["iterations"]


def test_session_idle_timeout_falls_back_to_launches(gold):
    h = make_handle(gold)
    ref = h.align(); tr_ref = h.trace()
    h.set_option("session_idle_timeout_ms", 1e-4)     # 100 ns: the resident kernel leaves before a command can arrive
    for _ in range(3):
        r = h.align()
        assert np.array_equal(h.trace(), tr_ref) and np.array_equal(r["T"], ref["T"])
    h.set_option("session_idle_timeout_ms", 50)
    r = h.align()
    assert np.array_equal(h.trace(), tr_ref)


def test_sessions_back_to_back_with_new_sources(gold):
    h = make_handle(gold)
    rng = np.random.default_rng(3)
    src = gold["source"]
    traces = []
    for k in range(4):
        sub = np.ascontiguousarray(src[rng.permutation(len(src))[: len(src) - 37 * k]])
        h.set_source(sub)
        h.set_option("sessions", 1); h.align(); a = h.trace()
        h.set_option("sessions", 0); h.align(); b = h.trace()
        assert np.array_equal(a, b)
        traces.append(a)
    assert len(traces) == 4


def test_direct_solver_stays_inside_the_parity_bar(gold):
    h = make_handle(gold)
    h.set_option("direct_solver", 0)
    r0 = h.align(); t0 = h.trace()
    h.set_option("direct_solver", 1)
    r1 = h.align(); t1 = h.trace()
    assert t0.shape == t1.shape and r0["iterations"] == r1["iterations"]
    for a, b in zip(t1, t0):
        assert np.abs(a[:6] - b[:6]).max() <= 1e-11
        assert abs(a[6] - b[6]) <= 1e-9 * abs(b[6]) and rel_inf(a[7:13], b[7:13]) <= 1e-9 and rel_inf(a[13:49], b[13:49]) <= 1e-9
    assert np.abs(r0["T"] - r1["T"]).max() <= 1e-6


def test_session_can_leave_sms_free_for_other_streams(street):
    """ "session_free_sms": the one-CTA-per-SM evaluation kernels run on fewer SMs, so another stream's kernel (here a second
    handle's map build) makes progress while an align is open; results stay inside the parity bar (other partial sums)."""
    import torch

    assert torch.cuda.get_device_properties(0).multi_processor_count == 148
    def handle(free):
        h = ndtb.Handle(0)
        h.set_params(resolution=1.0, search_method=ndtb.DIRECT7, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
        h.set_option("session_free_sms", free)
        assert h.set_target(street["target"]) == 0 and h.set_source(street["source"]) == 0
        return h
    full, part = handle(0), handle(16)
    ra, rb = full.align(), part.align()
    ta, tb = full.trace(), part.trace()
    assert ra["iterations"] == rb["iterations"] and ta.shape == tb.shape
    assert rel_inf(tb[:, 6:49], ta[:, 6:49]) <= 1e-9 and np.abs(ra["T"] - rb["T"]).max() <= 1e-6
    # session and one launch per evaluation still agree bit for bit on the reduced grid
    part.set_option("sessions", 0)
    rc = part.align()
    assert np.array_equal(part.trace(), tb) and np.array_equal(rc["T"], rb["T"])
    with pytest.raises(ndtb.NdtbError):
        part.set_option("session_free_sms", 500)


def test_device_newton_step_against_the_host_step():
    """device_newton.cuh, the experiment behind DESIGN 4.4: a warp-cooperative Newton step on the device reproduces the host step
    of ndtb_align BIT FOR BIT in everything that is IEEE arithmetic -- direction (pivoted elimination + refinement), step length,
    next pose, the decision to continue / stop / hand over to the host -- but not in the float transform and the angle tables:
    those go through sin / cos, and CUDA's routines do not round like glibc's sinf / sin (1 ulp apart in a few per cent of the
    entries).  That is why the step stays on the host: a 1-ulp transform moves every evaluation by ~1e-7 relative, outside the
    1e-9 per-evaluation bar against the oracle."""
    rng = np.random.default_rng(11)
    n_cont = 0
    n_entries = n_off = 0
    for case in range(300):
        A = rng.normal(size=(6, 6)) * np.array([1, 1, 1, 8, 8, 8])
        H = -(A @ A.T) * rng.uniform(10, 1e4)                             # NDT Hessians near an optimum are negative definite
        if case % 7 == 0:
            H = H + rng.normal(size=(6, 6)) * 1e-3                        # not exactly symmetric
        g = rng.normal(size=6) * rng.uniform(1e-3, 1e3)
        if case % 11 == 3:
            g = -g
        p = np.concatenate([rng.uniform(-50, 50, 3), rng.uniform(-3.1, 3.1, 3)])
        if case % 13 == 5:
            p[3:] = rng.uniform(-2e-4, 2e-4, 3)                           # around the 1e-4 snap of the angle tables
        first = case % 5 == 0
        nr = int(rng.integers(0, 40)); a_prev = float(rng.choice([0.1, 0.02, 0.004, 0.0009]))
        args = (g, H, p, 0.1, 0.005, 0.01, 35, first, nr, a_prev)
        d, h = ndtb.test_newton_step(0, *args), ndtb.test_newton_step(1, *args)
        assert d[0] == h[0] and d[1] == h[1], case
        if h[0] == 0:
            n_cont += 1
            assert d[2] == h[2] and np.array_equal(d[4], h[4]), case     # step length and next pose: exact
            du, hu = d[3].view(np.int32), h[3].view(np.int32)
            assert np.abs(d[3].astype(np.float64) - h[3].astype(np.float64)).max() <= 4e-7, case   # entries of rotations and tables (|.| <= 1): a few float ulps
            assert np.array_equal(d[3][[3, 7, 11]], h[3][[3, 7, 11]])                           # the translation column is float(p): exact
            n_entries += du.size; n_off += int((du != hu).sum())
    assert n_cont > 150
    assert 0 < n_off < 0.1 * n_entries                                    # libm differences exist, and they are rare
    # the cases the host has to decide: singular H, zero gradient, NaN
    Hs = -np.outer(np.arange(1, 7.0), np.arange(1, 7.0))
    for g, H in ((np.ones(6), Hs), (np.zeros(6), -np.eye(6)), (np.full(6, np.nan), -np.eye(6))):
        d = ndtb.test_newton_step(0, g, H, np.zeros(6), 0.1, 0.005, 0.01, 35, True, 0, 0.0)
        h = ndtb.test_newton_step(1, g, H, np.zeros(6), 0.1, 0.005, 0.01, 35, True, 0, 0.0)
        assert d[0] == h[0] == 2


def test_two_handles_from_two_threads(street):
    """Two registrars driven from two host threads at once (slam-sam runs the odometry and the loop-closure registration in separate
    threads): the cooperative sort kernel and the resident session kernel of one handle must not dead-lock the other's -- each
    needs the whole GPU, the launches serialise -- and every result equals the single-threaded one."""
    import threading

    def work(out, reps):
        h = ndtb.Handle(0)
        h.set_params(resolution=1.0, search_method=ndtb.DIRECT7, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
        for _ in range(reps):
            assert h.set_target(street["target"]) == 0 and h.set_source(street["source"]) == 0
            out.append(h.align()["T"])

    ref = []
    work(ref, 1)
    outs = [[], []]
    ts = [threading.Thread(target=work, args=(outs[i], 12)) for i in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
        assert not t.is_alive(), "dead-lock between two handles"
    for o in outs:
        assert len(o) == 12 and all(np.array_equal(T, ref[0]) for T in o)
