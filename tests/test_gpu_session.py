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
    """device_newton.cuh: a warp-cooperative Newton step on the device reproduces the host step of ndtb_align BIT FOR BIT --
    direction (pivoted elimination + refinement), step length, next pose, the decision to continue / stop / hand over to the
    host, and the float transform, whose sin / cos restate glibc's sincosf operation by operation (newton_math.cuh).  Only the
    angle tables go through CUDA's FP64 sin / cos, an ulp from glibc's for some arguments: that survives the rounding to float in
    ~1e-8 of the entries, so none is expected here -- and in an align the host compares every step anyway."""
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
            assert np.array_equal(du[:12], hu[:12]), case                # T: the sincosf restatement lands on glibc's bits
            assert np.abs(d[3].astype(np.float64) - h[3].astype(np.float64)).max() <= 2e-7, case   # table entries (|.| <= 1): at most a float ulp
            n_entries += du.size; n_off += int((du != hu).sum())
    assert n_cont > 150
    assert n_off == 0                                                     # 24 300 entries at ~1e-8 each
    # the cases the host has to decide: singular H, zero gradient, NaN
    Hs = -np.outer(np.arange(1, 7.0), np.arange(1, 7.0))
    for g, H in ((np.ones(6), Hs), (np.zeros(6), -np.eye(6)), (np.full(6, np.nan), -np.eye(6))):
        d = ndtb.test_newton_step(0, g, H, np.zeros(6), 0.1, 0.005, 0.01, 35, True, 0, 0.0)
        h = ndtb.test_newton_step(1, g, H, np.zeros(6), 0.1, 0.005, 0.01, 35, True, 0, 0.0)
        assert d[0] == h[0] == 2


def test_device_newton_in_the_resident_kernel_is_invisible(gold, street):
    """ "device_newton" (opt-in): the resident kernel takes the Newton steps itself and the host only confirms them -- same
    traces, poses and iteration counts, bit for bit, as with every step on the host, still one launch per align."""
    for data, res in ((gold, 2.0), (street, 1.0)):
        for search in (ndtb.DIRECT7, ndtb.DIRECT1):
            h = ndtb.Handle(0)
            h.set_params(resolution=res, search_method=search, max_iterations=35, trans_epsilon=0.01, step_size=0.1)
            h.set_target(data["target"]); h.set_source(data["source"])
            assert h.session_stats()["libm_port_ok"]
            h.set_option("device_newton", 1)
            s0 = h.session_stats(); n0 = h.launch_count()
            r_dev = h.align(); t_dev = h.trace()
            s1 = h.session_stats()
            assert h.launch_count() - n0 == 1
            assert s1["mismatches"] == s0["mismatches"]
            # every evaluation but the first was started by the device, except where it handed the step to the host (the SVD case,
            # or flow control: a host thread that was descheduled has not yet read the results the next evaluation would overwrite)
            assert (s1["accepted"] - s0["accepted"]) + (s1["declined"] - s0["declined"]) >= r_dev["evaluations"] - 1
            assert s1["accepted"] - s0["accepted"] >= 1
            h.set_option("device_newton", 0)
            r_host = h.align(); t_host = h.trace()
            assert h.session_stats()["accepted"] == s1["accepted"]
            assert np.array_equal(t_dev, t_host) and np.array_equal(r_dev["T"], r_host["T"])
            assert r_dev["iterations"] == r_host["iterations"] and r_dev["evaluations"] == r_host["evaluations"]
            h.set_option("device_newton", 1)
            for _ in range(3):                                            # back to back: sequence numbers and keys carry over
                r = h.align()
                assert np.array_equal(h.trace(), t_host) and np.array_equal(r["T"], r_host["T"])
            assert h.session_stats()["mismatches"] == s0["mismatches"]


def test_device_newton_disagreement_is_repaired_by_the_host(gold):
    """The host compares the command the device built for the next evaluation with its own, word for word.  A forced disagreement
    at the n-th comparison ("device_newton_fault_at"): the host lets the resident kernel run out, repeats that evaluation from its
    own state with one launch per evaluation, and the align returns the same bits."""
    h = make_handle(gold)
    h.set_option("device_newton", 0)
    ref = h.align(); tr_ref = h.trace()
    h.set_option("device_newton", 1)
    for fault_at in (1, 2, 3):
        h.set_option("device_newton_fault_at", fault_at)
        m0 = h.session_stats()["mismatches"]
        r = h.align()
        assert h.session_stats()["mismatches"] == m0 + 1
        assert np.array_equal(h.trace(), tr_ref) and np.array_equal(r["T"], ref["T"]) and r["iterations"] == ref["iterations"]
    h.set_option("device_newton_fault_at", 0)
    m0 = h.session_stats()
    r = h.align()
    m1 = h.session_stats()
    assert m1["mismatches"] == m0["mismatches"] and m1["accepted"] > m0["accepted"]
    assert np.array_equal(h.trace(), tr_ref) and np.array_equal(r["T"], ref["T"])


def test_device_newton_steps_aside(gold):
    """Steps the device does not take: with a regularisation pose (the host adds its terms to g and H), with the SVD forced
    ("direct_solver" 0) and with the line search, no command allows speculation -- and results equal the launch path's."""
    h = make_handle(gold)
    h.set_option("device_newton", 1)
    h.set_option("direct_solver", 0)
    a0 = h.session_stats()["accepted"]
    r1 = h.align(); t1 = h.trace()
    assert h.session_stats()["accepted"] == a0
    h.set_option("sessions", 0)
    r2 = h.align()
    assert np.array_equal(h.trace(), t1) and np.array_equal(r1["T"], r2["T"])
    h.set_option("sessions", 1); h.set_option("direct_solver", 1)
    h.set_params(resolution=2.0, search_method=ndtb.DIRECT7, max_iterations=35, trans_epsilon=0.01, step_size=0.1, regularization_scale_factor=0.001)
    h.set_regularization_pose(np.eye(4, dtype=np.float32))
    r3 = h.align(); t3 = h.trace()
    assert h.session_stats()["accepted"] == a0
    h.set_option("sessions", 0)
    r4 = h.align()
    assert np.array_equal(h.trace(), t3) and np.array_equal(r3["T"], r4["T"])


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
