"""More-Thuente step length of the oracle (ndt_omp_impl.hpp:805-1079, NdtParams::use_line_search): properties the
reference's algorithm guarantees, checked on the golden scan pair.  The GPU parity of the same path is in
tests/test_gpu_pclomp.py::test_align_with_line_search."""
import os

import numpy as np

import oracle_lib as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ndt_golden.npz")


def run(step, ls, guess):
    g = dict(np.load(GOLD))
    o = O.OracleNdt(resolution=2.0, max_iterations=35, trans_epsilon=0.01, step_size=step, use_line_search=ls, num_threads=4)
    o.set_target(g["target"]); o.set_source(g["source"])
    return o.align(guess)


def test_line_search_off_matches_golden_path(oracle):
    g = dict(np.load(GOLD))
    r = run(0.1, False, None)
    assert [r["iterations"], int(r["converged"])] == list(g["align_DIRECT7_iters"])
    assert np.array_equal(r["trace"], g["align_DIRECT7_trace"])     # the step-length code path is unchanged without line search


def test_line_search_iterates_and_satisfies_wolfe_like_progress(oracle):
    guess = np.eye(4, dtype=np.float32); guess[0, 3] = 0.8; guess[1, 3] = -0.5
    plain = run(1.0, False, guess)
    ls = run(1.0, True, guess)
    assert ls["converged"] and plain["converged"]
    # the search ran (extra trace rows: one per FP64 Hessian-only pass) and needed no more Newton iterations
    assert len(ls["trace"]) > ls["iterations"] + 1
    assert ls["iterations"] <= plain["iterations"]
    # both end in the same basin on this pair
    assert np.abs(ls["T"][:3, 3] - plain["T"][:3, 3]).max() < 0.01
    # an FP64 Hessian row repeats the pose of the row before it and carries a symmetric, negative-definite-ish Hessian
    tr = ls["trace"]
    dup = [i for i in range(1, len(tr)) if np.array_equal(tr[i, :6], tr[i - 1, :6])]
    assert dup
    for i in dup:
        H = tr[i, 13:49].reshape(6, 6)
        assert np.abs(H - H.T).max() <= 1e-9 * np.abs(H).max()
        assert np.all(np.diag(H) < 0)
