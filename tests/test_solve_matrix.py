"""oakb200_solve_matrix (the job of LRSNash::solve_fast, an absent submodule of the reference): host code, no GPU.
Equilibria are not unique, so the checks are the properties that define one, plus the game value against an
independent LP (scipy) and closed-form games."""
import numpy as np
import pytest

from oak_b200.search import solve_matrix


def _check(a, p1, p2, v, tol=2e-6):
    d = np.floor(a.astype(np.float32) * np.float32(256)).astype(np.float64) / 256.0  # what the solver saw
    assert p1.shape == (a.shape[0],) and p2.shape == (a.shape[1],)
    assert (p1 >= 0).all() and (p2 >= 0).all() and abs(p1.sum() - 1) < 1e-5 and abs(p2.sum() - 1) < 1e-5
    # no pure deviation gains: the row mix guarantees v against every column, the column mix concedes at most v
    assert (p1 @ d).min() >= v - tol and (d @ p2).max() <= v + tol
    return d


def test_closed_form_games():
    p1, p2, v = solve_matrix(np.array([[1, 0], [0, 1]], np.float32))  # matching pennies on {0, 1}
    assert np.allclose(p1, 0.5) and np.allclose(p2, 0.5) and abs(v - 0.5) < 1e-6
    p1, p2, v = solve_matrix(np.array([[0.5, 0.25], [0.75, 0.5]], np.float32))  # saddle point at (1, 1)
    assert np.allclose(p1, [0, 1]) and np.allclose(p2, [0, 1]) and abs(v - 0.5) < 1e-6
    rps = np.array([[0.5, 0, 1], [1, 0.5, 0], [0, 1, 0.5]], np.float32)
    p1, p2, v = solve_matrix(rps)
    assert np.allclose(p1, 1 / 3, atol=1e-6) and np.allclose(p2, 1 / 3, atol=1e-6) and abs(v - 0.5) < 1e-6
    p1, p2, v = solve_matrix(np.full((1, 1), 0.3, np.float32))
    assert p1[0] == 1 and p2[0] == 1 and abs(v - 76 / 256) < 1e-6
    p1, p2, v = solve_matrix(np.zeros((4, 7), np.float32))  # fully degenerate
    _check(np.zeros((4, 7), np.float32), p1, p2, v)
    assert abs(v) < 1e-6


def test_random_games_against_scipy():
    from scipy.optimize import linprog
    rng = np.random.default_rng(12)
    for trial in range(300):
        m, n = int(rng.integers(1, 10)), int(rng.integers(1, 10))
        a = rng.random((m, n)).astype(np.float32)
        if trial % 5 == 0:
            a = np.round(a * 4) / 4  # many ties: degenerate pivots
        p1, p2, v = solve_matrix(a)
        d = _check(a, p1, p2, v)
        # max v s.t. d^T x >= v, sum x = 1
        res = linprog(c=np.r_[np.zeros(m), -1.0], A_ub=np.c_[-d.T, np.ones(n)], b_ub=np.zeros(n), A_eq=np.r_[np.ones(m), 0.0][None, :],
                      b_eq=[1.0], bounds=[(0, None)] * m + [(None, None)])
        assert res.status == 0 and abs(-res.fun - v) < 1e-6, (trial, -res.fun, v)


def test_argument_checks():
    with pytest.raises(ValueError):
        solve_matrix(np.zeros(3, np.float32))
    from oak_b200.engine import OakB200Error
    with pytest.raises(OakB200Error):
        solve_matrix(np.zeros((10, 2), np.float32))


def test_runtime_policy_modes():
    from oak_b200.selfplay import runtime_policy
    emp = np.array([[0.5, 0.3, 0.2, 0, 0, 0, 0, 0, 0]])
    nash = np.array([[0.0, 1.0, 0, 0, 0, 0, 0, 0, 0]])
    assert np.allclose(runtime_policy(emp, nash, "e"), emp)
    assert np.allclose(runtime_policy(emp, nash, "n"), nash)
    assert np.allclose(runtime_policy(emp, nash, "x")[0, :3], [1, 0, 0])
    assert np.allclose(runtime_policy(emp, nash, "e0.5-n0.5")[0, :3], [0.25, 0.65, 0.1])
    p = runtime_policy(emp, nash, "e", temp=2.0)[0, :3]
    assert np.allclose(p, np.array([0.25, 0.09, 0.04]) / 0.38)
    assert np.allclose(runtime_policy(emp, nash, "e", min_prob=0.25)[0, :3], [0.625, 0.375, 0])
    with pytest.raises(RuntimeError):
        runtime_policy(emp, nash, "e", min_prob=0.9)
