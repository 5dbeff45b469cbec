"""Pins for the quintic-spline restatement (oracle/spline.py): the reference's own known-answer tests replayed
(src/Math/tests/SplineTest.cpp:21-167, bindings/python/Math/tests/test_quintic_spline.py), plus the properties that
define the spline (C4 at the interior knots, boundary conditions, clamping outside the knots)."""
import numpy as np
import pytest

from oracle.spline import cubic_spline_coefficients, quintic_spline_coefficients, spline_evaluate


def _poly(c, t, k=0):
    return sum(c[i] * np.prod(np.arange(i, i - k, -1)) * t ** (i - k) for i in range(k, 6))


@pytest.mark.parametrize("t0,t1,checks", [(0.32, 642.0, 1000), (0.32, 2.64, 10000)])
def test_spline_through_samples_of_a_quintic_is_that_quintic(t0, t1, checks):
    """SplineTest.cpp "Quintic spline": 100 knots on a random quintic in R^4 between 320 ms and 642 s (isApprox 1e-5);
    test_quintic_spline.py: the same between 0.32 s and 2.64 s (atol 1e-5)."""
    rng = np.random.default_rng(3)
    c = rng.uniform(-1.0, 1.0, (6, 4))
    times = t0 + (t1 - t0) / 99 * np.arange(100)
    x = np.stack([_poly(c, t) for t in times])
    co, v, a = quintic_spline_coefficients(times, x, _poly(c, t0, 1), _poly(c, t0, 2), _poly(c, times[-1], 1), _poly(c, times[-1], 2))
    for i in range(0, checks, 7):
        t = t0 + (times[-1] - t0) / checks * i
        pos, vel, acc = spline_evaluate(times, co, t)
        for got, k in ((pos, 0), (vel, 1), (acc, 2)):
            want = _poly(c, t, k)
            assert np.linalg.norm(got - want) <= 1e-5 * min(np.linalg.norm(got), np.linalg.norm(want)) + (1e-5 if t1 < 10 else 0.0)


def test_interior_knots_are_c4_and_ends_meet_the_boundary_conditions():
    rng = np.random.default_rng(5)
    times = np.sort(np.concatenate([[0.0, 2.0], rng.uniform(0.05, 1.95, 6)]))
    x = rng.uniform(-1, 1, (8, 3))
    bc = rng.uniform(-1, 1, (4, 3))
    co, v, a = quintic_spline_coefficients(times, x, *bc)
    T = np.diff(times)
    der = lambda c, t, k: sum(np.prod(np.arange(i, i - k, -1)) * c[i] * t ** (i - k) for i in range(k, 6))
    for i in range(7):
        assert np.abs(der(co[i], 0.0, 0) - x[i]).max() < 1e-13 and np.abs(der(co[i], T[i], 0) - x[i + 1]).max() < 1e-10
    for i in range(6):
        for k in range(1, 5):            # velocity ... fourth derivative continuous
            l, r = der(co[i], T[i], k), der(co[i + 1], 0.0, k)
            assert np.abs(l - r).max() <= 1e-8 * max(1.0, np.abs(r).max()), (i, k)
    assert np.abs(der(co[0], 0.0, 1) - bc[0]).max() < 1e-13 and np.abs(der(co[0], 0.0, 2) - bc[1]).max() < 1e-13
    assert np.abs(der(co[-1], T[-1], 1) - bc[2]).max() < 1e-9 and np.abs(der(co[-1], T[-1], 2) - bc[3]).max() < 1e-8
    # outside the knots the spline holds its end states, derivative included (Spline.h:607-625)
    p0, v0, _ = spline_evaluate(times, co, -1.0)
    p1, v1, _ = spline_evaluate(times, co, 5.0)
    assert np.abs(p0 - x[0]).max() < 1e-13 and np.abs(v0 - bc[0]).max() < 1e-13
    assert np.abs(p1 - x[-1]).max() < 1e-10 and np.abs(v1 - bc[2]).max() < 1e-9


def test_two_knots_at_rest_is_the_minimum_jerk_profile_and_bad_knots_are_refused():
    times = np.array([1.0, 1.5])
    x = np.array([[0.0, 1.0], [2.0, -1.0]])
    co, _, _ = quintic_spline_coefficients(times, x)
    for tau in (0.0, 0.2, 0.5, 0.9):
        s = 10 * tau ** 3 - 15 * tau ** 4 + 6 * tau ** 5
        pos, vel, _ = spline_evaluate(times, co, 1.0 + 0.5 * tau)
        assert np.abs(pos - (x[0] + s * (x[1] - x[0]))).max() < 1e-13
        assert np.abs(vel - 30 * tau ** 2 * (1 - tau) ** 2 / 0.5 * (x[1] - x[0])).max() < 1e-12
    with pytest.raises(ValueError):
        quintic_spline_coefficients([0.0], x[:1])
    with pytest.raises(ValueError):
        quintic_spline_coefficients([0.0, 0.0], x)


def test_cubic_spline_through_samples_of_a_cubic_is_that_cubic():
    """SplineTest.cpp "Cubic spline" (:185-330): 100 knots on a random cubic in R^4 between 320 ms and 642 s; position and
    velocity are checked (isApprox 1e-5); plus C2 at the interior knots of an uneven spline and the clamped end velocities."""
    rng = np.random.default_rng(9)
    c = np.concatenate([rng.uniform(-1.0, 1.0, (4, 4)), np.zeros((2, 4))])
    times = 0.32 + (642.0 - 0.32) / 99 * np.arange(100)
    x = np.stack([_poly(c, t) for t in times])
    co, _ = cubic_spline_coefficients(times, x, _poly(c, times[0], 1), _poly(c, times[-1], 1))
    for i in range(0, 1000, 7):
        t = 0.32 + (times[-1] - 0.32) / 1000 * i
        pos, vel, _ = spline_evaluate(times, co, t)
        for got, k in ((pos, 0), (vel, 1)):
            want = _poly(c, t, k)
            assert np.linalg.norm(got - want) <= 1e-5 * min(np.linalg.norm(got), np.linalg.norm(want))
    times = np.sort(np.concatenate([[0.0, 2.0], rng.uniform(0.05, 1.95, 6)]))
    x = rng.uniform(-1, 1, (8, 3))
    va, vb = rng.uniform(-1, 1, 3), rng.uniform(-1, 1, 3)
    co, v = cubic_spline_coefficients(times, x, va, vb)
    T = np.diff(times)
    der = lambda cc, t, k: sum(np.prod(np.arange(i, i - k, -1)) * cc[i] * t ** (i - k) for i in range(k, 6))
    for i in range(6):
        for k in range(0, 3):
            assert np.abs(der(co[i], T[i], k) - der(co[i + 1], 0.0, k)).max() < 1e-9, (i, k)
    assert np.abs(der(co[0], 0.0, 1) - va).max() < 1e-13 and np.abs(der(co[-1], T[-1], 1) - vb).max() < 1e-10
    assert np.all(co[:, 4:] == 0.0)
