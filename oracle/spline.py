"""CPU restatement of the reference's quintic spline (TEST INFRASTRUCTURE: only tests/, smoke() and bench.py's CPU legs
may import this; the product computes its splines on the GPU, see blfik_set_setpoint_spline).

Follows  src/Math/include/BipedalLocomotion/Math/Spline.h  and  QuinticSpline.h  of the reference:

* knots (t_i, x_i), boundary velocity / acceleration at the first and the last knot (Spline.h:323-383, 458-513);
* one quintic per interval, written in the local time of the interval, built from position, velocity and acceleration
  at its two ends (QuinticSpline.h:266-294);
* the velocities and accelerations of the interior knots are the ones that make the third and the fourth derivative
  continuous there (QuinticSpline.h:297-415: a block-tridiagonal system with 2 x 2 blocks, the same matrix for every
  coordinate; the reference solves it with Eigen::SparseQR, here a dense solve);
* evaluation: before the first knot the first polynomial at local time 0, from the last knot on the last polynomial at
  its end, otherwise the interval with t_i <= t < t_{i+1} (Spline.h:594-644).

Parity: pinned by the reference's own known-answer test for this class (src/Math/tests/SplineTest.cpp:21-167 and
bindings/python/Math/tests/test_quintic_spline.py: the knots sample a random quintic polynomial, the spline must return
that polynomial, its first and its second derivative) -- tests/test_oracle_spline.py replays both.
"""
import numpy as np


def segment_coefficients(T, x0, v0, a0, x1, v1, a1):
    """c[0..5] of x(t) = sum c_k t^k on [0, T] from the two end states (QuinticSpline.h:266-294)."""
    d = x1 - x0
    c3 = (20.0 * d - (8.0 * v1 + 12.0 * v0) * T - (3.0 * a0 - a1) * T * T) / (2.0 * T ** 3)
    c4 = (-30.0 * d + (14.0 * v1 + 16.0 * v0) * T + (3.0 * a0 - 2.0 * a1) * T * T) / (2.0 * T ** 4)
    c5 = (12.0 * d - 6.0 * (v1 + v0) * T - (a0 - a1) * T * T) / (2.0 * T ** 5)
    return np.stack([x0, v0, 0.5 * a0, c3, c4, c5])


def interior_system(times):
    """Blocks of the jerk / snap continuity equations at the interior knots (QuinticSpline.h:93-197 up to a row scale):
    row pair i couples (v, a) of knots i, i+1, i+2 through L[i], D[i], U[i]."""
    T = np.diff(np.asarray(times, dtype=np.float64))
    m = len(T) - 1
    L, D, U = np.zeros((m, 2, 2)), np.zeros((m, 2, 2)), np.zeros((m, 2, 2))
    for i in range(m):
        Tp, Tn = T[i], T[i + 1]
        L[i] = [[-24.0 / Tp ** 2, -3.0 / Tp], [-168.0 / Tp ** 3, -24.0 / Tp ** 2]]
        D[i] = [[-36.0 / Tp ** 2 + 36.0 / Tn ** 2, 9.0 / Tp + 9.0 / Tn],
                [-192.0 / Tp ** 3 - 192.0 / Tn ** 3, 36.0 / Tp ** 2 - 36.0 / Tn ** 2]]
        U[i] = [[24.0 / Tn ** 2, -3.0 / Tn], [-168.0 / Tn ** 3, 24.0 / Tn ** 2]]
    return T, L, D, U


def quintic_spline_coefficients(times, x, v0=None, a0=None, vf=None, af=None):
    """times [n], x [n, ...]; returns (coefficients [n-1, 6, ...], knot velocities [n, ...], knot accelerations [n, ...])."""
    times = np.asarray(times, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    n = len(times)
    if n < 2 or x.shape[0] != n:
        raise ValueError("at least two knots, one time per knot (Spline.h:338-349)")
    if np.any(np.diff(times) == 0.0):
        raise ValueError("two consecutive knots have the same time (Spline.h:444-448)")
    z = np.zeros(x.shape[1:])
    v0, a0, vf, af = [z if b is None else np.asarray(b, dtype=np.float64) for b in (v0, a0, vf, af)]
    v, a = np.zeros_like(x), np.zeros_like(x)
    v[0], a[0], v[-1], a[-1] = v0, a0, vf, af
    T, L, D, U = interior_system(times)
    m = n - 2
    if m > 0:
        r = np.zeros((m, 2) + x.shape[1:])
        for i in range(m):
            dp, dn = x[i + 1] - x[i], x[i + 2] - x[i + 1]
            r[i, 0] = -60.0 * dp / T[i] ** 3 + 60.0 * dn / T[i + 1] ** 3
            r[i, 1] = -360.0 * dp / T[i] ** 4 - 360.0 * dn / T[i + 1] ** 4
        r[0] -= np.einsum("ab,b...->a...", L[0], np.stack([v0, a0]))
        r[-1] -= np.einsum("ab,b...->a...", U[-1], np.stack([vf, af]))
        A = np.zeros((2 * m, 2 * m))
        for i in range(m):
            A[2 * i:2 * i + 2, 2 * i:2 * i + 2] = D[i]
            if i > 0:
                A[2 * i:2 * i + 2, 2 * i - 2:2 * i] = L[i]
            if i < m - 1:
                A[2 * i:2 * i + 2, 2 * i + 2:2 * i + 4] = U[i]
        sol = np.linalg.solve(A, r.reshape(2 * m, -1)).reshape((m, 2) + x.shape[1:])
        v[1:-1], a[1:-1] = sol[:, 0], sol[:, 1]
    coeffs = np.stack([segment_coefficients(T[i], x[i], v[i], a[i], x[i + 1], v[i + 1], a[i + 1]) for i in range(n - 1)])
    return coeffs, v, a


def cubic_spline_coefficients(times, x, v0=None, vf=None):
    """Math::CubicSpline (CubicSpline.h:77-93 the interval polynomial, :95-330 the knot velocities from the continuity of the
    acceleration: 2/Tp v_{i-1} + 4 (1/Tp + 1/Tn) v_i + 2/Tn v_{i+1} = 6 (dp/Tp^2 + dn/Tn^2)).  Returns the same layout as the
    quintic, [n-1, 6, ...] with c_4 = c_5 = 0, and the knot velocities."""
    times = np.asarray(times, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    n = len(times)
    if n < 2 or x.shape[0] != n:
        raise ValueError("at least two knots, one time per knot (Spline.h:338-349)")
    T = np.diff(times)
    if np.any(T == 0.0):
        raise ValueError("two consecutive knots have the same time (Spline.h:444-448)")
    z = np.zeros(x.shape[1:])
    v = np.zeros_like(x)
    v[0], v[-1] = (z if v0 is None else np.asarray(v0, dtype=np.float64)), (z if vf is None else np.asarray(vf, dtype=np.float64))
    m = n - 2
    if m > 0:
        A = np.zeros((m, m))
        r = np.zeros((m,) + x.shape[1:])
        for i in range(m):
            Tp, Tn = T[i], T[i + 1]
            A[i, i] = 4.0 / Tp + 4.0 / Tn
            r[i] = 6.0 * ((x[i + 1] - x[i]) / Tp ** 2 + (x[i + 2] - x[i + 1]) / Tn ** 2)
            if i > 0:
                A[i, i - 1] = 2.0 / Tp
            else:
                r[i] -= 2.0 / Tp * v[0]
            if i < m - 1:
                A[i, i + 1] = 2.0 / Tn
            else:
                r[i] -= 2.0 / Tn * v[-1]
        v[1:-1] = np.linalg.solve(A, r.reshape(m, -1)).reshape(r.shape)
    Tc = T.reshape((-1,) + (1,) * (x.ndim - 1))
    d = x[1:] - x[:-1]
    zero = np.zeros_like(d)
    coeffs = np.stack([x[:-1], v[:-1], (3.0 * d - (2.0 * v[:-1] + v[1:]) * Tc) / Tc ** 2, (-2.0 * d + (v[:-1] + v[1:]) * Tc) / Tc ** 3,
                       zero, zero], axis=1)
    return coeffs, v


def spline_evaluate(times, coeffs, t):
    """position, velocity, acceleration at time t (Spline.h:271-320, 594-644)."""
    times = np.asarray(times, dtype=np.float64)
    if t < times[0]:
        i, tl = 0, 0.0
    elif t >= times[-1]:
        i, tl = len(times) - 2, times[-1] - times[-2]
    else:
        i = int(np.searchsorted(times, t, side="right")) - 1
        tl = t - times[i]
    c = coeffs[i]
    pos = sum(c[k] * tl ** k for k in range(6))
    vel = sum(k * c[k] * tl ** (k - 1) for k in range(1, 6))
    acc = sum(k * (k - 1) * c[k] * tl ** (k - 2) for k in range(2, 6))
    return pos, vel, acc
