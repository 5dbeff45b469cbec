"""bindings/python/Math/src/Constants.cpp"""
StandardAccelerationOfGravitation = 9.80665


# ---------------------------------------------------------------------------------------------------------------------
# Splines: bindings/python/Math/src/Spline.cpp:37-110 (TrajectoryPoint, Spline, QuinticSpline with Eigen::VectorXd knots)
# over src/Math/include/BipedalLocomotion/Math/Spline.h and QuinticSpline.h.  This is the host-side, one-trajectory
# class with the reference's method names; a batch of robots gets its splines evaluated on the GPU through
# IKLinearTask.set_set_point_spline (blfik_set_setpoint_spline), to which an object of this class can be handed.
import numpy as _np

from ._log import log as _log


def _seconds(t):
    """the reference's bindings take datetime.timedelta (std::chrono::nanoseconds); plain seconds are accepted too"""
    return t.total_seconds() if hasattr(t, "total_seconds") else float(t)


class TrajectoryPoint:
    """Spline.h:31-38"""

    def __init__(self, n=0):
        self.position = _np.zeros(n)
        self.velocity = _np.zeros(n)
        self.acceleration = _np.zeros(n)
        self.time = 0.0


class QuinticSpline:
    """Math::QuinticSpline<Eigen::VectorXd> (QuinticSpline.h:26-90)"""

    def __init__(self):
        self._times = None
        self._knots = None
        self._ic = None           # (velocity, acceleration) at the first knot
        self._fc = None
        self._coeffs = None       # [n - 1, 6, dim], local time of every interval
        self._dt = None
        self._advance_time = 0.0
        self._point = TrajectoryPoint()
        self._output_valid = False

    # ---- Spline.h:323-383
    def set_knots(self, position, time):
        prefix = "[Spline::setKnots]"
        if len(position) < 2:
            _log().error("%s The number of knots has to be at least 2.", prefix)
            return False
        if len(time) != len(position):
            _log().error("%s The number of points has to be equal to the length of the time vector.", prefix)
            return False
        knots = [_np.asarray(k, dtype=_np.float64).reshape(-1) for k in position]
        for k in knots:
            if k.size != knots[0].size:
                _log().error("%s The size of the knot is different from the one expected. Expected: %d Retrieved %d.", prefix, knots[0].size, k.size)
                return False
        self._knots = _np.stack(knots)
        self._times = _np.array([_seconds(t) for t in time], dtype=_np.float64)
        self._point = TrajectoryPoint(knots[0].size)
        self._point.position = self._knots[0].copy()
        self._advance_time = self._times[0]
        self._coeffs = None
        return True

    # ---- Spline.h:385-436
    def set_initial_conditions(self, velocity, acceleration):
        v, a = _np.asarray(velocity, dtype=_np.float64).reshape(-1), _np.asarray(acceleration, dtype=_np.float64).reshape(-1)
        if v.size != a.size:
            _log().error("[Spline::setInitialConditions] The size of the velocity is different from the one of the acceleration. "
                         "Expected: %d Retrieved: %d.", v.size, a.size)
            return False
        self._ic, self._coeffs = (v.copy(), a.copy()), None
        return True

    def set_final_conditions(self, velocity, acceleration):
        v, a = _np.asarray(velocity, dtype=_np.float64).reshape(-1), _np.asarray(acceleration, dtype=_np.float64).reshape(-1)
        if v.size != a.size:
            _log().error("[Spline::setFinalConditions] The size of the velocity is different from the one of the acceleration. "
                         "Expected: %d Retrieved: %d.", v.size, a.size)
            return False
        self._fc, self._coeffs = (v.copy(), a.copy()), None
        return True

    def set_advance_time_step(self, dt):
        dt = _seconds(dt)
        if dt <= 0:
            _log().error("[Spline::setAdvanceTimeStep] The time step has to be greater than zero.")
            return False
        self._dt = dt
        return True

    # ---- Spline.h:516-577, QuinticSpline.h:266-415
    def _compute_coefficients(self):
        prefix = "[Spline::computeCoefficients]"
        if self._coeffs is not None:
            return True
        if self._knots is None:
            _log().error("%s Please set the knots before computing the coefficients.", prefix)
            return False
        dim = self._knots.shape[1]
        for what, bc in (("initial", self._ic), ("final", self._fc)):
            if bc is None or bc[0].size != dim or bc[1].size != dim:
                _log().error("[Spline::setBoundaryVelocitiesAndAcceleration] The size of the %s velocity or acceleration is different from "
                             "the expected one. Expected: %d.", what, dim)
                _log().error("%s Unable to set the boundary conditions related to the velocity and acceleration.", prefix)
                return False
        T = _np.diff(self._times)
        if _np.any(T == 0.0):
            _log().error("[Splines::computePhasesDuration] Two consecutive points have the same time coordinate.")
            _log().error("%s Unable to compute the phase duration.", prefix)
            return False
        x, n = self._knots, len(self._times)
        v, a = _np.zeros_like(x), _np.zeros_like(x)
        v[0], a[0], v[-1], a[-1] = self._ic[0], self._ic[1], self._fc[0], self._fc[1]
        m = n - 2
        if m > 0:
            # third and fourth derivative continuous at the interior knots: block-tridiagonal in (v_i, a_i), the blocks
            # depend on the interval lengths only; eliminated block by block, all coordinates at once
            Tp, Tn = T[:-1], T[1:]
            L = _np.array([[-24 / Tp ** 2, -3 / Tp], [-168 / Tp ** 3, -24 / Tp ** 2]]).transpose(2, 0, 1)
            D = _np.array([[-36 / Tp ** 2 + 36 / Tn ** 2, 9 / Tp + 9 / Tn],
                           [-192 / Tp ** 3 - 192 / Tn ** 3, 36 / Tp ** 2 - 36 / Tn ** 2]]).transpose(2, 0, 1)
            U = _np.array([[24 / Tn ** 2, -3 / Tn], [-168 / Tn ** 3, 24 / Tn ** 2]]).transpose(2, 0, 1)
            dp, dn = x[1:-1] - x[:-2], x[2:] - x[1:-1]
            r = _np.stack([60 * (dn / Tn[:, None] ** 3 - dp / Tp[:, None] ** 3),
                           -360 * (dp / Tp[:, None] ** 4 + dn / Tn[:, None] ** 4)], axis=1)      # [m, 2, dim]
            r[0] -= L[0] @ _np.stack([v[0], a[0]])
            r[-1] -= U[-1] @ _np.stack([v[-1], a[-1]])
            Dp = D.copy()
            for i in range(1, m):
                M = L[i] @ _np.linalg.inv(Dp[i - 1])
                Dp[i] = D[i] - M @ U[i - 1]
                r[i] = r[i] - M @ r[i - 1]
            sol = _np.zeros_like(r)
            sol[-1] = _np.linalg.solve(Dp[-1], r[-1])
            for i in range(m - 2, -1, -1):
                sol[i] = _np.linalg.solve(Dp[i], r[i] - U[i] @ sol[i + 1])
            v[1:-1], a[1:-1] = sol[:, 0], sol[:, 1]
        Tc = T[:, None]
        d = x[1:] - x[:-1]
        v0, v1, a0, a1 = v[:-1], v[1:], a[:-1], a[1:]
        self._coeffs = _np.stack([
            x[:-1], v0, 0.5 * a0,
            (20 * d - (8 * v1 + 12 * v0) * Tc - (3 * a0 - a1) * Tc ** 2) / (2 * Tc ** 3),
            (-30 * d + (14 * v1 + 16 * v0) * Tc + (3 * a0 - 2 * a1) * Tc ** 2) / (2 * Tc ** 4),
            (12 * d - 6 * (v1 + v0) * Tc - (a0 - a1) * Tc ** 2) / (2 * Tc ** 5)], axis=1)
        return True

    def _evaluate(self, t):
        times, c = self._times, self._coeffs
        if t < times[0]:                              # Spline.h:607-614
            i, tl = 0, 0.0
        elif t >= times[-1]:                          # :616-625
            i, tl = len(times) - 2, times[-1] - times[-2]
        else:
            i = int(_np.searchsorted(times, t, side="right")) - 1
            tl = t - times[i]
        k = c[i]
        pos = k[0] + tl * (k[1] + tl * (k[2] + tl * (k[3] + tl * (k[4] + tl * k[5]))))
        vel = k[1] + tl * (2 * k[2] + tl * (3 * k[3] + tl * (4 * k[4] + tl * 5 * k[5])))
        acc = 2 * k[2] + tl * (6 * k[3] + tl * (12 * k[4] + tl * 20 * k[5]))
        return pos, vel, acc

    # ---- Spline.h:594-644; like the reference's binding the results are written into the arrays passed in
    def evaluate_point(self, t, position, velocity=None, acceleration=None):
        if not self._compute_coefficients():
            _log().error("[Spline::evaluatePoint] Unable to compute the coefficients of the spline.")
            return False
        pos, vel, acc = self._evaluate(_seconds(t))
        if isinstance(position, TrajectoryPoint):
            position.position, position.velocity, position.acceleration, position.time = pos, vel, acc, _seconds(t)
            return True
        _np.asarray(position)[...] = pos
        if velocity is not None:
            _np.asarray(velocity)[...] = vel
        if acceleration is not None:
            _np.asarray(acceleration)[...] = acc
        return True

    # ---- Spline.h:646-711: returns (ok, positions, velocities, accelerations) as the reference's binding does
    def evaluate_ordered_points(self, t):
        if not self._compute_coefficients():
            _log().error("[Spline::evaluateOrderedPoints] Unable to compute the coefficients of the spline.")
            return False, [], [], []
        out = [self._evaluate(_seconds(ti)) for ti in t]
        return True, [o[0] for o in out], [o[1] for o in out], [o[2] for o in out]

    # ---- Spline.h:730-749
    def advance(self):
        self._output_valid = False
        if self._dt is None or not self._compute_coefficients():
            _log().error("[Spline::advance] Unable to evaluate the spline at the current time.")
            return False
        p = self._point
        p.position, p.velocity, p.acceleration = self._evaluate(self._advance_time)
        p.time = self._advance_time
        self._advance_time = min(self._advance_time + self._dt, self._times[-1])
        self._output_valid = True
        return True

    def is_output_valid(self):
        return self._output_valid

    def get_output(self):
        return self._point

    _cubic = False

    # ---- batch extension: what blfik_set_setpoint_spline needs from one trajectory
    def knots(self):
        """(times [n], positions [n, dim], boundary [4, dim]); None where unset"""
        bc = None
        if self._ic is not None and self._fc is not None:
            bc = _np.stack([self._ic[0], self._ic[1], self._fc[0], self._fc[1]])
        return self._times, self._knots, bc


class CubicSpline(QuinticSpline):
    """Math::CubicSpline<Eigen::VectorXd> (CubicSpline.h:26-330): same interface; the knot velocities make the acceleration
    continuous, the boundary accelerations are accepted and not used (as in the reference)."""
    _cubic = True

    def _compute_coefficients(self):
        prefix = "[Spline::computeCoefficients]"
        if self._coeffs is not None:
            return True
        if self._knots is None:
            _log().error("%s Please set the knots before computing the coefficients.", prefix)
            return False
        dim = self._knots.shape[1]
        for what, bc in (("initial", self._ic), ("final", self._fc)):
            if bc is None or bc[0].size != dim or bc[1].size != dim:
                _log().error("[Spline::setBoundaryVelocitiesAndAcceleration] The size of the %s velocity or acceleration is different from "
                             "the expected one. Expected: %d.", what, dim)
                _log().error("%s Unable to set the boundary conditions related to the velocity and acceleration.", prefix)
                return False
        T = _np.diff(self._times)
        if _np.any(T == 0.0):
            _log().error("[Splines::computePhasesDuration] Two consecutive points have the same time coordinate.")
            _log().error("%s Unable to compute the phase duration.", prefix)
            return False
        x, n = self._knots, len(self._times)
        v = _np.zeros_like(x)
        v[0], v[-1] = self._ic[0], self._fc[0]
        m = n - 2
        if m > 0:
            # 2/Tp v_{i-1} + 4 (1/Tp + 1/Tn) v_i + 2/Tn v_{i+1} = 6 (dp/Tp^2 + dn/Tn^2): tridiagonal, eliminated in place
            Tp, Tn = T[:-1], T[1:]
            lo, di, up = 2 / Tp, 4 / Tp + 4 / Tn, 2 / Tn
            r = 6 * ((x[1:-1] - x[:-2]) / Tp[:, None] ** 2 + (x[2:] - x[1:-1]) / Tn[:, None] ** 2)
            r[0] -= lo[0] * v[0]
            r[-1] -= up[-1] * v[-1]
            for i in range(1, m):
                w = lo[i] / di[i - 1]
                di[i] -= w * up[i - 1]
                r[i] -= w * r[i - 1]
            sol = _np.zeros_like(r)
            sol[-1] = r[-1] / di[-1]
            for i in range(m - 2, -1, -1):
                sol[i] = (r[i] - up[i] * sol[i + 1]) / di[i]
            v[1:-1] = sol
        Tc = T[:, None]
        d = x[1:] - x[:-1]
        zero = _np.zeros_like(d)
        self._coeffs = _np.stack([x[:-1], v[:-1], (3 * d - (2 * v[:-1] + v[1:]) * Tc) / Tc ** 2, (-2 * d + (v[:-1] + v[1:]) * Tc) / Tc ** 3,
                                  zero, zero], axis=1)
        return True
