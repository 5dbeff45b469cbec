"""``Planners::SO3Planner`` (src/Planners/include/BipedalLocomotion/Planners/SO3Planner.h:23-161, SO3Planner.tpp:24-310): the
minimum-jerk rotation trajectory between two orientations that the reference's callers evaluate every period and pass to
``SO3Task.set_set_point`` / ``SE3Task.set_set_point``.  (The reference binds it only inside ``SwingFootPlanner``; the class
is offered here because it is what ``blfik_set_setpoint_minimum_jerk`` evaluates on the device for a whole batch:
``SO3PlannerInertial`` with zero boundary conditions -- tests/test_python_bindings.py checks the two against each other.)

    R(t) = Exp(s(t) d) R_0,  d = Log(R_1 R_0^-1)     "Right" trivialisation, inertial-frame velocity  w = sdot d
    R(t) = R_0 Exp(s(t) d),  d = Log(R_0^-1 R_1)     "Left" trivialisation, body-frame velocity
with s the quintic that meets the boundary velocities and accelerations, which must be parallel to d (:51-99)."""
from __future__ import annotations

import numpy as np

from . import manif
from ._log import log


def _seconds(t):
    return t.total_seconds() if hasattr(t, "total_seconds") else float(t)


class SO3PlannerState:
    """SO3Planner.h:31-38"""

    def __init__(self):
        self.rotation = manif.SO3.Identity()
        self.velocity = manif.SO3Tangent.Zero()
        self.acceleration = manif.SO3Tangent.Zero()


class _SO3Planner:
    _right = True

    def __init__(self):
        self._R0 = manif.SO3.Identity()
        self._distance = np.zeros(3)
        self._T = 0.0
        self._dt = 0.0
        self._advance_time = 0.0
        self._state = SO3PlannerState()
        self._ic = None
        self._fc = None
        self._poly = None

    @classmethod
    def _compute_distance(cls, R0, R1):
        """:276-287"""
        return ((R1 * R0.inverse()) if cls._right else (R0.inverse() * R1)).log().coeffs().copy()

    @classmethod
    def project_tangent_vector(cls, initial_rotation, final_rotation, vector):
        """:289-305"""
        d = cls._compute_distance(initial_rotation, final_rotation)
        n = np.linalg.norm(d)
        if n < 1e-4:
            return manif.SO3Tangent.Zero()
        return manif.SO3Tangent(d.dot(vector.coeffs()) / n * d)

    def set_rotations(self, initial_rotation, final_rotation, duration):
        T = _seconds(duration)
        if T <= 0:
            log().error("[SO3Planner::setRotation] The trajectory duration must be a strictly positive number.")
            return False
        self._R0, self._T = initial_rotation, T
        self._distance = self._compute_distance(initial_rotation, final_rotation)
        self._advance_time = 0.0
        self._state = SO3PlannerState()
        self._state.rotation = initial_rotation
        self._poly = None
        return True

    def set_initial_conditions(self, velocity, acceleration):
        self._ic, self._poly = (velocity.coeffs().copy(), acceleration.coeffs().copy()), None
        return True

    def set_final_conditions(self, velocity, acceleration):
        self._fc, self._poly = (velocity.coeffs().copy(), acceleration.coeffs().copy()), None
        return True

    def _compute_coefficients(self):
        prefix = "[SO3Planner::computeCoefficients]"
        if self._poly is not None:
            return True
        if self._ic is None:
            log().error("%s The initial boundary condition is not set.", prefix)
            return False
        if self._fc is None:
            log().error("%s The final boundary condition is not set.", prefix)
            return False

        def coefficient(vector):
            """the scalar k with vector = k * distance, None when they are not parallel (:51-86, tolerance 1e-3 on the cosine)"""
            d = self._distance
            if not np.any(d) or not np.any(vector):
                return 0.0
            c = d.dot(vector) / (np.linalg.norm(d) * np.linalg.norm(vector))
            if abs(abs(c) - 1.0) > 1e-3:
                return None
            for di, vi in zip(d, vector):
                if di != 0.0 and np.isfinite(vi / di):
                    return vi / di
            return None

        ks = [coefficient(v) for v in (self._ic[0], self._ic[1], self._fc[0], self._fc[1])]
        if any(k is None for k in ks):
            log().error("%s The velocity and the boundary acceleration must be a linear dependent to the vector representing the axis "
                        "between the two rotations.", prefix)
            return False
        kv0, ka0, kv1, ka1 = ks
        T = self._T
        # the quintic s with s(0) = 0, s(T) = 1 and the four boundary slopes (:101-111)
        self._poly = (0.0, kv0, ka0 / 2.0,
                      (20.0 - (8.0 * kv1 + 12.0 * kv0) * T - (3.0 * ka0 - ka1) * T * T) / (2.0 * T ** 3),
                      (-30.0 + (14.0 * kv1 + 16.0 * kv0) * T + (3.0 * ka0 - 2.0 * ka1) * T * T) / (2.0 * T ** 4),
                      (12.0 - 6.0 * (kv1 + kv0) * T - (ka0 - ka1) * T * T) / (2.0 * T ** 5))
        return True

    def evaluate_point(self, time, state=None):
        """:171-232; returns the state when none is passed in, False on failure"""
        prefix = "[SO3Planner::evaluatePoint]"
        t = _seconds(time)
        if t < 0 or t > self._T:
            log().error("%s The time has to be a real positive number between 0 and %s.", prefix, self._T)
            return False
        if not self._compute_coefficients():
            log().error("%s Unable to compute the coefficients.", prefix)
            return False
        a = self._poly
        s = a[0] + t * (a[1] + t * (a[2] + t * (a[3] + t * (a[4] + t * a[5]))))
        ds = a[1] + t * (2 * a[2] + t * (3 * a[3] + t * (4 * a[4] + t * 5 * a[5])))
        dds = 2 * a[2] + t * (6 * a[3] + t * (12 * a[4] + t * 20 * a[5]))
        out = state if state is not None else SO3PlannerState()
        step = manif.SO3Tangent(self._distance * s).exp()
        out.rotation = step * self._R0 if self._right else self._R0 * step
        out.velocity = manif.SO3Tangent(self._distance * ds)
        out.acceleration = manif.SO3Tangent(self._distance * dds)
        return out if state is None else True

    def set_advance_time_step(self, dt):
        if _seconds(dt) <= 0:
            log().error("[SO3Planner::setAdvanceTimeStep] The time step of the advance time has to be a strictly positive number.")
            return False
        self._dt = _seconds(dt)
        return True

    def is_output_valid(self):
        return self._dt != 0.0

    def advance(self):
        if not self.is_output_valid():
            log().error("[SO3Planner::advance] The advance capabilities cannot be used. Did you set the advance time step?")
            return False
        self._advance_time = min(self._advance_time + self._dt, self._T)
        return self.evaluate_point(self._advance_time, self._state) is True

    def get_output(self):
        return self._state


class SO3PlannerInertial(_SO3Planner):
    """SO3Planner<LieGroupTrivialization::Right> (SO3Planner.h:152)"""
    _right = True


class SO3PlannerBody(_SO3Planner):
    """SO3Planner<LieGroupTrivialization::Left> (SO3Planner.h:158)"""
    _right = False
