"""bindings/python/ContinuousDynamicalSystem/src/FloatingBaseSystemVelocityKinematics.cpp, ForwardEuler / Integrator
bindings: ``FloatingBaseSystemVelocityKinematics`` with the ``state`` property and ``set_control_input``, and
``FloatingBaseSystemVelocityKinematicsForwardEulerIntegrator`` with ``integration_step``, ``get_solution`` and
``integrate(initial_time, final_time)``.  One Euler step: p += v dt, R <- Exp(w dt) R (left-plus, ForwardEuler.h:82),
s += sdot dt -- computed by the CUDA integrator (blfik_integrate_state); a state may carry a leading batch dimension."""
from __future__ import annotations

import ctypes as C
from datetime import timedelta

import numpy as np

from .. import capi
from . import manif
from ._log import log


class FloatingBaseSystemVelocityKinematics:
    def __init__(self):
        self._p = self._q = self._s = None
        self._twist = self._sdot = None

    def initialize(self, handler=None):
        return True   # `rho` is accepted and ignored (FloatingBaseSystemVelocityKinematics.cpp:14-31)

    # state = (base position, base rotation (manif.SO3 or quaternion array), joint positions)
    def set_state(self, state):
        p, R, s = state
        p, s = np.asarray(p, dtype=np.float64), np.asarray(s, dtype=np.float64)
        q = R.quat() if isinstance(R, manif.SO3) else np.asarray(R, dtype=np.float64)
        self._batched = p.ndim == 2
        self._p, self._q, self._s = np.atleast_2d(p).copy(), np.atleast_2d(q).copy(), np.atleast_2d(s).copy()
        return True

    def get_state(self):
        if self._p is None:
            return None
        if self._batched:
            return self._p, self._q, self._s
        return self._p[0], manif.SO3(quaternion=self._q[0]), self._s[0]

    state = property(get_state, set_state)

    def set_control_input(self, control_input):
        twist, sdot = control_input
        twist = twist.coeffs() if isinstance(twist, manif.SE3Tangent) else np.asarray(twist, dtype=np.float64)
        sdot = np.asarray(sdot, dtype=np.float64)
        if self._s is not None and sdot.shape[-1] != self._s.shape[-1]:
            log().error("[FloatingBaseSystemVelocityKinematics::setControlInput] The size of the joint velocity is not coherent with the state.")
            return False
        if twist.shape[-1] != 6:
            log().error("[FloatingBaseSystemVelocityKinematics::setControlInput] The base twist must have 6 components.")
            return False
        self._twist, self._sdot = np.atleast_2d(twist).copy(), np.atleast_2d(sdot).copy()
        return True


class FloatingBaseSystemVelocityKinematicsForwardEulerIntegrator:
    def __init__(self):
        self._system, self._dt = None, None

    def set_dynamical_system(self, system):
        if self._system is not None:
            log().error("[Integrator::setDynamicalSystem] The dynamical system has been already set.")   # Integrator.h:94-99
            return False
        self._system = system
        return True

    def dynamical_system(self):
        return self._system

    def _get_step(self):
        return self._dt

    def _set_step(self, dt):
        if not isinstance(dt, timedelta) or dt <= timedelta(0):
            raise ValueError("the integration step must be a positive datetime.timedelta")
        self._dt = dt

    integration_step = property(_get_step, _set_step)

    def get_solution(self):
        return self._system.get_state()

    def integrate(self, initial_time, final_time):
        prefix = "[FixedStepIntegrator::integrate]"
        sys = self._system
        if sys is None:
            log().error("%s Please set the dynamical system before calling this function.", prefix)
            return False
        if self._dt is None:
            log().error("%s Please set the integration step.", prefix)
            return False
        if final_time < initial_time:
            log().error("%s The final time has to be greater than the initial one.", prefix)
            return False
        if sys._p is None or sys._twist is None:
            log().error("%s The state or the control input of the system has not been set.", prefix)
            return False
        dt = self._dt.total_seconds()
        span = (final_time - initial_time).total_seconds()
        steps = int(round(span / dt))
        if steps < 1:   # FixedStepIntegrator.h:117-119 underflows here; this implementation refuses
            log().error("%s The integration interval is shorter than the integration step.", prefix)
            return False
        N, n = sys._p.shape[0], sys._s.shape[1]
        nu = np.ascontiguousarray(np.hstack([np.broadcast_to(sys._twist, (N, 6)), np.broadcast_to(sys._sdot, (N, n))]))
        L = capi.load()
        dp = C.POINTER(C.c_double)
        rc = L.blfik_integrate_state(C.c_int32(n), C.c_int64(N), sys._p.ctypes.data_as(dp), sys._q.ctypes.data_as(dp),
                                     sys._s.ctypes.data_as(dp), nu.ctypes.data_as(dp), C.c_double(dt), C.c_int32(steps),
                                     C.c_int32(capi.MEM_HOST), C.c_int32(0))
        if rc != 0:
            log().error("%s %s", prefix, L.blfik_last_error().decode())
            return False
        return True


class _FixedStepIntegrator:
    """Integrator.h:41-129 / FixedStepIntegrator.h:38-140 as the bindings expose them (set_dynamical_system,
    set_integration_step / integration_step, get_solution, integrate)."""

    def __init__(self):
        self._system, self._dt = None, None

    def set_dynamical_system(self, system):
        if self._system is not None:
            log().error("[Integrator::setDynamicalSystem] The dynamical system has been already set.")
            return False
        if system is None:
            log().error("[Integrator::setDynamicalSystem] The dynamical system passed to the function is corrupted.")
            return False
        self._system = system
        return True

    def dynamical_system(self):
        return self._system

    def set_integration_step(self, dt):
        if not isinstance(dt, timedelta) or dt <= timedelta(0):
            log().error("[FixedStepIntegrator::setIntegrationStep] The integration must be a strictly positive number.")
            return False
        self._dt = dt
        return True

    def _get_step(self):
        return self._dt

    def _set_step(self, dt):
        if not self.set_integration_step(dt):
            raise ValueError("the integration step must be a positive datetime.timedelta")

    integration_step = property(_get_step, _set_step)

    def get_solution(self):
        return self._system.get_state()

    def _steps(self, initial_time, final_time):
        """(number of full steps, length of the last one) or None: (iterations - 1) steps of dT, then one that lands on the
        final time (FixedStepIntegrator.h:94-140)"""
        prefix = "[FixedStepIntegrator::integrate]"
        if self._system is None:
            log().error("%s Please set the dynamical system before call this function.", prefix)
            return None
        if self._dt is None:
            log().error("%s Please set the integration step.", prefix)
            return None
        if final_time < initial_time:
            log().error("%s The final time has to be greater than the initial one.", prefix)
            return None
        iterations = (final_time - initial_time) // self._dt
        if iterations < 1:
            log().error("%s The integration interval is shorter than the integration step.", prefix)
            return None
        last = final_time - (initial_time + self._dt * (iterations - 1))
        return iterations - 1, last.total_seconds()


class LinearTimeInvariantSystem:
    """LinearTimeInvariantSystem.h:33-118, .cpp:14-115: dx = A x + B u; state / input / derivative are 1-tuples."""

    def __init__(self):
        self._A = self._B = None
        self._x = np.zeros(0)
        self._u = np.zeros(0)

    def initialize(self, handler=None):
        return True

    def set_system_matrices(self, A, B):
        prefix = "[LinearTimeInvariantSystem::setSystemMatrices]"
        A = np.asarray(A, dtype=np.float64)
        B = np.asarray(B, dtype=np.float64)
        if B.ndim == 1:
            B = B.reshape(-1, 1)
        if A.ndim != 2 or A.shape[0] != B.shape[0]:
            log().error("%s A and B must have the same number of rows.", prefix)
            return False
        if A.shape[0] != A.shape[1]:
            log().error("%s A must be a square matrix.", prefix)
            return False
        self._A, self._B = A.copy(), B.copy()
        return True

    def set_state(self, state):
        self._x = np.array(state[0], dtype=np.float64).reshape(-1)
        return True

    def get_state(self):
        return (self._x,)

    state = property(get_state, set_state)

    def set_control_input(self, control_input):
        self._u = np.array(control_input[0], dtype=np.float64).reshape(-1)
        return True

    def _derivative(self, x):
        prefix = "[LinearTimeInvariantSystem::dynamics]"
        if self._A is None:
            log().error("%s Please initialize the matrices. Call setSystemMatrices()", prefix)
            return None
        if x.shape[0] != self._A.shape[0]:
            log().error("%s The size of the vector 'state' is not coherent with the system matrices.", prefix)
            return None
        if self._u.shape[0] != self._B.shape[1]:
            log().error("%s The size of the vector 'control input' is not coherent with the system matrices.", prefix)
            return None
        return self._A @ x + self._B @ self._u

    def dynamics(self, time=None):
        dx = self._derivative(self._x)
        return None if dx is None else (dx,)


class LinearTimeInvariantSystemForwardEulerIntegrator(_FixedStepIntegrator):
    """ForwardEuler.h:49-124 over a LinearTimeInvariantSystem (a handful of doubles: host arithmetic)"""

    def _one(self, dt):
        sys = self._system
        dx = sys._derivative(sys._x)
        if dx is None:
            log().error("[ForwardEuler::oneStepIntegration] Unable to compute the system dynamics.")
            return False
        sys._x = sys._x + dt * dx
        return True

    def integrate(self, initial_time, final_time):
        plan = self._steps(initial_time, final_time)
        if plan is None:
            return False
        full, last = plan
        dt = self._dt.total_seconds()
        return all(self._one(dt) for _ in range(full)) and self._one(last)


class LinearTimeInvariantSystemRK4Integrator(LinearTimeInvariantSystemForwardEulerIntegrator):
    """RK4.h:37-175: x+ = x + dT / 6 (k1 + 2 k2 + 2 k3 + k4), the input held over the step"""

    def _one(self, dt):
        sys = self._system
        x = sys._x
        k1 = sys._derivative(x)
        if k1 is None:
            log().error("[RK4::oneStepIntegration] Unable to compute the system dynamics.")
            return False
        k2 = sys._derivative(x + 0.5 * dt * k1)
        k3 = sys._derivative(x + 0.5 * dt * k2)
        k4 = sys._derivative(x + dt * k3)
        sys._x = x + dt / 6.0 * (k1 + 2.0 * k2 + 2.0 * k3 + k4)
        return True


class FloatingBaseSystemAccelerationKinematics:
    """FloatingBaseSystemAccelerationKinematics.h:30-150, .cpp:14-98: state = (p, R, s, dp, omega, ds), input = (base
    acceleration in R6 (mixed), joint accelerations); a state may carry a leading batch dimension."""

    def __init__(self):
        self._p = self._q = self._s = self._dp = self._w = self._ds = None
        self._acc = None
        self._batched = False

    def initialize(self, handler=None):
        return True

    def set_state(self, state):
        p, R, s, dp, w, ds = state
        q = R.quat() if isinstance(R, manif.SO3) else np.asarray(R, dtype=np.float64)
        w = w.coeffs() if hasattr(w, "coeffs") else np.asarray(w, dtype=np.float64)
        arr = [np.asarray(a, dtype=np.float64) for a in (p, q, s, dp, w, ds)]
        if arr[2].shape != arr[5].shape:
            log().error("[FloatingBaseSystemAccelerationKinematics::setState] Wrong size of the joint vectors.")
            return False
        self._batched = arr[0].ndim == 2
        self._p, self._q, self._s, self._dp, self._w, self._ds = (np.atleast_2d(a).copy() for a in arr)
        return True

    def get_state(self):
        if self._p is None:
            return None
        if self._batched:
            return self._p, self._q, self._s, self._dp, self._w, self._ds
        return self._p[0], manif.SO3(quaternion=self._q[0]), self._s[0], self._dp[0], self._w[0], self._ds[0]

    state = property(get_state, set_state)

    def set_control_input(self, control_input):
        dtwist, dds = control_input
        dtwist, dds = np.asarray(dtwist, dtype=np.float64), np.asarray(dds, dtype=np.float64)
        if dtwist.shape[-1] != 6 or (self._s is not None and dds.shape[-1] != self._s.shape[-1]):
            log().error("[FloatingBaseSystemAccelerationKinematics::dynamics] Wrong size of the vectors.")
            return False
        self._acc = (np.atleast_2d(dtwist).copy(), np.atleast_2d(dds).copy())
        return True


class FloatingBaseSystemAccelerationKinematicsForwardEulerIntegrator(_FixedStepIntegrator):
    """every component advances with the derivative at the OLD state (ForwardEuler.h:102-121), on the GPU
    (blfik_integrate_acceleration_state)"""

    def integrate(self, initial_time, final_time):
        plan = self._steps(initial_time, final_time)
        if plan is None:
            return False
        sys = self._system
        if sys._p is None or sys._acc is None:
            log().error("[ForwardEuler::oneStepIntegration] Unable to compute the system dynamics.")
            return False
        N, n = sys._p.shape[0], sys._s.shape[1]
        acc = np.ascontiguousarray(np.hstack([np.broadcast_to(sys._acc[0], (N, 6)), np.broadcast_to(sys._acc[1], (N, n))]))
        L = capi.load()
        dp = C.POINTER(C.c_double)
        full, last = plan
        for dt, steps in ((self._dt.total_seconds(), full), (last, 1)):
            if steps < 1:
                continue
            rc = L.blfik_integrate_acceleration_state(C.c_int32(n), C.c_int64(N), *(a.ctypes.data_as(dp) for a in (sys._p, sys._q, sys._s, sys._dp, sys._w, sys._ds)),
                                                      acc.ctypes.data_as(dp), C.c_double(dt), C.c_int32(steps), C.c_int32(0))
            if rc != 0:
                log().error("[ForwardEuler::oneStepIntegration] %s", L.blfik_last_error().decode())
                return False
        return True
