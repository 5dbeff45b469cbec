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
