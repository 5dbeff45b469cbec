"""Minimal stand-in for the ``manifpy`` types the reference bindings exchange (SO3, SE3 and their tangents).
Quaternions are stored (x, y, z, w) as in manif / Eigen; ``tangent + group`` is the left-plus ``Exp(tau) * X`` the
integrator uses (ForwardEuler.h:82); ``X - Y`` is the right-minus ``Log(Y^-1 X)`` the reference tests use."""
from __future__ import annotations

import numpy as np


def _quat_mul(a, b):
    ax, ay, az, aw = a
    bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz])


class SO3Tangent:
    def __init__(self, coeffs=(0.0, 0.0, 0.0)):
        self._c = np.asarray(coeffs, dtype=np.float64).reshape(3).copy()

    def coeffs(self):
        return self._c

    @staticmethod
    def Zero():
        return SO3Tangent()

    def exp(self) -> "SO3":
        th = np.linalg.norm(self._c)
        k = 0.5 - th * th / 48.0 if th < 1e-8 else np.sin(th / 2) / th
        return SO3(quaternion=np.concatenate([k * self._c, [np.cos(th / 2)]]))

    def __add__(self, X):
        if isinstance(X, SO3Tangent):           # tangent + tangent
            return SO3Tangent(self._c + X._c)
        return self.exp() * X                   # left-plus: Exp(tau) * X

    def __mul__(self, s: float):
        return SO3Tangent(self._c * s)

    __rmul__ = __mul__


class SO3:
    def __init__(self, quaternion=(0.0, 0.0, 0.0, 1.0)):
        q = np.asarray(quaternion, dtype=np.float64).reshape(4)
        self._q = q / np.linalg.norm(q)

    @staticmethod
    def Identity():
        return SO3()

    @staticmethod
    def Random(rng=None):
        q = (rng or np.random.default_rng()).normal(size=4)
        return SO3(quaternion=q)

    def __add__(self, tau: "SO3Tangent") -> "SO3":   # right-plus: X * Exp(tau)
        return self * tau.exp()

    def quat(self):
        return self._q

    def coeffs(self):
        return self._q

    def rotation(self):
        x, y, z, w = self._q
        return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                         [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                         [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])

    def inverse(self):
        return SO3(quaternion=self._q * np.array([-1.0, -1.0, -1.0, 1.0]))

    def __mul__(self, other: "SO3") -> "SO3":
        return SO3(quaternion=_quat_mul(self._q, other._q))

    def log(self) -> SO3Tangent:
        q = self._q if self._q[3] >= 0 else -self._q
        n = np.linalg.norm(q[:3])
        if n < 1e-10:
            return SO3Tangent(2.0 / q[3] * (1.0 - n * n / (3.0 * q[3] * q[3])) * q[:3])
        return SO3Tangent(2.0 * np.arctan2(n, q[3]) / n * q[:3])

    def __sub__(self, other: "SO3") -> SO3Tangent:   # right-minus
        return (other.inverse() * self).log()


class SE3Tangent:
    def __init__(self, coeffs=(0.0,) * 6):
        self._c = np.asarray(coeffs, dtype=np.float64).reshape(6).copy()

    def coeffs(self):
        return self._c

    def lin(self):
        return self._c[:3]

    def ang(self):
        return self._c[3:]

    @staticmethod
    def Zero():
        return SE3Tangent()


class SE3:
    def __init__(self, position=(0.0, 0.0, 0.0), quaternion=(0.0, 0.0, 0.0, 1.0)):
        self._p = np.asarray(position, dtype=np.float64).reshape(3).copy()
        self._R = SO3(quaternion=quaternion)

    @staticmethod
    def Identity():
        return SE3()

    def translation(self):
        return self._p

    def quat(self):
        return self._R.quat()

    def rotation(self):
        return self._R.rotation()

    def asSO3(self):
        return self._R
