"""Stand-in for ``idyntree.swig.KinDynComputations`` as the reference bindings use it (``set_kin_dyn(kindyn)``), holding a
BATCH of N robots on one kinematic tree.  Method names follow iDynTree's (``setRobotState``, ``getWorldTransform``,
``getFrameFreeFloatingJacobian``, ``getCenterOfMassPosition`` ...); the velocity representation is MIXED, the only one the
IK tasks accept (SE3Task.cpp:92-102).  Queries are evaluated for the whole batch by the CUDA kinematics kernel."""
from __future__ import annotations

import numpy as np

from .. import capi, tasks as T
from ..models import KinematicModel
from ._log import log


def _rot_to_quat(R):
    from ..workloads import rot_to_quat
    return rot_to_quat(np.asarray(R, dtype=np.float64).reshape(-1, 3, 3))


class KinDynComputations:
    def __init__(self, model: KinematicModel, batch_size: int = 1, device: int = 0):
        self._km = model
        self.N, self.n, self.device = int(batch_size), model.n, device
        self._p = np.zeros((self.N, 3))
        self._q = np.tile(np.array([0.0, 0.0, 0.0, 1.0]), (self.N, 1))
        self._s = np.zeros((self.N, self.n))
        self._version, self._kin_version, self._kin, self._model = 1, 0, None, None
        self._nu = np.zeros((self.N, self.n + 6))     # generalised velocity, MIXED (set by setRobotState's velocity arguments)
        self._gravity = 9.80665

    # ---- iDynTree-style model queries
    def model(self):
        return self._km

    def isValid(self):
        return True

    def getNrOfDegreesOfFreedom(self):
        return self.n

    def getFrameIndex(self, name):
        try:
            return self._km.frame_index(name)
        except (KeyError, ValueError):
            return -1

    def getFrameName(self, index):
        return self._km.frame_names[index]

    def getFloatingBase(self):
        return self._km.frame_names[next(i for i, l in enumerate(self._km.frame_link) if l == 0)]

    def getFrameVelocityRepresentation(self):
        return "MIXED_REPRESENTATION"

    # ---- state
    def setRobotState(self, world_T_base, s, base_velocity=None, s_dot=None, world_gravity=None, instance=None):
        """``world_T_base`` 4x4 (or [N,4,4]), ``s`` [n] (or [N,n]); ``base_velocity`` [6] (MIXED: linear and angular velocity of the
        base in world coordinates), ``s_dot`` [n] and ``world_gravity`` [3] (its norm is used, along -z) are what the dynamic
        queries below read; the IK path does not need them.  ``instance``: one robot of the batch (default: all of them)."""
        H = np.asarray(world_T_base, dtype=np.float64)
        s = np.asarray(s, dtype=np.float64)
        if H.shape[-2:] != (4, 4) or s.shape[-1] != self.n:
            log().error("[KinDynComputations::setRobotState] Wrong size of the base transform or of the joint vector.")
            return False
        sel = slice(None) if instance is None else instance
        self._p[sel] = H[..., :3, 3]
        self._q[sel] = _rot_to_quat(H[..., :3, :3]) if H.ndim == 3 else _rot_to_quat(H[:3, :3])[0]
        self._s[sel] = s
        if base_velocity is not None:
            self._nu[sel, :6] = np.asarray(base_velocity, dtype=np.float64)
        if s_dot is not None:
            self._nu[sel, 6:] = np.asarray(s_dot, dtype=np.float64)
        if world_gravity is not None:
            self._gravity = float(np.linalg.norm(np.asarray(world_gravity, dtype=np.float64)))
        self._version += 1
        return True

    def set_robot_state_batch(self, base_position, base_quaternion, joint_positions):
        self._p[:] = np.asarray(base_position, dtype=np.float64).reshape(self.N, 3)
        self._q[:] = np.asarray(base_quaternion, dtype=np.float64).reshape(self.N, 4)
        self._s[:] = np.asarray(joint_positions, dtype=np.float64).reshape(self.N, self.n)
        self._version += 1
        return True

    def state(self):
        return self._p, self._q, self._s

    def state_version(self):
        return self._version

    def native_model(self) -> capi.Model:
        if self._model is None:
            self._model = capi.Model(self._km)
        return self._model

    # ---- kinematic queries (whole batch on the GPU; instance selects what is returned)
    def _problem(self):
        if self._kin is None:
            dummy = T.joint_tracking_task(self._km, "state_holder", np.ones(self.n), priority=1, weight=np.ones(self.n))
            self._kin = capi.Problem(self.native_model(), [dummy], self.N, self.device)
        if self._kin_version != self._version:
            self._kin.set_state(self._p, self._q, self._s)
            self._kin_version = self._version
        return self._kin

    def _frame(self, frame):
        idx = frame if isinstance(frame, (int, np.integer)) else self.getFrameIndex(frame)
        if idx < 0 or idx >= len(self._km.frame_names):
            raise ValueError(f"unknown frame {frame!r}")
        return int(idx)

    def getWorldTransform(self, frame, instance=0):
        k = self._problem().kinematics([self._frame(frame)], jac=False, com=False)
        H = np.eye(4)
        H[:3, :3], H[:3, 3] = k["frame_rot"][instance, 0], k["frame_pos"][instance, 0]
        return H

    def getFrameFreeFloatingJacobian(self, frame, instance=0):
        return self._problem().kinematics([self._frame(frame)], jac=True, com=False)["frame_jac"][instance, 0]

    def getCenterOfMassPosition(self, instance=0):
        return self._problem().kinematics([], jac=False, com=True)["com"][instance]

    def getCenterOfMassJacobian(self, instance=0):
        return self._problem().kinematics([], jac=True, com=True)["com_jac"][instance]

    # ---- dynamic quantities the TSID tasks read (iDynTree names; MIXED representation; blfik_dynamics / blfik_bias_forces)
    def getFreeFloatingMassMatrix(self, instance=0):
        return self._problem().dynamics(self._gravity)[0][instance]

    def generalizedGravityForces(self, instance=0):
        return self._problem().dynamics(self._gravity)[1][instance]

    def generalizedBiasForces(self, instance=0):
        """C(q, nu) nu + g(q) (src/TSID/src/BaseDynamicsTask.cpp:178, JointDynamicsTask.cpp:233)"""
        return self._problem().bias_forces(self._nu, (), self._gravity)[0][instance]

    def getFrameBiasAcc(self, frame, instance=0):
        """Jdot nu of the frame, linear then angular (src/TSID/src/SE3Task.cpp:254)"""
        return self._problem().bias_forces(self._nu, [self._frame(frame)], self._gravity)[1][instance, 0]

    def getFrameVel(self, frame, instance=0):
        """J nu of the frame, linear then angular"""
        prob = self._problem()
        prob.bias_forces(self._nu, [self._frame(frame)], self._gravity)
        return prob.last_frame_velocities[instance, 0]

    def getCenterOfMassBiasAcc(self, instance=0):
        return self._problem().bias_forces(self._nu, (), self._gravity)[2][instance]

    def batch_kinematics(self, frames, jac=True, com=True):
        """all robots at once: dict of ``frame_pos [N,F,3]``, ``frame_rot``, ``frame_jac [N,F,6,nv]``, ``com``, ``com_jac``"""
        return self._problem().kinematics([self._frame(f) for f in frames], jac=jac, com=com)
