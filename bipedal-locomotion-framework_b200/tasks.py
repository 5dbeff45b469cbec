"""Task descriptions for the batched QPInverseKinematics problem.

One ``TaskSpec`` per reference task object; the fields are the reference's
parameter-handler keys:

* ``SE3Task``  -- ``frame_name, kp_linear, kp_angular, mask, use_position_exogenous_feedback,
  use_orientation_exogenous_feedback``                              (src/IK/src/SE3Task.cpp:114-214)
* ``CoMTask``  -- ``kp_linear, mask, use_exogenous_feedback``       (src/IK/src/CoMTask.cpp:107-162)
* ``SO3Task``  -- ``frame_name, kp_angular``                        (src/IK/src/SO3Task.cpp:67-151)
* ``R3Task``   -- ``frame_name, kp_linear, mask``                   (src/IK/src/R3Task.cpp:60-176)
* ``AngularMomentumTask`` -- ``mask``                               (src/IK/src/AngularMomentumTask.cpp:67-135)
* ``JointTrackingTask`` -- ``kp``                                   (src/IK/src/JointTrackingTask.cpp:90-112)
* ``JointLimitsTask``   -- ``klim, sampling_time, use_model_limits, upper_limits, lower_limits``
                                                                    (src/IK/src/JointLimitsTask.cpp:136-198)
* ``JointVelocityLimitsTask`` -- ``upper_limits, lower_limits``     (src/IK/src/JointVelocityLimitsTask.cpp:111-124)

``priority`` 0 = hard constraint, 1 = weighted cost with ``weight``
(QPInverseKinematics::addTask, src/IK/src/QPInverseKinematics.cpp:167-249).
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional, Sequence

import numpy as np

from .models import KinematicModel

# numeric ids shared with include/blfik.h
TASK_SE3 = 0
TASK_COM = 1
TASK_SO3 = 2
TASK_JOINT_TRACKING = 3
TASK_JOINT_LIMITS = 4
TASK_JOINT_VELOCITY_LIMITS = 5
TASK_R3 = 6
TASK_DISTANCE = 7
TASK_GRAVITY = 8
TASK_ANGULAR_MOMENTUM = 9

TYPE_NAMES = {TASK_SE3: "SE3Task", TASK_COM: "CoMTask", TASK_SO3: "SO3Task", TASK_JOINT_TRACKING: "JointTrackingTask",
              TASK_JOINT_LIMITS: "JointLimitsTask", TASK_JOINT_VELOCITY_LIMITS: "JointVelocityLimitsTask", TASK_R3: "R3Task",
              TASK_DISTANCE: "DistanceTask", TASK_GRAVITY: "GravityTask", TASK_ANGULAR_MOMENTUM: "AngularMomentumTask"}


def _vec3(v) -> np.ndarray:
    a = np.atleast_1d(np.asarray(v, dtype=np.float64))
    if a.size == 1:
        a = np.repeat(a, 3)
    if a.size != 3:
        raise ValueError("gain must be a scalar or a 3-vector")
    return a.copy()


@dataclasses.dataclass
class TaskSpec:
    name: str
    type: int
    priority: int = 0
    frame: int = -1
    mask: Sequence[bool] = (True, True, True)
    controller_enabled: bool = True
    kp_linear: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(3))
    kp_angular: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(3))
    kp_joint: Optional[np.ndarray] = None
    weight: Optional[np.ndarray] = None
    klim: Optional[np.ndarray] = None
    upper: Optional[np.ndarray] = None
    lower: Optional[np.ndarray] = None
    sampling_time: float = 0.0
    ref_frame: int = -1          # DistanceTask: reference frame (-1: the world origin)
    threshold: float = 0.001     # DistanceTask: distance_numeric_threshold
    exo_position: bool = False     # SE3Task use_position_exogenous_feedback / CoMTask use_exogenous_feedback
    exo_orientation: bool = False  # SE3Task use_orientation_exogenous_feedback

    def rows(self, n: int) -> int:
        nm = int(sum(bool(m) for m in self.mask))
        return {TASK_SE3: nm + 3, TASK_COM: nm, TASK_R3: nm, TASK_SO3: 3, TASK_DISTANCE: 1, TASK_GRAVITY: 2, TASK_ANGULAR_MOMENTUM: nm, TASK_JOINT_TRACKING: n,
                TASK_JOINT_LIMITS: 2 * n, TASK_JOINT_VELOCITY_LIMITS: 2 * n}[self.type]

    def feedback_len(self) -> int:
        """doubles of exogenous feedback appended to every set-point record: SE3 p(3) q(4), CoM p(3)"""
        if self.type == TASK_SE3 and (self.exo_position or self.exo_orientation):
            return 7
        if self.type == TASK_COM and self.exo_position:
            return 3
        return 0

    def record_len(self, n: int) -> int:
        return self.setpoint_len(n) + self.feedback_len()

    def setpoint_len(self, n: int) -> int:
        return {TASK_SE3: 13, TASK_COM: 6, TASK_SO3: 7, TASK_R3: 6, TASK_DISTANCE: 1, TASK_GRAVITY: 6, TASK_ANGULAR_MOMENTUM: 3, TASK_JOINT_TRACKING: 2 * n,
                TASK_JOINT_LIMITS: 0, TASK_JOINT_VELOCITY_LIMITS: 0}[self.type]

    @property
    def is_inequality(self) -> bool:
        return self.type in (TASK_JOINT_LIMITS, TASK_JOINT_VELOCITY_LIMITS)


def se3_task(model: KinematicModel, name: str, frame_name: str, kp_linear, kp_angular, priority=0,
             weight=None, mask=(True, True, True), use_position_exogenous_feedback=False,
             use_orientation_exogenous_feedback=False) -> TaskSpec:
    t = TaskSpec(name, TASK_SE3, priority, model.frame_index(frame_name), tuple(mask), True, _vec3(kp_linear), _vec3(kp_angular))
    t.exo_position, t.exo_orientation = bool(use_position_exogenous_feedback), bool(use_orientation_exogenous_feedback)
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def com_task(model: KinematicModel, name: str, kp_linear, priority=0, weight=None, mask=(True, True, True),
             use_exogenous_feedback=False) -> TaskSpec:
    t = TaskSpec(name, TASK_COM, priority, -1, tuple(mask), True, _vec3(kp_linear))
    t.exo_position = bool(use_exogenous_feedback)
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def so3_task(model: KinematicModel, name: str, frame_name: str, kp_angular, priority=1, weight=None) -> TaskSpec:
    t = TaskSpec(name, TASK_SO3, priority, model.frame_index(frame_name), (True, True, True), True, np.zeros(3), _vec3(kp_angular))
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def r3_task(model: KinematicModel, name: str, frame_name: str, kp_linear, priority=0, weight=None, mask=(True, True, True)) -> TaskSpec:
    t = TaskSpec(name, TASK_R3, priority, model.frame_index(frame_name), tuple(mask), True, _vec3(kp_linear))
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def distance_task(model: KinematicModel, name: str, target_frame_name: str, kp: float, priority=1, weight=None,
                  reference_frame_name: str = "", distance_numeric_threshold: float = 0.001) -> TaskSpec:
    """keys: kp, target_frame_name, reference_frame_name (optional), distance_numeric_threshold (src/IK/src/DistanceTask.cpp:95-175)."""
    t = TaskSpec(name, TASK_DISTANCE, priority, model.frame_index(target_frame_name), (True, True, True), True, _vec3(kp))
    t.ref_frame = model.frame_index(reference_frame_name) if reference_frame_name else -1
    t.threshold = float(distance_numeric_threshold)
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def gravity_task(model: KinematicModel, name: str, target_frame_name: str, kp: float, priority=1, weight=None) -> TaskSpec:
    """keys: kp, target_frame_name (src/IK/src/GravityTask.cpp:76-147)."""
    t = TaskSpec(name, TASK_GRAVITY, priority, model.frame_index(target_frame_name), (True, True, True), True, _vec3(kp))
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def angular_momentum_task(model: KinematicModel, name: str, priority=1, weight=None, mask=(True, True, True)) -> TaskSpec:
    """key: mask (optional); A = rows 3..5 of the centroidal momentum matrix, b = desired angular momentum
    (src/IK/src/AngularMomentumTask.cpp:67-185).  No gain: the set-point is the momentum itself."""
    t = TaskSpec(name, TASK_ANGULAR_MOMENTUM, priority, -1, tuple(mask), True)
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def joint_tracking_task(model: KinematicModel, name: str, kp, priority=1, weight=None) -> TaskSpec:
    kp = np.asarray(kp, dtype=np.float64)
    if kp.size != model.n:
        raise ValueError("kp size must equal the number of DoFs")  # JointTrackingTask.cpp:100-110
    t = TaskSpec(name, TASK_JOINT_TRACKING, priority, kp_joint=kp.copy())
    t.weight = None if weight is None else np.asarray(weight, dtype=np.float64).copy()
    return t


def joint_limits_task(model: KinematicModel, name: str, klim, sampling_time: float, use_model_limits=True,
                      upper_limits=None, lower_limits=None) -> TaskSpec:
    klim = np.asarray(klim, dtype=np.float64)
    if klim.size != model.n or np.any(klim < 0) or np.any(klim > 1):
        raise ValueError("klim must have n entries in [0, 1]")  # JointLimitsTask.cpp:150-166
    if sampling_time <= 0:
        raise ValueError("sampling_time must be positive")  # JointLimitsTask.cpp:168-176
    up = model.q_upper if use_model_limits else np.asarray(upper_limits, dtype=np.float64)
    lo = model.q_lower if use_model_limits else np.asarray(lower_limits, dtype=np.float64)
    return TaskSpec(name, TASK_JOINT_LIMITS, 0, klim=klim.copy(), upper=up.copy(), lower=lo.copy(), sampling_time=float(sampling_time))


def joint_velocity_limits_task(model: KinematicModel, name: str, upper_limits, lower_limits) -> TaskSpec:
    up = np.asarray(upper_limits, dtype=np.float64)
    lo = np.asarray(lower_limits, dtype=np.float64)
    if up.size != model.n or lo.size != model.n:
        raise ValueError("limits must have n entries")
    return TaskSpec(name, TASK_JOINT_VELOCITY_LIMITS, 0, upper=up.copy(), lower=lo.copy())


def validate_tasks(model: KinematicModel, tasks: List[TaskSpec]) -> None:
    """addTask rules of QPInverseKinematics.cpp:167-249."""
    names = set()
    for t in tasks:
        if t.name in names:
            raise ValueError(f"task '{t.name}' already exists")
        names.add(t.name)
        if t.priority not in (0, 1):
            raise ValueError("priority must be 0 or 1")
        if t.is_inequality and t.priority != 0:
            raise ValueError("inequality tasks can only be hard constraints")
        if t.priority == 1:
            if t.weight is None or np.asarray(t.weight).size != t.rows(model.n):
                raise ValueError(f"task '{t.name}': a priority-1 task needs a weight of size {t.rows(model.n)}")
