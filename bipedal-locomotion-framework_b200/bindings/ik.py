"""bindings/python/IK/src/*.cpp: the IK tasks, ``IntegrationBasedIKState``, ``QPInverseKinematics`` (``initialize``,
``add_task``, ``set_task_weight``, ``finalize``, ``advance``, ``get_output``, ``is_output_valid``, ``get_task_names``,
``get_task``, ``get_raw_solution``, ``build``) and ``QPFixedBaseInverseKinematics``.

A task object only records its configuration and set-points (per robot); ``QPInverseKinematics.finalize`` turns the task
list into one ``blfik_problem`` and ``advance`` is one launch of the tick kernel for the whole batch.  Parameter keys,
method names, defaults and the "return False + log" behaviour follow the C++ classes (file:line in each docstring)."""
from __future__ import annotations

from datetime import timedelta

import numpy as np

from .. import capi, tasks as T
from . import manif
from ._log import log
from .system import ConstantWeightProvider, VariablesHandler, WeightProvider


class IKLinearTask:
    """src/IK/include/BipedalLocomotion/IK/IKLinearTask.h, System/LinearTask.h"""

    class Type:
        equality, inequality = "equality", "inequality"

    _sp_len = 0

    def __init__(self):
        self._kin_dyn = None
        self._initialized = False
        self._set_point_set = False
        self._velocity_variable = ""
        self._sp = None
        self._sp_version = 0
        self._gains_version = 0
        self._fb = None            # exogenous feedback [N, _fb_len] (tasks created with a use_*_exogenous_feedback key)
        self._fb_version = 0
        self._description = "Generic Linear Task Element"

    def set_kin_dyn(self, kin_dyn):
        if kin_dyn is None or not kin_dyn.isValid():
            log().error("[%s::setKinDyn] Invalid kinDyn object.", type(self).__name__)
            return False
        self._kin_dyn = kin_dyn
        return True

    def _common_initialize(self, param_handler):
        name = type(self).__name__
        self._initialized = False
        if self._kin_dyn is None:
            log().error("[%s::initialize] KinDynComputations object is not valid.", name)
            return False
        if param_handler is None:
            log().error("[%s::initialize] The parameter handler is not valid.", name)
            return False
        v = param_handler._opt(param_handler.get_parameter_string, "robot_velocity_variable_name")
        if v is None:
            log().error("[%s::initialize] Error while retrieving the robot velocity variable.", name)
            return False
        self._velocity_variable = v
        return True

    def set_variables_handler(self, variables_handler: VariablesHandler):
        name = type(self).__name__
        if not self._initialized:
            log().error("[%s::setVariablesHandler] The task is not initialized. Please call initialize method.", name)
            return False
        var = variables_handler.get_variable(self._velocity_variable)
        if not var.is_valid():
            log().error("[%s::setVariablesHandler] Unable to get the variable named %s.", name, self._velocity_variable)
            return False
        if var.size != self._kin_dyn.getNrOfDegreesOfFreedom() + 6:
            log().error("[%s::setVariablesHandler] The size of the robot velocity variable is different from the one expected. "
                        "Expected size: %d. Given size: %d.", name, self._kin_dyn.getNrOfDegreesOfFreedom() + 6, var.size)
            return False
        return True

    def size(self):
        raise NotImplementedError

    def type(self):
        return IKLinearTask.Type.equality

    def is_valid(self):
        return self._initialized and (self._sp_len == 0 or self._set_point_set)

    def get_description(self):
        return self._description

    # ---- batch plumbing
    def _store(self, record, instance=None):
        """record: [len] (every robot, or robot `instance`) or [N, len]"""
        N = self._kin_dyn.N
        rec = np.asarray(record, dtype=np.float64)
        if self._sp is None:
            self._sp = np.zeros((N, self._sp_len))
        if rec.ndim == 2:
            if rec.shape != (N, self._sp_len):
                raise ValueError(f"expected a set-point batch of shape {(N, self._sp_len)}, got {rec.shape}")
            self._sp[:] = rec
        elif instance is None:
            self._sp[:] = rec
        else:
            self._sp[instance] = rec
        self._sp_version += 1
        self._set_point_set = True
        return True

    _spline_ok = False         # SE3Task, CoMTask, R3Task: the origin of the set-point can follow a quintic spline on the device
    _min_jerk_ok = None        # None: as _spline_ok

    def _spline_dim(self):
        """coordinates of that spline: 3, or the number of joints for a JointTrackingTask (QuinticSpline over R^n)"""
        return 3

    def set_set_point_spline(self, spline, positions=None, boundary=None, cubic=False):
        """Batch extension (no counterpart in the reference, where a Math::QuinticSpline is evaluated on the host and its
        point passed to setSetPoint every period): the origin of this task's set-point follows a quintic spline evaluated on
        the device at the solver's time (QPInverseKinematics.set_time; blfik_set_setpoint_spline).  ``spline`` is a
        ``math.QuinticSpline`` or ``math.CubicSpline`` (every robot the same trajectory), a list of N of them with the same
        knot times, or the knot times [n_knots] together with ``positions`` [N, n_knots, 3] and ``boundary`` [N, 4, 3]
        ([N, 2, 3] with ``cubic=True``); None switches it off."""
        name = type(self).__name__
        if not self._spline_ok:
            log().error("[%s::setSetPointSpline] The task has no position set-point a spline could drive.", name)
            return False
        if spline is None:
            self._spline, self._spline_version = None, self._spline_version + 1
            return True
        if self._kin_dyn is None:
            log().error("[%s::setSetPointSpline] Please call setKinDyn first.", name)
            return False
        N, dim = self._kin_dyn.N, self._spline_dim()
        if hasattr(spline, "knots") or (isinstance(spline, (list, tuple)) and len(spline) and hasattr(spline[0], "knots")):
            group = [spline] * N if hasattr(spline, "knots") else list(spline)
            if len(group) != N:
                log().error("[%s::setSetPointSpline] One spline per robot is expected. Expected: %d. Given: %d.", name, N, len(group))
                return False
            parts = [g.knots() for g in group]
            if any(k[0] is None or k[1] is None or k[1].shape[1] != dim for k in parts):
                log().error("[%s::setSetPointSpline] Every spline needs its knots set, in R^%d.", name, dim)
                return False
            if any(len(k[0]) != len(parts[0][0]) or np.any(k[0] != parts[0][0]) for k in parts):
                log().error("[%s::setSetPointSpline] The splines of the robots must share their knot times.", name)
                return False
            if any(g._cubic != group[0]._cubic for g in group):
                log().error("[%s::setSetPointSpline] The splines of the robots must be of one kind.", name)
                return False
            cubic = group[0]._cubic
            times = parts[0][0]
            positions = np.stack([k[1] for k in parts])
            boundary = np.stack([np.zeros((4, dim)) if k[2] is None else k[2] for k in parts])
            if cubic:
                boundary = np.ascontiguousarray(boundary[:, [0, 2]])     # velocities only
        else:
            times = np.asarray(spline, dtype=np.float64).reshape(-1)
            positions = np.asarray(positions, dtype=np.float64)
            boundary = None if boundary is None else np.asarray(boundary, dtype=np.float64)
            if positions.shape != (N, len(times), dim) or (boundary is not None and boundary.shape != (N, 2 if cubic else 4, dim)):
                log().error("[%s::setSetPointSpline] positions must be [N, n_knots, 3] and boundary [N, 4, 3] ([N, 2, 3] for a cubic spline).", name)
                return False
        if len(times) < 2:
            log().error("[Spline::setKnots] The number of knots has to be at least 2.")
            return False
        if np.any(np.diff(times) <= 0.0):
            log().error("[Splines::computePhasesDuration] Two consecutive points have the same time coordinate, or the times decrease.")
            return False
        self._spline, self._spline_version = (times.copy(), positions.copy(), None if boundary is None else boundary.copy(), bool(cubic)), self._spline_version + 1
        return True

    _spline = None
    _spline_version = 0

    def set_set_point_minimum_jerk(self, target, duration):
        """Batch extension (the reference evaluates a Planners::SO3Planner / a two-knot QuinticSpline on the host every period and
        passes the point to setSetPoint): from the stored set-point to ``target`` -- set-point records [N, len] or one [len] for
        every robot, of which the pose part is read -- along the minimum-jerk profile over ``duration`` (seconds or timedelta),
        evaluated on the device from the solver's time at the next advance() on (blfik_set_setpoint_minimum_jerk; equals
        ``planners.SO3PlannerInertial`` with zero boundary conditions for the rotation).  SE3 / SO3 / CoM / R3 tasks; a
        non-positive duration switches it off."""
        name = type(self).__name__
        if not ((self._spline_ok if self._min_jerk_ok is None else self._min_jerk_ok) or isinstance(self, SO3Task)):
            log().error("[%s::setSetPointMinimumJerk] The task has no pose set-point such a generator could drive.", name)
            return False
        T_ = duration.total_seconds() if hasattr(duration, "total_seconds") else float(duration)
        if T_ <= 0:
            self._min_jerk, self._min_jerk_version = None, self._min_jerk_version + 1
            return True
        N = self._kin_dyn.N
        rec = np.asarray(target, dtype=np.float64)
        if rec.ndim == 1:
            rec = np.tile(rec, (N, 1))
        if rec.shape != (N, self._sp_len):
            log().error("[%s::setSetPointMinimumJerk] The target must be a set-point record. Expected shape: %s. Given: %s.", name, (N, self._sp_len), rec.shape)
            return False
        self._min_jerk, self._min_jerk_version = (np.ascontiguousarray(rec), T_), self._min_jerk_version + 1
        return True

    _min_jerk = None
    _min_jerk_version = 0

    _fb_len = 0
    _fb_init = ()

    def _store_feedback(self, values, first, instance=None):
        """values: [k] (every robot, or robot `instance`) or [N, k], written at column `first` of the feedback record"""
        N = self._kin_dyn.N
        if self._fb is None:
            self._fb = np.tile(np.asarray(self._fb_init, dtype=np.float64), (N, 1))
        v = np.asarray(values, dtype=np.float64)
        sel = slice(None) if (instance is None or v.ndim == 2) else instance
        self._fb[sel, first:first + v.shape[-1]] = v
        self._fb_version += 1
        return True

    def _spec(self, name, priority, weight) -> T.TaskSpec:
        raise NotImplementedError


def _gain3(h, key):
    v = h._opt(h.get_parameter_float, key)
    if v is not None:
        return np.full(3, v)
    v = h._opt(h.get_parameter_vector_float, key)
    if v is not None and len(v) == 3:
        return np.asarray(v, dtype=np.float64)
    return None


def _mask(h):
    m = h._opt(h.get_parameter_vector_bool, "mask")
    if m is None:
        return (True, True, True)
    if len(m) != 3:
        return None
    return tuple(bool(x) for x in m)


def _vec(v, n):
    a = np.asarray(v, dtype=np.float64)
    if a.shape[-1] != n:
        raise ValueError(f"expected {n} components, got {a.shape[-1]}")
    return a


def _quat_of(R):
    if isinstance(R, (manif.SO3, manif.SE3)):
        return R.quat()
    return np.asarray(R, dtype=np.float64)


class SE3Task(IKLinearTask):
    """SE3Task.cpp:114-214 (keys frame_name, kp_linear, kp_angular, mask), :231-321 (update), :251-258 (controller mode)"""
    _sp_len = 13
    _spline_ok = True

    class Mode:
        Enable, Disable = "Enable", "Disable"

    def __init__(self):
        super().__init__()
        self._mode = SE3Task.Mode.Enable

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        frame = param_handler._opt(param_handler.get_parameter_string, "frame_name")
        self._frame = -1 if frame is None else self._kin_dyn.getFrameIndex(frame)
        if self._frame < 0:
            log().error("[SE3Task::initialize] Error while retrieving the frame that should be controlled.")
            return False
        self._kp_lin, self._kp_ang = _gain3(param_handler, "kp_linear"), _gain3(param_handler, "kp_angular")
        if self._kp_lin is None or self._kp_ang is None:
            log().error("[SE3Task::initialize] Unable to get the gains kp_linear / kp_angular.")
            return False
        self._mask3 = _mask(param_handler)
        if self._mask3 is None:
            log().error("[SE3Task::initialize] The size of the mask must be 3.")
            return False
        # optional, default false (SE3Task.cpp:196-214)
        self._exo_pos = bool(param_handler._opt(param_handler.get_parameter_bool, "use_position_exogenous_feedback") or False)
        self._exo_ori = bool(param_handler._opt(param_handler.get_parameter_bool, "use_orientation_exogenous_feedback") or False)
        self._fb_len = 7 if (self._exo_pos or self._exo_ori) else 0
        self._fb_init = (0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0)   # the controllers' initial state: p = 0, q = identity
        self._description = f"SE3Task Optimal Control Element - Frame name: {frame}."
        self._initialized = True
        return True

    def set_feedback(self, I_H_F=None, I_p_F=None, I_R_F=None, instance=None):
        """SE3Task.cpp:338-366: measured pose (or only its position / rotation) for the controllers; a part whose
        use_*_exogenous_feedback key is off is ignored.  Batch form: I_p_F [N,3], I_R_F [N,4] quaternions (x, y, z, w)."""
        if isinstance(I_H_F, manif.SE3):
            I_p_F, I_R_F = I_H_F.translation(), I_H_F.asSO3()
        elif I_H_F is not None:
            pose = np.asarray(I_H_F, dtype=np.float64)
            I_p_F, I_R_F = pose[..., :3], pose[..., 3:7]
        if self._kin_dyn is None:
            return False
        if I_p_F is not None and self._exo_pos:
            self._store_feedback(_vec(I_p_F, 3), 0, instance)
        if I_R_F is not None and self._exo_ori:
            self._store_feedback(_vec(_quat_of(I_R_F), 4), 3, instance)
        return True

    def set_set_point(self, I_H_F, mixed_velocity=None, instance=None):
        if isinstance(I_H_F, manif.SE3):
            v = np.zeros(6) if mixed_velocity is None else (mixed_velocity.coeffs() if isinstance(mixed_velocity, manif.SE3Tangent) else _vec(mixed_velocity, 6))
            return self._store(np.concatenate([I_H_F.translation(), I_H_F.quat(), v]), instance)
        return self._store(I_H_F, instance)    # raw [N, 13] batch: p(3) q(4) v(3) w(3)

    def set_task_controller_mode(self, mode):
        self._mode = mode
        self._gains_version += 1

    def get_task_controller_mode(self):
        return self._mode

    def size(self):
        return sum(self._mask3) + 3

    def _spec(self, name, priority, weight):
        t = T.se3_task(self._kin_dyn.model(), name, self._kin_dyn.getFrameName(self._frame), self._kp_lin, self._kp_ang, priority, weight, self._mask3)
        t.controller_enabled = self._mode == SE3Task.Mode.Enable
        t.exo_position, t.exo_orientation = self._exo_pos, self._exo_ori
        return t


class CoMTask(IKLinearTask):
    """CoMTask.cpp:107-162 (keys kp_linear, mask), :171-235"""
    _sp_len = 6
    _spline_ok = True

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        self._kp_lin = _gain3(param_handler, "kp_linear")
        if self._kp_lin is None:
            log().error("[CoMTask::initialize] Unable to get the gain kp_linear.")
            return False
        self._mask3 = _mask(param_handler)
        if self._mask3 is None:
            log().error("[CoMTask::initialize] The size of the mask must be 3.")
            return False
        self._exo_pos = bool(param_handler._opt(param_handler.get_parameter_bool, "use_exogenous_feedback") or False)   # CoMTask.cpp:116-123
        self._fb_len = 3 if self._exo_pos else 0
        self._fb_init = (0.0, 0.0, 0.0)
        self._description = "CoM Task"
        self._initialized = True
        return True

    def set_feedback(self, feedback, instance=None):
        """CoMTask.cpp:262-273: measured CoM; ignored unless use_exogenous_feedback"""
        if self._kin_dyn is None:
            return False
        if self._exo_pos:
            self._store_feedback(_vec(feedback, 3), 0, instance)
        return True

    def set_set_point(self, position, velocity=None, instance=None):
        position = np.asarray(position, dtype=np.float64)
        if position.shape[-1] == 6:
            return self._store(position, instance)
        velocity = np.zeros_like(position) if velocity is None else np.asarray(velocity, dtype=np.float64)
        return self._store(np.concatenate([position, velocity], axis=-1), instance)

    def size(self):
        return sum(self._mask3)

    def _spec(self, name, priority, weight):
        return T.com_task(self._kin_dyn.model(), name, self._kp_lin, priority, weight, self._mask3, use_exogenous_feedback=self._exo_pos)


class SO3Task(IKLinearTask):
    """SO3Task.cpp:67-151 (keys frame_name, kp_angular), :153-192"""
    _sp_len = 7

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        frame = param_handler._opt(param_handler.get_parameter_string, "frame_name")
        self._frame = -1 if frame is None else self._kin_dyn.getFrameIndex(frame)
        if self._frame < 0:
            log().error("[SO3Task::initialize] Error while retrieving the frame that should be controlled.")
            return False
        self._kp_ang = _gain3(param_handler, "kp_angular")
        if self._kp_ang is None:
            log().error("[SO3Task::initialize] Unable to get the gain kp_angular.")
            return False
        self._description = f"SO3Task Optimal Control Element - Frame name: {frame}."
        self._initialized = True
        return True

    def set_set_point(self, I_R_F, angular_velocity=None, instance=None):
        if isinstance(I_R_F, manif.SO3):
            w = np.zeros(3) if angular_velocity is None else (angular_velocity.coeffs() if isinstance(angular_velocity, manif.SO3Tangent) else _vec(angular_velocity, 3))
            return self._store(np.concatenate([I_R_F.quat(), w]), instance)
        return self._store(I_R_F, instance)

    def size(self):
        return 3

    def _spec(self, name, priority, weight):
        return T.so3_task(self._kin_dyn.model(), name, self._kin_dyn.getFrameName(self._frame), self._kp_ang, priority, weight)


class R3Task(IKLinearTask):
    """R3Task.cpp:60-176 (keys frame_name, kp_linear, mask), :178-243"""
    _sp_len = 6
    _spline_ok = True

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        frame = param_handler._opt(param_handler.get_parameter_string, "frame_name")
        self._frame = -1 if frame is None else self._kin_dyn.getFrameIndex(frame)
        if self._frame < 0:
            log().error("[R3Task::initialize] Error while retrieving the frame that should be controlled.")
            return False
        self._kp_lin = _gain3(param_handler, "kp_linear")
        if self._kp_lin is None:
            log().error("[R3Task::initialize] Unable to get the gain kp_linear.")
            return False
        self._mask3 = _mask(param_handler)
        if self._mask3 is None:
            log().error("[R3Task::initialize] The size of the mask must be 3.")
            return False
        self._initialized = True
        return True

    def set_set_point(self, I_p_F, velocity=None, instance=None):
        p = np.asarray(I_p_F, dtype=np.float64)
        if p.shape[-1] == 6:
            return self._store(p, instance)
        v = np.zeros_like(p) if velocity is None else np.asarray(velocity, dtype=np.float64)
        return self._store(np.concatenate([p, v], axis=-1), instance)

    def size(self):
        return sum(self._mask3)

    def _spec(self, name, priority, weight):
        return T.r3_task(self._kin_dyn.model(), name, self._kin_dyn.getFrameName(self._frame), self._kp_lin, priority, weight, self._mask3)


class DistanceTask(IKLinearTask):
    """DistanceTask.cpp:95-175 (keys kp, target_frame_name, reference_frame_name, distance_numeric_threshold), :177-243"""
    _sp_len = 1

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        h = param_handler
        self._kp = h._opt(h.get_parameter_float, "kp")
        if self._kp is None:
            log().error("[DistanceTask::initialize] Failed to get the proportional linear gain.")
            return False
        self._ref_name = h._opt(h.get_parameter_string, "reference_frame_name") or ""
        self._target_name = h._opt(h.get_parameter_string, "target_frame_name")
        if self._target_name is None:
            log().error("[DistanceTask::initialize] Failed to get the end effector frame name.")
            return False
        thr = h._opt(h.get_parameter_float, "distance_numeric_threshold")
        self._threshold = 0.001 if thr is None else thr
        if self._ref_name and self._kin_dyn.getFrameIndex(self._ref_name) < 0:
            log().error("[DistanceTask::initialize] The specified base name (%s) does not seem to exist.", self._ref_name)
            return False
        if self._kin_dyn.getFrameIndex(self._target_name) < 0:
            log().error("[DistanceTask::initialize] The specified target name (%s) does not seem to exist.", self._target_name)
            return False
        self._initialized = True
        return True

    def set_desired_distance(self, desired_distance, instance=None):
        d = np.abs(np.asarray(desired_distance, dtype=np.float64))   # DistanceTask.cpp:240
        return self._store(d.reshape(-1, 1) if d.ndim >= 1 and d.size > 1 else d.reshape(1), instance)

    def set_set_point(self, desired_distance, instance=None):
        return self.set_desired_distance(desired_distance, instance)

    def size(self):
        return 1

    def _spec(self, name, priority, weight):
        return T.distance_task(self._kin_dyn.model(), name, self._target_name, self._kp, priority, weight, self._ref_name, self._threshold)


class GravityTask(IKLinearTask):
    """GravityTask.cpp:76-147 (keys kp, target_frame_name), :149-250"""
    _sp_len = 6

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        h = param_handler
        self._kp = h._opt(h.get_parameter_float, "kp")
        if self._kp is None:
            log().error("[GravityTask::initialize] Failed to get the proportional linear gain.")
            return False
        self._target_name = h._opt(h.get_parameter_string, "target_frame_name")
        if self._target_name is None or self._kin_dyn.getFrameIndex(self._target_name) < 0:
            log().error("[GravityTask::initialize] Failed to get the target frame name.")
            return False
        self._initialized = True
        return True

    def _part(self, v, first, instance):
        if self._sp is None:
            self._sp = np.zeros((self._kin_dyn.N, 6))
            self._sp[:, 2] = 1.0      # m_desiredZDirectionBody starts at e_z (GravityTask.cpp:71)
        v = np.asarray(v, dtype=np.float64)
        sel = slice(None) if (instance is None or v.ndim == 2) else instance
        self._sp[sel, first:first + 3] = v
        self._sp_version += 1
        self._set_point_set = True
        return True

    def set_desired_gravity_direction_in_target_frame(self, desired_gravity_direction, instance=None):
        g = np.asarray(desired_gravity_direction, dtype=np.float64)
        nrm = np.linalg.norm(g, axis=-1, keepdims=True)
        return self._part(g / np.where(nrm > 0, nrm, 1.0), 0, instance)   # normalised, GravityTask.cpp:204-206

    def set_feedforward_velocity_in_target_frame(self, feedforward_velocity, instance=None):
        return self._part(feedforward_velocity, 3, instance)

    def set_set_point(self, desired_gravity_direction, feedforward_velocity=(0.0, 0.0, 0.0), instance=None):
        return self.set_desired_gravity_direction_in_target_frame(desired_gravity_direction, instance) and \
            self.set_feedforward_velocity_in_target_frame(feedforward_velocity, instance)

    def size(self):
        return 2

    def _spec(self, name, priority, weight):
        return T.gravity_task(self._kin_dyn.model(), name, self._target_name, self._kp, priority, weight)


class AngularMomentumTask(IKLinearTask):
    """AngularMomentumTask.cpp:67-135 (key mask, optional), :137-210"""
    _sp_len = 3

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        m = param_handler._opt(param_handler.get_parameter_vector_bool, "mask")
        self._mask3 = tuple(bool(x) for x in m) if (m is not None and len(m) == 3) else (True, True, True)
        self._description = "Angular momentum task."
        self._initialized = True
        return True

    def set_set_point(self, desired_angular_momentum, instance=None):
        return self._store(_vec(desired_angular_momentum, 3), instance)

    def size(self):
        return sum(self._mask3)

    def _spec(self, name, priority, weight):
        return T.angular_momentum_task(self._kin_dyn.model(), name, priority, weight, self._mask3)


class JointTrackingTask(IKLinearTask):
    """JointTrackingTask.cpp:90-112 (key kp, size-checked), :127-155, setSetPoint :157-210"""

    _spline_ok = True          # the joint positions can follow a spline over R^n evaluated on the device
    _min_jerk_ok = False

    def _spline_dim(self):
        return self._kin_dyn.getNrOfDegreesOfFreedom()

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        kp = param_handler._opt(param_handler.get_parameter_vector_float, "kp")
        n = self._kin_dyn.getNrOfDegreesOfFreedom()
        if kp is None:
            log().error("[JointTrackingTask::initialize] Error while retrieving the proportional gain.")
            return False
        if len(kp) != n:
            log().error("[JointTrackingTask::initialize] The size of the kp gain does not match with the number of actuated joints. "
                        "Expected: %d. Given: %d.", n, len(kp))
            return False
        self._kp = np.asarray(kp, dtype=np.float64)
        self._sp_len = 2 * n
        self._initialized = True
        return True

    def set_set_point(self, joint_position, joint_velocity=None, instance=None):
        n = self._kin_dyn.getNrOfDegreesOfFreedom()
        s = np.asarray(joint_position, dtype=np.float64)
        if s.shape[-1] == 2 * n:
            return self._store(s, instance)
        v = np.zeros_like(s) if joint_velocity is None else np.asarray(joint_velocity, dtype=np.float64)
        if s.shape[-1] != n or v.shape[-1] != n:
            log().error("[JointTrackingTask::setSetPoint] Wrong size of the desired reference trajectory: Expected size: %d. "
                        "Joint position size: %d. Joint velocity size: %d.", n, s.shape[-1], v.shape[-1])
            return False
        return self._store(np.concatenate([s, v], axis=-1), instance)

    def size(self):
        return self._kin_dyn.getNrOfDegreesOfFreedom()

    def _spec(self, name, priority, weight):
        return T.joint_tracking_task(self._kin_dyn.model(), name, self._kp, priority, weight)


class JointLimitsTask(IKLinearTask):
    """JointLimitsTask.cpp:136-198 (keys klim, sampling_time, use_model_limits, upper_limits, lower_limits), :211-259"""

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        h, n, prefix = param_handler, self._kin_dyn.getNrOfDegreesOfFreedom(), "[JointLimitsTask::initialize]"
        klim = h._opt(h.get_parameter_vector_float, "klim")
        if klim is None or len(klim) != n or np.any(np.asarray(klim) < 0) or np.any(np.asarray(klim) > 1):
            log().error("%s 'klim' must have one entry in [0, 1] per joint.", prefix)
            return False
        dt = h._opt(h.get_parameter_datetime, "sampling_time")
        if dt is None or dt <= timedelta(0):
            log().error("%s The sampling time must be a strictly positive number.", prefix)
            return False
        use_model = h._opt(h.get_parameter_bool, "use_model_limits")
        if use_model is None:
            log().error("%s Unable to get the parameter 'use_model_limits'.", prefix)   # mandatory, :178-183
            return False
        up = lo = None
        if not use_model:
            up, lo = h._opt(h.get_parameter_vector_float, "upper_limits"), h._opt(h.get_parameter_vector_float, "lower_limits")
            if up is None or lo is None or len(up) != n or len(lo) != n:
                log().error("%s Unable to get the joint limits 'upper_limits' / 'lower_limits' (one entry per joint).", prefix)
                return False
        self._klim, self._dt, self._use_model, self._up, self._lo = np.asarray(klim, dtype=np.float64), dt.total_seconds(), use_model, up, lo
        self._initialized = True
        return True

    def size(self):
        return 2 * self._kin_dyn.getNrOfDegreesOfFreedom()

    def type(self):
        return IKLinearTask.Type.inequality

    def _spec(self, name, priority, weight):
        return T.joint_limits_task(self._kin_dyn.model(), name, self._klim, self._dt, self._use_model, self._up, self._lo)


class JointVelocityLimitsTask(IKLinearTask):
    """JointVelocityLimitsTask.cpp:111-124 (keys upper_limits, lower_limits), :134-145"""

    def initialize(self, param_handler):
        if not self._common_initialize(param_handler):
            return False
        h, n = param_handler, self._kin_dyn.getNrOfDegreesOfFreedom()
        up, lo = h._opt(h.get_parameter_vector_float, "upper_limits"), h._opt(h.get_parameter_vector_float, "lower_limits")
        if up is None or lo is None or len(up) != n or len(lo) != n:
            log().error("[JointVelocityLimitsTask::initialize] Unable to get 'upper_limits' / 'lower_limits' (one entry per joint).")
            return False
        self._up, self._lo = np.asarray(up, dtype=np.float64), np.asarray(lo, dtype=np.float64)
        self._initialized = True
        return True

    def size(self):
        return 2 * self._kin_dyn.getNrOfDegreesOfFreedom()

    def type(self):
        return IKLinearTask.Type.inequality

    def _spec(self, name, priority, weight):
        return T.joint_velocity_limits_task(self._kin_dyn.model(), name, self._up, self._lo)


class IntegrationBasedIKState:
    """IntegrationBasedIK.h:27-33"""

    def __init__(self, n=0):
        self.joint_velocity = np.zeros(n)
        self.base_velocity = manif.SE3Tangent()


_TASK_TYPES = {c.__name__: c for c in (SE3Task, CoMTask, SO3Task, R3Task, DistanceTask, GravityTask, AngularMomentumTask, JointTrackingTask,
                                        JointLimitsTask, JointVelocityLimitsTask)}


class QPInverseKinematics:
    """QPInverseKinematics.cpp:156-806"""

    def __init__(self):
        self._initialized = self._finalized = self._output_valid = False
        self._velocity_variable = ""
        self._tasks = []          # dicts: task, name, priority, provider, uploaded weight / set-point / gains versions
        self._problem = None
        self._state = IntegrationBasedIKState()
        self._nu = self._status = None

    def initialize(self, handler):
        if handler is None:
            log().error("[QPInverseKinematics::initialize] The parameter handler is not valid.")
            return False
        v = handler._opt(handler.get_parameter_string, "robot_velocity_variable_name")
        if v is None:
            log().error("[QPInverseKinematics::initialize] Unable to get the parameter 'robot_velocity_variable_name'.")
            return False
        self._velocity_variable, self._initialized = v, True
        return True

    def _entry(self, name):
        return next((e for e in self._tasks if e["name"] == name), None)

    def add_task(self, task, task_name, priority, weight=None, weight_provider=None):
        prefix = "[QPInverseKinematics::addTask]"
        if not self._initialized:
            log().error("%s Please call initialize() before addTask().", prefix)
            return False
        if self._entry(task_name) is not None:
            log().error("%s The task named %s already exist.", prefix, task_name)
            return False
        if priority not in (0, 1):
            log().error("%s - [Task name: '%s'] For the time being we support only priority equal to 0 or 1.", prefix, task_name)
            return False
        provider = weight_provider
        if weight is not None:
            provider = ConstantWeightProvider(weight)
        if priority == 1:
            if provider is None:
                log().error("%s - [Task name: '%s'] Invalid weight provider. For priority equal to 1 the weight is mandatory.", prefix, task_name)
                return False
            if task.type() == IKLinearTask.Type.inequality:
                log().error("%s - [Task name: '%s'] This implementation of the inverse kinematics cannot handle inequality tasks with "
                            "priority equal to 1.", prefix, task_name)
                return False
            if len(provider.get_output()) != task.size():
                log().error("%s - [Task name: '%s'] The size of the weight is not coherent with the size of the task. Expected: %d. "
                            "Given: %d.", prefix, task_name, task.size(), len(provider.get_output()))
                return False
        if not isinstance(task, tuple(_TASK_TYPES.values())):
            # the reference accepts any IKLinearTask subclass (a "custom task" supplies A and b itself); the batched kernels
            # only know the task kinds of src/IK
            if not getattr(task, "is_valid", lambda: False)():
                log().error("%s - [Task name: '%s'] The task is not valid.", prefix, task_name)
                return False
            self._tasks.append(dict(task=task, name=task_name, priority=priority, provider=None, custom=True))
            self._finalized = False
            return True
        self._tasks.append(dict(task=task, name=task_name, priority=priority, provider=provider if priority == 1 else None, custom=False,
                                sp_version=-1, gains_version=-1, weight=None))
        self._finalized = False   # finalize() must be called again (:245-246)
        return True

    def set_task_weight(self, task_name, weight=None, weight_provider=None):
        prefix = "[QPInverseKinematics::setTaskWeight]"
        e = self._entry(task_name)
        if e is None:
            log().error("%s The task named %s does not exist.", prefix, task_name)
            return False
        if e["priority"] != 1:
            log().error("%s - [Task name: '%s'] The weight can be set only to tasks with priority equal to 1.", prefix, task_name)
            return False
        provider = ConstantWeightProvider(weight) if weight is not None else weight_provider
        if provider is None:
            log().error("%s - [Task name: '%s'] Invalid weight provider.", prefix, task_name)
            return False
        if len(provider.get_output()) != e["task"].size():
            log().error("%s - [Task name: '%s'] The size of the weight is not coherent with the size of the task.", prefix, task_name)
            return False
        e["provider"] = provider
        return True

    def get_task_weight_provider(self, task_name):
        e = self._entry(task_name)
        return None if e is None else e["provider"]

    def get_task_names(self):
        return [e["name"] for e in self._tasks]

    def get_task(self, name):
        e = self._entry(name)
        return None if e is None else e["task"]

    def finalize(self, handler: VariablesHandler):
        prefix = "[QPInverseKinematics::finalize]"
        if not self._initialized:
            log().error("%s Please call initialize() before finalize().", prefix)
            return False
        var = handler.get_variable(self._velocity_variable)
        if not var.is_valid():
            log().error("%s Unable to find the variable named %s in the variables handler.", prefix, self._velocity_variable)
            return False
        for e in self._tasks:
            if not e["task"].set_variables_handler(handler):
                log().error("%s Error while setting the variable handler in the task named %s.", prefix, e["name"])
                return False
        native = [e for e in self._tasks if not e["custom"]]
        if self._problem is not None:
            self._problem.close()
            self._problem = None
        if native:
            kd = native[0]["task"]._kin_dyn
            if handler.get_number_of_variables() != kd.getNrOfDegreesOfFreedom() + 6:
                log().error("%s The batched solver optimises the robot velocity only: the handler must hold that one variable.", prefix)
                return False
            specs = [e["task"]._spec(e["name"], e["priority"], None if e["provider"] is None else e["provider"].get_output()) for e in native]
            try:
                T.validate_tasks(kd.model(), specs)
                self._problem = capi.Problem(kd.native_model(), specs, kd.N, kd.device)
            except (ValueError, capi.BlfikError) as err:
                log().error("%s %s", prefix, err)
                return False
            for e, sp in zip(native, specs):
                e["sp_version"], e["gains_version"], e["weight"] = -1, e["task"]._gains_version, None if sp.weight is None else np.array(sp.weight)
                e["fb_version"] = e["spline_version"] = e["min_jerk_version"] = 0
            self._kin_dyn, self._state_version = kd, -1
        self._finalized = True
        return True

    def advance(self):
        prefix = "[QPInverseKinematics::advance]"
        self._output_valid = False
        if not self._finalized:
            log().error("%s Please call finalize() before advance().", prefix)   # :421-425
            return False
        native = [e for e in self._tasks if not e["custom"]]
        if any(e["custom"] for e in self._tasks) or self._problem is None:
            log().error("%s Tasks that supply their own A and b cannot be solved by the batched kernels.", prefix)
            return False
        prob, kd = self._problem, self._kin_dyn
        for t, e in enumerate(native):
            task = e["task"]
            if task._sp_len and not task._set_point_set:
                log().error("%s Unable to update the task named %s: the set-point has not been set at least once.", prefix, e["name"])   # :429-439
                return False
            if task._sp_len and e["sp_version"] != task._sp_version:
                prob.set_setpoint(t, task._sp)
                e["sp_version"] = task._sp_version
            if e.get("min_jerk_version", 0) != task._min_jerk_version:
                try:
                    if task._min_jerk is None:
                        prob.set_setpoint_minimum_jerk(t, None, 0.0)
                    else:
                        prob.set_setpoint_minimum_jerk(t, *task._min_jerk)
                except capi.BlfikError as err:
                    log().error("%s Unable to set the minimum-jerk generator of the task named %s: %s", prefix, e["name"], err)
                    return False
                e["min_jerk_version"] = task._min_jerk_version
            if task._spline_ok and e.get("spline_version", 0) != task._spline_version:
                try:
                    if task._spline is None:
                        prob.set_setpoint_spline(t, None, None)
                    else:
                        prob.set_setpoint_spline(t, *task._spline)
                except capi.BlfikError as err:
                    log().error("%s Unable to set the spline of the task named %s: %s", prefix, e["name"], err)
                    return False
                e["spline_version"] = task._spline_version
            if task._fb_len and task._fb is not None and e.get("fb_version", 0) != task._fb_version:
                prob.set_feedback(t, task._fb)
                e["fb_version"] = task._fb_version
            if e["provider"] is not None:
                w = np.asarray(e["provider"].get_output(), dtype=np.float64)   # read on every tick (:450-451)
                if e["weight"] is None or not np.array_equal(w, e["weight"]):
                    prob.set_weight(t, w)
                    e["weight"] = w.copy()
            if e["gains_version"] != task._gains_version:
                if isinstance(task, SE3Task):
                    prob.set_controller_mode(t, task._mode == SE3Task.Mode.Enable)
                e["gains_version"] = task._gains_version
        if self._state_version != kd.state_version():
            prob.set_state(*kd.state())
            self._state_version = kd.state_version()
        try:
            prob.tick(1, 0.0, integrate=False, warm_start=True)
            self._nu = prob.get_output()
            self._status, _ = prob.get_status()
        except capi.BlfikError as err:
            log().error("%s %s", prefix, err)
            return False
        failed = int(np.sum(self._status == 2))   # Solved and SolvedInaccurate are accepted (:518-523)
        if failed and kd.N == 1:
            log().error("%s osqp was not able to find a feasible solution.", prefix)
            return False
        if failed:   # a batch is not all-or-nothing: the call succeeded, each robot answers for itself (is_output_valid(i))
            log().warning("%s osqp was not able to find a feasible solution for %d of %d robots.", prefix, failed, kd.N)
        self._state.base_velocity = manif.SE3Tangent(self._nu[0, :6])
        self._state.joint_velocity = self._nu[0, 6:].copy()
        self._output_valid = True
        return True

    def get_output(self):
        return self._state

    def is_output_valid(self, instance=None):
        """without an argument: the reference's isOutputValid(); with a robot index: that robot's last solution is usable"""
        if instance is None:
            return self._output_valid
        return bool(self._output_valid and self._status is not None and self._status[int(instance)] != 2)

    def get_raw_solution(self):
        return None if self._nu is None else self._nu[0]

    # ---- batch extensions
    def set_time(self, t):
        """time at which the on-device set-point generators (set_set_point_spline) are evaluated by the next advance()"""
        if self._problem is None:
            log().error("[QPInverseKinematics::setTime] Please call finalize() first.")
            return False
        self._problem.set_time(t.total_seconds() if hasattr(t, "total_seconds") else float(t))
        return True

    def get_batch_output(self):
        """[N, 6 + n] velocities of the last advance()"""
        return self._nu

    def get_batch_status(self):
        return self._status

    def __str__(self):
        return "====== QPInverseKinematics class ======\n" + "\n".join(
            f"\t - {e['name']}: {e['task'].get_description()} Priority: {e['priority']}." for e in self._tasks)

    @staticmethod
    def build(param_handler, kin_dyn):
        """QPInverseKinematics.cpp:649-806: ``tasks`` (names of the task groups), group ``IK``, per-task group with ``type``,
        ``priority``, ``weight_provider_type`` (optional) + the provider's keys.  Returns (variables_handler, weights, ik);
        ik is None on failure."""
        prefix = "[QPInverseKinematics::build]"
        fail = (VariablesHandler(), {}, None)
        try:
            ik_group, names = param_handler.get_group("IK"), param_handler.get_parameter_vector_string("tasks")
        except ValueError as err:
            log().error("%s %s", prefix, err)
            return fail
        solver = QPInverseKinematics()
        if not solver.initialize(ik_group):
            return fail
        variables = VariablesHandler()
        try:
            if not variables.initialize(param_handler.get_group("VARIABLES")):
                return fail
        except ValueError:
            variables.add_variable(solver._velocity_variable, kin_dyn.getNrOfDegreesOfFreedom() + 6)
        weights = {}
        for gname in names:
            try:
                g = param_handler.get_group(gname)
                ttype, priority = g.get_parameter_string("type"), g.get_parameter_int("priority")
            except ValueError as err:
                log().error("%s %s", prefix, err)
                return fail
            if ttype not in _TASK_TYPES:
                log().error("%s The task type '%s' is not registered.", prefix, ttype)
                return fail
            task = _TASK_TYPES[ttype]()
            if not task.set_kin_dyn(kin_dyn) or not task.initialize(g):
                log().error("%s Unable to initialize the task named '%s'.", prefix, gname)
                return fail
            provider = None
            if priority == 1:
                ptype = g._opt(g.get_parameter_string, "weight_provider_type") or "ConstantWeightProvider"
                if ptype != "ConstantWeightProvider":
                    log().error("%s The weight provider '%s' is not available from Python.", prefix, ptype)
                    return fail
                provider = ConstantWeightProvider()
                if not provider.initialize(g):
                    return fail
                weights[gname] = provider
            if not solver.add_task(task, gname, priority, weight_provider=provider):
                return fail
        if not solver.finalize(variables):
            return fail
        return variables, weights, solver


class QPFixedBaseInverseKinematics(QPInverseKinematics):
    """QPFixedBaseInverseKinematics.cpp:17-122: adds "base_se3_task" (priority 0, zero gains, controller disabled, zero
    feed-forward) on the floating-base link."""

    def __init__(self):
        super().__init__()
        self._base_task, self._base_link = SE3Task(), None

    def set_kin_dyn(self, kin_dyn):
        if kin_dyn is None:
            log().error("[QPFixedBaseInverseKinematics::setKinDyn] The kinDyn object is not valid.")
            return False
        if not self._base_task.set_kin_dyn(kin_dyn):
            return False
        self._base_link = kin_dyn.getFloatingBase()
        return True

    def initialize(self, handler):
        prefix = "[QPFixedBaseInverseKinematics::initialize]"
        if self._base_link is None:
            log().error("%s The object kinDyn is not valid.", prefix)
            return False
        if handler is None:
            log().error("%s The parameter handler is not valid.", prefix)
            return False
        v = handler._opt(handler.get_parameter_string, "robot_velocity_variable_name")
        if v is None:
            log().error("%s Unable to get the parameter 'robot_velocity_variable_name'", prefix)
            return False
        from .parameters_handler import StdParametersHandler
        h = StdParametersHandler()
        h.set_parameter_string("robot_velocity_variable_name", v)
        h.set_parameter_float("kp_linear", 0.0)
        h.set_parameter_float("kp_angular", 0.0)
        h.set_parameter_string("frame_name", self._base_link)
        if not self._base_task.initialize(h):
            log().error("%s Error creating SE3task for the base.", prefix)
            return False
        self._base_task.set_task_controller_mode(SE3Task.Mode.Disable)
        self._base_task.set_set_point(manif.SE3.Identity(), manif.SE3Tangent.Zero())
        if not QPInverseKinematics.initialize(self, handler):
            return False
        return self.add_task(self._base_task, "base_se3_task", 0)
