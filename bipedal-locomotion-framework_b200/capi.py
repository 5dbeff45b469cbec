"""ctypes binding of the C ABI in ``include/blfik.h`` (``libblfik_cuda.so``).

This is the only way Python reaches the GPU path: there is no CPU fallback and the
library is loaded eagerly by ``load()``; a missing ``.so`` raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence

import numpy as np

from .models import KinematicModel
from .tasks import TaskSpec

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BLFIK_LIBRARY") or os.path.join(_HERE, "libblfik_cuda.so")   # BLFIK_LIBRARY: an alternative build (A/B timing)
_lib = None

MEM_HOST, MEM_DEVICE = 0, 1
TICK_INTEGRATE, TICK_WARM_START, TICK_NO_FEEDFORWARD = 1, 2, 4
TICK_STRICT = 16
DEBUG_H, DEBUG_G, DEBUG_CEQ, DEBUG_BEQ, DEBUG_LO, DEBUG_HI = range(6)

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int32)

# every symbol include/blfik.h declares
EXPORTS = [
    "blfik_model_create", "blfik_model_destroy", "blfik_problem_create", "blfik_problem_destroy",
    "blfik_default_settings", "blfik_set_solver_settings", "blfik_set_state", "blfik_get_state",
    "blfik_set_setpoint", "blfik_setpoint_len", "blfik_set_weight", "blfik_set_gains",
    "blfik_set_controller_mode", "blfik_tick", "blfik_get_output", "blfik_get_status",
    "blfik_count_unsolved", "blfik_kinematics", "blfik_integrate", "blfik_debug_enable", "blfik_debug_get",
    "blfik_num_dofs", "blfik_num_equality_rows", "blfik_num_instances", "blfik_sync", "blfik_get_stream",
    "blfik_set_stream", "blfik_launch_count", "blfik_last_error", "blfik_measure_fp64_peak", "blfik_kinematics_noalloc",
    "blfik_set_kernel_path", "blfik_get_kernel_path", "blfik_integrate_state", "blfik_tick_host",
    "blfik_packed_record_len", "blfik_packed_offset", "blfik_tick_packed",
    "blfik_set_setpoint_trajectory", "blfik_set_time", "blfik_get_time", "blfik_set_limits",
    "blfik_set_feedback", "blfik_feedback_len", "blfik_record_len", "blfik_set_setpoint_minimum_jerk", "blfik_set_setpoint_spline", "blfik_set_setpoint_cubic_spline", "blfik_evaluate_setpoint",
    "blfik_packed_record_len_ex", "blfik_packed_offset_ex", "blfik_packed_part_len",
    "blfik_set_setpoint_shared", "blfik_kernel_timing", "blfik_get_kernel_times", "blfik_host_alloc", "blfik_host_free",
    "blfik_set_weight_schedule", "blfik_get_weight", "blfik_integrate_acceleration_state", "blfik_dynamics", "blfik_bias_forces",
]


class ModelDesc(C.Structure):
    _fields_ = [("n_dofs", C.c_int32), ("n_frames", C.c_int32), ("parent", c_int_p), ("dof", c_int_p),
                ("X_rot", c_double_p), ("X_pos", c_double_p), ("axis", c_double_p), ("mass", c_double_p),
                ("com", c_double_p), ("frame_link", c_int_p), ("frame_rot", c_double_p), ("frame_pos", c_double_p),
                ("inertia", c_double_p)]


class TaskDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("priority", C.c_int32), ("frame", C.c_int32), ("mask", C.c_int32 * 3),
                ("controller_enabled", C.c_int32), ("kp_linear", C.c_double * 3), ("kp_angular", C.c_double * 3),
                ("kp_joint", c_double_p), ("weight", c_double_p), ("klim", c_double_p),
                ("upper_limits", c_double_p), ("lower_limits", c_double_p), ("sampling_time", C.c_double),
                ("ref_frame", C.c_int32), ("distance_threshold", C.c_double),
                ("use_position_exogenous_feedback", C.c_int32), ("use_orientation_exogenous_feedback", C.c_int32)]


class SolverSettings(C.Structure):
    _fields_ = [("rho", C.c_double), ("sigma", C.c_double), ("alpha", C.c_double), ("eps", C.c_double),
                ("admm_iters", C.c_int32), ("max_rounds", C.c_int32), ("polish_iters", C.c_int32),
                ("eq_iters", C.c_int32)]


class BlfikError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"blfik error {code}: {msg}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile the CUDA library in-tree (nvcc, sm_100a)."""
    srcdir = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(srcdir, f) for f in os.listdir(srcdir)] + [os.path.join(_HERE, "..", "include", "blfik.h")]
    if force or not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs):
        subprocess.check_call(["make", "-C", srcdir, "-s", "NVFLAGS=-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC"])
    return LIB_PATH


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                    "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.blfik_last_error.restype = C.c_char_p
        _lib.blfik_get_stream.restype = C.c_void_p
        _lib.blfik_get_stream.argtypes = [C.c_void_p]
        _lib.blfik_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        _lib.blfik_launch_count.restype = C.c_int64
        _lib.blfik_launch_count.argtypes = [C.c_void_p]
        _lib.blfik_num_instances.restype = C.c_int64
        for name in ("blfik_num_instances", "blfik_num_dofs", "blfik_num_equality_rows", "blfik_problem_destroy",
                     "blfik_model_destroy", "blfik_sync"):
            getattr(_lib, name).argtypes = [C.c_void_p]
        _lib.blfik_problem_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_void_p]
        _lib.blfik_tick.argtypes = [C.c_void_p, C.c_int32, C.c_double, C.c_uint32]
        _lib.blfik_integrate.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_double, C.c_int32]
        _lib.blfik_set_state.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        _lib.blfik_get_state.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        _lib.blfik_set_setpoint.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
        _lib.blfik_get_output.argtypes = [C.c_void_p, C.c_void_p, C.c_int32]
        _lib.blfik_get_status.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        _lib.blfik_debug_get.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
        _lib.blfik_debug_enable.argtypes = [C.c_void_p, C.c_int32]
        _lib.blfik_kinematics.argtypes = [C.c_void_p, C.c_void_p, C.c_int32] + [C.c_void_p] * 5 + [C.c_int32]
        _lib.blfik_kinematics_noalloc.argtypes = _lib.blfik_kinematics.argtypes
        _lib.blfik_set_weight.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
        _lib.blfik_set_gains.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.blfik_set_controller_mode.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
        _lib.blfik_count_unsolved.argtypes = [C.c_void_p, C.c_void_p]
        _lib.blfik_set_solver_settings.argtypes = [C.c_void_p, C.c_void_p]
        _lib.blfik_setpoint_len.argtypes = [C.c_void_p, C.c_int32]
        _lib.blfik_set_kernel_path.argtypes = [C.c_void_p, C.c_int32]
        _lib.blfik_get_kernel_path.argtypes = [C.c_void_p]
    return _lib


def integrate_acceleration_state(p, q, s, dp, omega, ds, acc, dt: float, nsteps: int = 1, device: int = 0):
    """FloatingBaseSystemAccelerationKinematics + ForwardEuler for a batch; returns the new (p, q, s, dp, omega, ds), the
    arguments are left alone"""
    lib = load()
    N, n = s.shape
    arrs = [np.array(a, dtype=np.float64, order="C", copy=True) for a in (p, q, s, dp, omega, ds, acc)]
    lib.blfik_integrate_acceleration_state.argtypes = [C.c_int32, C.c_int64] + [C.c_void_p] * 7 + [C.c_double, C.c_int32, C.c_int32]
    _check(lib.blfik_integrate_acceleration_state(int(n), int(N), *[a.ctypes.data for a in arrs], float(dt), int(nsteps), int(device)))
    return arrs[:6]


def measure_fp64_peak(device: int = 0, millis: float = 200.0) -> float:
    """Sustained DFMA throughput of ``device`` in TFLOP/s (roofline denominator)."""
    out = C.c_double(0.0)
    lib = load()
    lib.blfik_measure_fp64_peak.argtypes = [C.c_int32, C.c_double, C.c_void_p]
    _check(lib.blfik_measure_fp64_peak(int(device), float(millis), C.byref(out)))
    return float(out.value)


def _check(rc: int) -> None:
    if rc != 0:
        raise BlfikError(rc, load().blfik_last_error().decode())


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a):
    """numpy array / torch tensor / int device pointer / None -> (void*, is_device)."""
    if a is None:
        return None, False
    if isinstance(a, int):
        return C.c_void_p(a), True
    if isinstance(a, np.ndarray):
        assert a.dtype in (np.float64, np.int32) and a.flags.c_contiguous
        return C.c_void_p(a.ctypes.data), False
    # torch tensor (plumbing only: device memory + pinned host memory)
    assert a.is_contiguous()
    return C.c_void_p(a.data_ptr()), bool(a.is_cuda)


class Model:
    """Device-independent kinematic model handle (``blfik_model``)."""

    def __init__(self, km: KinematicModel):
        L = load()
        self.km = km
        k = dict(parent=np.ascontiguousarray(km.parent, dtype=np.int32), dof=np.ascontiguousarray(km.dof, dtype=np.int32),
                 Xr=_f64(km.Xr), Xp=_f64(km.Xp), axis=_f64(km.axis), mass=_f64(km.mass), com=_f64(km.com),
                 fl=np.ascontiguousarray(km.frame_link, dtype=np.int32), Fr=_f64(km.Fr), Fp=_f64(km.Fp),
                 I6=_f64(np.stack([km.inertia[:, 0, 0], km.inertia[:, 0, 1], km.inertia[:, 0, 2], km.inertia[:, 1, 1],
                                   km.inertia[:, 1, 2], km.inertia[:, 2, 2]], axis=1)))
        d = ModelDesc(km.n, len(km.frame_names), k["parent"].ctypes.data_as(c_int_p), k["dof"].ctypes.data_as(c_int_p),
                      k["Xr"].ctypes.data_as(c_double_p), k["Xp"].ctypes.data_as(c_double_p), k["axis"].ctypes.data_as(c_double_p),
                      k["mass"].ctypes.data_as(c_double_p), k["com"].ctypes.data_as(c_double_p), k["fl"].ctypes.data_as(c_int_p),
                      k["Fr"].ctypes.data_as(c_double_p), k["Fp"].ctypes.data_as(c_double_p), k["I6"].ctypes.data_as(c_double_p))
        self.handle = C.c_void_p()
        _check(L.blfik_model_create(C.byref(d), C.byref(self.handle)))

    def __del__(self):
        if getattr(self, "handle", None) and _lib is not None:
            _lib.blfik_model_destroy(self.handle)
            self.handle = None


def task_descs(tasks: Sequence[TaskSpec]):
    keep: List[np.ndarray] = []
    arr = (TaskDesc * max(1, len(tasks)))()

    def dp(a):
        if a is None:
            return None
        b = _f64(a)
        keep.append(b)
        return b.ctypes.data_as(c_double_p)

    for i, t in enumerate(tasks):
        d = arr[i]
        d.type, d.priority, d.frame = int(t.type), int(t.priority), int(t.frame)
        for c in range(3):
            d.mask[c] = int(bool(t.mask[c]))
            d.kp_linear[c] = float(t.kp_linear[c])
            d.kp_angular[c] = float(t.kp_angular[c])
        d.controller_enabled = int(bool(t.controller_enabled))
        d.kp_joint, d.weight, d.klim = dp(t.kp_joint), dp(t.weight), dp(t.klim)
        d.upper_limits, d.lower_limits = dp(t.upper), dp(t.lower)
        d.sampling_time = float(t.sampling_time)
        d.ref_frame = int(getattr(t, "ref_frame", -1))
        d.distance_threshold = float(getattr(t, "threshold", 0.001))
        d.use_position_exogenous_feedback = int(bool(getattr(t, "exo_position", False)))
        d.use_orientation_exogenous_feedback = int(bool(getattr(t, "exo_orientation", False)))
    return arr, keep


class Problem:
    """A batch of N robots sharing one model and one task list on one GPU (``blfik_problem``)."""

    def __init__(self, model: Model, tasks: Sequence[TaskSpec], n_instances: int, device: int = 0):
        L = load()
        self.L = L
        self.model = model
        self.tasks = list(tasks)
        self.N = int(n_instances)
        self.n = model.km.n
        self.nv = self.n + 6
        arr, keep = task_descs(self.tasks)
        self.handle = C.c_void_p()
        _check(L.blfik_problem_create(model.handle, arr, len(self.tasks), self.N, int(device), C.byref(self.handle)))
        self.me = L.blfik_num_equality_rows(self.handle)
        self.device = device

    def close(self):
        if getattr(self, "handle", None):
            self.L.blfik_problem_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- settings
    def set_solver(self, **kw) -> SolverSettings:
        s = SolverSettings()
        self.L.blfik_default_settings(C.byref(s))
        for k, v in kw.items():
            if not hasattr(s, k):
                raise AttributeError(k)
            setattr(s, k, v)
        _check(self.L.blfik_set_solver_settings(self.handle, C.byref(s)))
        return s

    # ---- inputs
    def set_state(self, base_pos, base_quat, s):
        (p, dev), (q, _), (sj, _) = _ptr(base_pos), _ptr(base_quat), _ptr(s)
        _check(self.L.blfik_set_state(self.handle, p, q, sj, MEM_DEVICE if dev else MEM_HOST))

    def get_state(self):
        p, q, s = np.empty((self.N, 3)), np.empty((self.N, 4)), np.empty((self.N, self.n))
        _check(self.L.blfik_get_state(self.handle, p.ctypes.data, q.ctypes.data, s.ctypes.data, MEM_HOST))
        return p, q, s

    def set_setpoint(self, task: int, data):
        ptr, dev = _ptr(data)
        _check(self.L.blfik_set_setpoint(self.handle, int(task), ptr, MEM_DEVICE if dev else MEM_HOST))

    def set_weight_schedule(self, task: int, target, settling_time: float):
        """first-order approach of the task's weight to ``target`` (a = 3 / settling_time), one ForwardEuler step per tick on the
        device side of a multi-tick launch (MultiStateWeightProvider / FirstOrderSmoother); settling_time <= 0: off"""
        tg = None if target is None else _f64(np.asarray(target, dtype=np.float64).reshape(-1))
        self.L.blfik_set_weight_schedule.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_double]
        _check(self.L.blfik_set_weight_schedule(self.handle, int(task), None if tg is None else tg.ctypes.data, float(settling_time)))

    def get_weight(self, task: int, rows: int):
        w = np.empty(rows)
        self.L.blfik_get_weight.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
        _check(self.L.blfik_get_weight(self.handle, int(task), w.ctypes.data))
        return w

    def set_setpoint_shared(self, task: int, record):
        """one set-point for every robot of the batch (a single record of setpoint_len doubles); the task then leaves the packed
        host records.  set_setpoint goes back to one set-point per robot."""
        r = _f64(np.asarray(record, dtype=np.float64).reshape(-1))
        self.L.blfik_set_setpoint_shared.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
        _check(self.L.blfik_set_setpoint_shared(self.handle, int(task), r.ctypes.data, MEM_HOST))

    def kernel_timing(self, enabled: bool = True):
        self.L.blfik_kernel_timing.argtypes = [C.c_void_p, C.c_int32]
        _check(self.L.blfik_kernel_timing(self.handle, int(bool(enabled))))

    def kernel_times(self):
        """(assemble ms, solve ms) of the last full-batch tick on the two-launch path (kernel_timing must be on)"""
        a, b = C.c_double(0.0), C.c_double(0.0)
        self.L.blfik_get_kernel_times.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _check(self.L.blfik_get_kernel_times(self.handle, C.byref(a), C.byref(b)))
        return float(a.value), float(b.value)

    def set_feedback(self, task: int, data):
        """exogenous feedback of an SE3 ([N,7] = p, q) or CoM ([N,3]) task created with the use_*_exogenous_feedback option"""
        ptr, dev = _ptr(data)
        _check(self.L.blfik_set_feedback(self.handle, int(task), ptr, MEM_DEVICE if dev else MEM_HOST))

    def set_inputs(self, inp):
        """``inp``: workloads.Inputs (host numpy)."""
        self.set_state(_f64(inp.base_pos), _f64(inp.base_quat), _f64(inp.s))
        for t, sp in enumerate(inp.setpoints):
            if sp is not None:
                self.set_setpoint(t, _f64(sp))

    def set_setpoint_trajectory(self, task: int, amplitude, period: float, phase):
        """x_d(t) = x_0 + A sin(2 pi t / period + phase_i) on the device; amplitude = [linear(3), angular(3)], phase [N]."""
        a, ph = _f64(np.asarray(amplitude, dtype=np.float64).reshape(6)), _f64(phase)
        self.L.blfik_set_setpoint_trajectory.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_double, C.c_void_p, C.c_int32]
        _check(self.L.blfik_set_setpoint_trajectory(self.handle, int(task), a.ctypes.data, float(period), ph.ctypes.data, MEM_HOST))

    def set_setpoint_minimum_jerk(self, task: int, target, duration: float):
        """minimum-jerk move of the task's set-point from its stored value to ``target`` [N, setpoint_len] over ``duration`` s,
        evaluated on the device (SO3Planner / two-knot QuinticSpline semantics); duration <= 0 switches it off"""
        tg = None if target is None else _f64(target)
        self.L.blfik_set_setpoint_minimum_jerk.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_double, C.c_int32]
        _check(self.L.blfik_set_setpoint_minimum_jerk(self.handle, int(task), None if tg is None else tg.ctypes.data, float(duration), MEM_HOST))

    def set_setpoint_spline(self, task: int, times, positions, boundary=None, cubic=False):
        """quintic spline through ``positions`` [N, n_knots, 3] at ``times`` [n_knots] (the problem's clock) for the linear part
        of an SE3 / CoM / R3 task -- or [N, n_knots, n] for the joint positions of a JointTrackingTask --, evaluated on the device
        at every tick (Math::QuinticSpline semantics); ``boundary``
        [N, 4, 3] = velocity, acceleration at the first knot, velocity, acceleration at the last (None: zero);
        ``times=None`` switches it off.  ``cubic=True``: Math::CubicSpline instead, ``boundary`` [N, 2, 3] = velocity at the
        first and at the last knot"""
        fn = self.L.blfik_set_setpoint_cubic_spline if cubic else self.L.blfik_set_setpoint_spline
        fn.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        if times is None:
            _check(fn(self.handle, int(task), 0, None, None, None, MEM_HOST))
            return
        tm, x = _f64(times).reshape(-1), _f64(positions)
        # a JointTrackingTask takes a joint-space spline (Spline<VectorXd>): n coordinates instead of 3
        joint = 0 <= int(task) < len(self.tasks) and int(self.tasks[int(task)].type) == 3   # BLFIK_TASK_JOINT_TRACKING
        dim = self.model.km.n if joint else 3
        if x.shape != (self.N, len(tm), dim):
            raise ValueError(f"positions must be [N, n_knots, {dim}] = {(self.N, len(tm), dim)}, got {x.shape}")
        bc = None if boundary is None else _f64(boundary)
        if bc is not None and bc.shape != (self.N, 2 if cubic else 4, dim):
            raise ValueError(f"boundary must be [N, {2 if cubic else 4}, {dim}], got {bc.shape}")
        _check(fn(self.handle, int(task), len(tm), tm.ctypes.data, x.ctypes.data, None if bc is None else bc.ctypes.data, MEM_HOST))

    def evaluate_setpoint(self, task: int, time: float) -> np.ndarray:
        """[N, setpoint_len] set-point of ``task`` as a tick at ``time`` reads it (stored set-point + attached generators)"""
        self.L.blfik_setpoint_len.argtypes = [C.c_void_p, C.c_int32]
        self.L.blfik_setpoint_len.restype = C.c_int32
        out = np.empty((self.N, int(self.L.blfik_setpoint_len(self.handle, int(task)))))
        self.L.blfik_evaluate_setpoint.argtypes = [C.c_void_p, C.c_int32, C.c_double, C.c_void_p, C.c_int32]
        _check(self.L.blfik_evaluate_setpoint(self.handle, int(task), float(time), out.ctypes.data, MEM_HOST))
        return out

    def set_time(self, t: float):
        self.L.blfik_set_time.argtypes = [C.c_void_p, C.c_double]
        _check(self.L.blfik_set_time(self.handle, float(t)))

    @property
    def time(self) -> float:
        self.L.blfik_get_time.argtypes = [C.c_void_p]
        self.L.blfik_get_time.restype = C.c_double
        return float(self.L.blfik_get_time(self.handle))

    def set_weight(self, task: int, weight):
        w = _f64(weight)
        _check(self.L.blfik_set_weight(self.handle, int(task), w.ctypes.data))

    def set_gains(self, task: int, kp_linear=None, kp_angular=None, kp_joint=None):
        a = [None if v is None else _f64(v) for v in (kp_linear, kp_angular, kp_joint)]
        _check(self.L.blfik_set_gains(self.handle, int(task), *[None if v is None else v.ctypes.data for v in a]))

    def set_limits(self, task: int, upper=None, lower=None, klim=None, sampling_time: float = 0.0):
        a = [None if v is None else _f64(v) for v in (upper, lower, klim)]
        self.L.blfik_set_limits.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double]
        _check(self.L.blfik_set_limits(self.handle, int(task), *[None if v is None else v.ctypes.data for v in a], float(sampling_time)))

    def set_controller_mode(self, task: int, enabled: bool):
        _check(self.L.blfik_set_controller_mode(self.handle, int(task), int(enabled)))

    # ---- tick
    def tick(self, nticks: int = 1, dt: float = 0.0, integrate: bool = False, warm_start: bool = False, strict: bool = False):
        """``strict``: a tick that ends without a solution inside the iteration budget counts as FAILED (BLFIK_TICK_STRICT)"""
        flags = (TICK_INTEGRATE if integrate else 0) | (TICK_WARM_START if warm_start else 0) | (TICK_STRICT if strict else 0)
        _check(self.L.blfik_tick(self.handle, int(nticks), float(dt), flags))

    def tick_host(self, base_pos, base_quat, s, setpoints, nu, status=None, dt: float = 0.0, integrate: bool = False,
                  warm_start: bool = False, nchunks: int = 8):
        """One tick with HOST inputs and outputs, copies pipelined with compute (``blfik_tick_host``).  Arrays are numpy
        arrays or (pinned) torch CPU tensors; ``setpoints`` is a list with one batch (or None) per task."""
        flags = (TICK_INTEGRATE if integrate else 0) | (TICK_WARM_START if warm_start else 0)

        def hp(a):
            if a is None:
                return None
            ptr, dev = _ptr(a)
            assert not dev, "blfik_tick_host takes host buffers"
            return ptr.value

        arr = (C.c_void_p * max(1, len(self.tasks)))(*[hp(sp) for sp in setpoints]) if setpoints is not None else None
        self.L.blfik_tick_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_double, C.c_uint32, C.c_int32]
        _check(self.L.blfik_tick_host(self.handle, hp(base_pos), hp(base_quat), hp(s), arr, hp(nu), hp(status), float(dt), flags, int(nchunks)))

    # ---- packed host records: p(3) q(4) s(n) then the set-point of every task that has one
    @property
    def packed_record_len(self) -> int:
        self.L.blfik_packed_record_len.argtypes = [C.c_void_p]
        return int(self.L.blfik_packed_record_len(self.handle))

    def packed_offset(self, task: int) -> int:
        self.L.blfik_packed_offset.argtypes = [C.c_void_p, C.c_int32]
        return int(self.L.blfik_packed_offset(self.handle, int(task)))

    def pack_inputs(self, inp, compact: bool = False) -> np.ndarray:
        """``workloads.Inputs`` -> [N][record length] host array.  compact: the layout of BLFIK_TICK_NO_FEEDFORWARD (pose part
        of every set-point only); raises if a feed-forward term that would be dropped is not zero."""
        fl = TICK_NO_FEEDFORWARD if compact else 0
        L = self.L
        L.blfik_packed_record_len_ex.argtypes = [C.c_void_p, C.c_uint32]
        L.blfik_packed_offset_ex.argtypes = [C.c_void_p, C.c_int32, C.c_uint32]
        L.blfik_packed_part_len.argtypes = [C.c_void_p, C.c_int32, C.c_uint32]
        rec = np.zeros((self.N, int(L.blfik_packed_record_len_ex(self.handle, fl))))
        rec[:, 0:3], rec[:, 3:7], rec[:, 7:7 + self.n] = inp.base_pos, inp.base_quat, inp.s
        for t, sp in enumerate(inp.setpoints):
            if sp is None:
                continue
            o = int(L.blfik_packed_offset_ex(self.handle, t, fl))
            if o < 0:      # batch-shared set-point: not part of the records
                continue
            if not compact:
                rec[:, o:o + sp.shape[1]] = sp
                continue
            base = int(L.blfik_packed_part_len(self.handle, t, fl)) - int(L.blfik_feedback_len(self.handle, t))
            full = int(L.blfik_setpoint_len(self.handle, t))
            if np.any(sp[:, base:full] != 0.0):
                raise ValueError(f"task {t}: the compact record drops non-zero feed-forward terms")
            rec[:, o:o + base] = sp[:, :base]
            if sp.shape[1] > full:   # exogenous feedback travels behind the pose part
                rec[:, o + base:o + base + sp.shape[1] - full] = sp[:, full:]
        return rec

    def tick_packed(self, records, nu, status=None, dt: float = 0.0, integrate: bool = False, warm_start: bool = False,
                    nchunks: int = 0, compact: bool = False):
        """One tick from packed HOST records (numpy or pinned torch CPU tensor).  nchunks = 0: zero-copy (the kernel reads
        pinned host memory directly; falls back to staged copies for pageable buffers); nchunks >= 1: that many
        copy / compute / copy pipeline chunks."""
        flags = (TICK_INTEGRATE if integrate else 0) | (TICK_WARM_START if warm_start else 0) | (TICK_NO_FEEDFORWARD if compact else 0)
        (r, d0), (o, d1) = _ptr(records), _ptr(nu)
        st = None if status is None else _ptr(status)[0]
        assert not d0 and not d1, "blfik_tick_packed takes host buffers"
        self.L.blfik_tick_packed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_uint32, C.c_int32]
        _check(self.L.blfik_tick_packed(self.handle, r, o, st, float(dt), flags, int(nchunks)))

    def get_output(self, out=None):
        if out is None:
            out = np.empty((self.N, self.nv))
        ptr, dev = _ptr(out)
        _check(self.L.blfik_get_output(self.handle, ptr, MEM_DEVICE if dev else MEM_HOST))
        return out

    def get_status(self):
        st, it = np.empty(self.N, dtype=np.int32), np.empty(self.N, dtype=np.int32)
        _check(self.L.blfik_get_status(self.handle, st.ctypes.data, it.ctypes.data, MEM_HOST))
        return st, it

    def dynamics(self, gravity: float = 9.80665):
        """free-floating mass matrix [N, nv, nv] and generalised gravity forces [N, nv] of the current state (MIXED)"""
        M, G = np.empty((self.N, self.nv, self.nv)), np.empty((self.N, self.nv))
        self.L.blfik_dynamics.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_int32]
        _check(self.L.blfik_dynamics(self.handle, M.ctypes.data, G.ctypes.data, float(gravity), MEM_HOST))
        return M, G

    def bias_forces(self, nu, frames: Sequence[int] = (), gravity: float = 9.80665):
        """generalised bias forces h = C(q, nu) nu + g(q) [N, nv], the bias accelerations Jdot nu of ``frames`` [N, nframes, 6] and
        of the CoM [N, 3] for the current state and the generalised velocity ``nu`` [N, nv] (MIXED)"""
        v = _f64(nu)
        if v.shape != (self.N, self.nv):
            raise ValueError(f"nu must be [N, nv] = {(self.N, self.nv)}, got {v.shape}")
        fr = np.ascontiguousarray(frames, dtype=np.int32)
        h, fb, cb = np.empty((self.N, self.nv)), np.empty((self.N, len(fr), 6)), np.empty((self.N, 3))
        fv = np.empty((self.N, len(fr), 6))
        self.L.blfik_bias_forces.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        _check(self.L.blfik_bias_forces(self.handle, v.ctypes.data, float(gravity), h.ctypes.data, fr.ctypes.data if len(fr) else None, len(fr),
                                        fb.ctypes.data if len(fr) else None, fv.ctypes.data if len(fr) else None, cb.ctypes.data, MEM_HOST))
        self.last_frame_velocities = fv   # J nu of ``frames``, [N, nframes, 6]
        return h, fb, cb

    def count_unsolved(self) -> int:
        c = C.c_int64(0)
        _check(self.L.blfik_count_unsolved(self.handle, C.byref(c)))
        return int(c.value)

    # ---- stand-alone stages
    def kinematics(self, frames: Sequence[int], jac: bool = True, com: bool = True):
        fr = np.ascontiguousarray(frames, dtype=np.int32)
        nf = len(fr)
        fp, fR = np.empty((self.N, nf, 3)), np.empty((self.N, nf, 3, 3))
        fJ = np.empty((self.N, nf, 6, self.nv)) if jac else None
        c = np.empty((self.N, 3)) if com else None
        cJ = np.empty((self.N, 3, self.nv)) if (com and jac) else None
        _check(self.L.blfik_kinematics(self.handle, fr.ctypes.data if nf else None, nf, fp.ctypes.data if nf else None,
                                       fR.ctypes.data if nf else None, None if fJ is None or not nf else fJ.ctypes.data,
                                       None if c is None else c.ctypes.data, None if cJ is None else cJ.ctypes.data, MEM_HOST))
        return dict(frame_pos=fp, frame_rot=fR, frame_jac=fJ, com=c, com_jac=cJ)

    def integrate(self, nu=None, dt: float = 0.01, nsteps: int = 1):
        ptr, dev = _ptr(None if nu is None else (nu if not isinstance(nu, np.ndarray) else _f64(nu)))
        _check(self.L.blfik_integrate(self.handle, ptr, MEM_DEVICE if dev else MEM_HOST, float(dt), int(nsteps)))

    # ---- introspection
    def debug_enable(self, on: bool = True):
        _check(self.L.blfik_debug_enable(self.handle, int(on)))

    def debug_get(self, what: int) -> np.ndarray:
        shape = {DEBUG_H: (self.N, self.nv, self.nv), DEBUG_G: (self.N, self.nv), DEBUG_CEQ: (self.N, self.me, self.nv),
                 DEBUG_BEQ: (self.N, self.me), DEBUG_LO: (self.N, self.n), DEBUG_HI: (self.N, self.n)}[what]
        out = np.empty(shape)
        _check(self.L.blfik_debug_get(self.handle, int(what), out.ctypes.data, MEM_HOST))
        return out

    def sync(self):
        _check(self.L.blfik_sync(self.handle))

    @property
    def stream(self) -> int:
        return int(self.L.blfik_get_stream(self.handle) or 0)

    def set_stream(self, cuda_stream: int):
        _check(self.L.blfik_set_stream(self.handle, C.c_void_p(cuda_stream)))

    def set_kernel_path(self, path: int):
        """0 = automatic, 1 = force the generic shared-memory kernel."""
        _check(self.L.blfik_set_kernel_path(self.handle, int(path)))

    @property
    def kernel_path(self) -> int:
        """2 = register-resident kernel, 3 = factored-form kernel (33 to 64 velocity variables), 1 = generic kernel."""
        return int(self.L.blfik_get_kernel_path(self.handle))

    @property
    def launch_count(self) -> int:
        return int(self.L.blfik_launch_count(self.handle))
