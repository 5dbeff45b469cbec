"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE ONLY -- see blfik_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  PARITY UNPINNED (no golden vectors exist in the
reference tree; see the header of blfik_oracle.h).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libblfik_oracle.so")
_lib = None

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class OrcModel(C.Structure):
    _fields_ = [("n", C.c_int), ("nf", C.c_int), ("parent", c_int_p), ("dof", c_int_p),
                ("Xr", c_double_p), ("Xp", c_double_p), ("axis", c_double_p), ("mass", c_double_p),
                ("com", c_double_p), ("frame_link", c_int_p), ("Fr", c_double_p), ("Fp", c_double_p),
                ("inertia", c_double_p)]


class OrcTask(C.Structure):
    _fields_ = [("type", C.c_int), ("priority", C.c_int), ("frame", C.c_int), ("mask", C.c_int * 3),
                ("controller_enabled", C.c_int), ("kp_lin", C.c_double * 3), ("kp_ang", C.c_double * 3),
                ("kp_joint", c_double_p), ("weight", c_double_p), ("klim", c_double_p),
                ("upper", c_double_p), ("lower", c_double_p), ("sampling_time", C.c_double),
                ("ref_frame", C.c_int), ("threshold", C.c_double), ("exo_pos", C.c_int), ("exo_ori", C.c_int)]


class OrcSettings(C.Structure):
    _fields_ = [("rho", C.c_double), ("sigma", C.c_double), ("alpha", C.c_double),
                ("eps_abs", C.c_double), ("eps_rel", C.c_double), ("max_iter", C.c_int),
                ("check_termination", C.c_int), ("adaptive_rho_interval", C.c_int), ("polish", C.c_int),
                ("delta", C.c_double), ("polish_refine_iter", C.c_int), ("warm_start", C.c_int)]


class OrcInfo(C.Structure):
    _fields_ = [("iters", C.c_int), ("status", C.c_int), ("polished", C.c_int), ("rho_updates", C.c_int),
                ("r_prim", C.c_double), ("r_dual", C.c_double)]


def build(force: bool = False) -> str:
    src = [os.path.join(_HERE, f) for f in ("blfik_oracle.c", "blfik_oracle.h", "Makefile")]
    if force or not os.path.exists(_LIB_PATH) or any(
            os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src if os.path.exists(s)):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_qp_kkt.restype = C.c_double
    return _lib


def _dp(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(c_double_p)


def _ip(a: np.ndarray):
    return a.ctypes.data_as(c_int_p)


class Oracle:
    """Single-robot reference semantics evaluated over a batch on the CPU."""

    def __init__(self, model, tasks: Sequence = ()):  # model: blf_b200.models.KinematicModel
        self.L = lib()
        self.model = model
        self.n, self.nv = model.n, model.n + 6
        self._keep = dict(
            parent=np.ascontiguousarray(model.parent, dtype=np.int32), dof=np.ascontiguousarray(model.dof, dtype=np.int32),
            Xr=np.ascontiguousarray(model.Xr, dtype=np.float64), Xp=np.ascontiguousarray(model.Xp, dtype=np.float64),
            axis=np.ascontiguousarray(model.axis, dtype=np.float64), mass=np.ascontiguousarray(model.mass, dtype=np.float64),
            com=np.ascontiguousarray(model.com, dtype=np.float64), fl=np.ascontiguousarray(model.frame_link, dtype=np.int32),
            Fr=np.ascontiguousarray(model.Fr, dtype=np.float64), Fp=np.ascontiguousarray(model.Fp, dtype=np.float64),
            I6=np.ascontiguousarray(np.stack([model.inertia[:, 0, 0], model.inertia[:, 0, 1], model.inertia[:, 0, 2], model.inertia[:, 1, 1],
                                              model.inertia[:, 1, 2], model.inertia[:, 2, 2]], axis=1), dtype=np.float64))
        k = self._keep
        self.cm = OrcModel(model.n, len(model.frame_names), _ip(k["parent"]), _ip(k["dof"]), _dp(k["Xr"]), _dp(k["Xp"]),
                           _dp(k["axis"]), _dp(k["mass"]), _dp(k["com"]), _ip(k["fl"]), _dp(k["Fr"]), _dp(k["Fp"]), _dp(k["I6"]))
        self.set_tasks(tasks)

    # ---- tasks
    def set_tasks(self, tasks: Sequence) -> None:
        self.tasks = list(tasks)
        self._tkeep = []
        arr = (OrcTask * max(1, len(self.tasks)))()
        for i, t in enumerate(self.tasks):
            def keep(a):
                if a is None:
                    return None
                b = np.ascontiguousarray(a, dtype=np.float64)
                self._tkeep.append(b)
                return _dp(b)
            ct = arr[i]
            ct.type, ct.priority, ct.frame = int(t.type), int(t.priority), int(t.frame)
            for c in range(3):
                ct.mask[c] = int(bool(t.mask[c]))
                ct.kp_lin[c] = float(t.kp_linear[c])
                ct.kp_ang[c] = float(t.kp_angular[c])
            ct.controller_enabled = int(bool(t.controller_enabled))
            ct.kp_joint, ct.weight, ct.klim = keep(t.kp_joint), keep(t.weight), keep(t.klim)
            ct.upper, ct.lower = keep(t.upper), keep(t.lower)
            ct.sampling_time = float(t.sampling_time)
            ct.ref_frame = int(getattr(t, 'ref_frame', -1))
            ct.threshold = float(getattr(t, 'threshold', 0.001))
            ct.exo_pos = int(bool(getattr(t, 'exo_position', False)))
            ct.exo_ori = int(bool(getattr(t, 'exo_orientation', False)))
        self.ctasks = arr
        self.m = sum(t.rows(self.n) for t in self.tasks if t.priority == 0)

    # ---- kinematics
    def fk(self, p, q, s):
        p, q, s = (np.ascontiguousarray(a, dtype=np.float64) for a in (p, q, s))
        R = np.empty((self.n + 1, 3, 3)); P = np.empty((self.n + 1, 3))
        self.L.orc_fk(C.byref(self.cm), _dp(p), _dp(q), _dp(s), _dp(R), _dp(P))
        return R, P

    def frame_pose(self, R, P, frame: int):
        Rf = np.empty((3, 3)); pf = np.empty(3)
        self.L.orc_frame_pose(C.byref(self.cm), _dp(R), _dp(P), int(frame), _dp(Rf), _dp(pf))
        return Rf, pf

    def frame_jacobian(self, R, P, frame: int):
        J = np.empty((6, self.nv))
        self.L.orc_frame_jacobian(C.byref(self.cm), _dp(R), _dp(P), int(frame), _dp(J))
        return J

    def com(self, R, P):
        c = np.empty(3); J = np.empty((3, self.nv))
        self.L.orc_com(C.byref(self.cm), _dp(R), _dp(P), _dp(c), _dp(J))
        return c, J

    def angular_momentum_matrix(self, R, P):
        A = np.empty((3, self.nv))
        self.L.orc_angular_momentum_matrix(C.byref(self.cm), _dp(R), _dp(P), _dp(A))
        return A

    # ---- tasks / assembly
    def task_update(self, ti: int, R, P, s, setpoint):
        t = self.tasks[ti]
        rows = t.rows(self.n)
        A = np.empty((rows, self.nv)); b = np.empty(rows)
        s = np.ascontiguousarray(s, dtype=np.float64)
        sp = None if setpoint is None else np.ascontiguousarray(setpoint, dtype=np.float64)
        r = self.L.orc_task_update(C.byref(self.cm), C.byref(self.ctasks[ti]), _dp(R), _dp(P), _dp(s), _dp(sp), _dp(A), _dp(b))
        assert r == rows
        return A, b

    def assemble(self, p, q, s, setpoints: Sequence[Optional[np.ndarray]]):
        p, q, s = (np.ascontiguousarray(a, dtype=np.float64) for a in (p, q, s))
        nv, m = self.nv, max(self.m, 1)
        H = np.empty((nv, nv)); g = np.empty(nv); Cm = np.zeros((m, nv)); l = np.zeros(m); u = np.zeros(m)
        sps = [None if sp is None else np.ascontiguousarray(sp, dtype=np.float64) for sp in setpoints]
        ptrs = (c_double_p * max(1, len(sps)))(*[_dp(sp) for sp in sps])
        rows = self.L.orc_assemble(C.byref(self.cm), self.ctasks, len(self.tasks), ptrs, _dp(p), _dp(q), _dp(s),
                                   _dp(H), _dp(g), _dp(Cm), _dp(l), _dp(u))
        assert rows == self.m
        return H, g, Cm[:rows], l[:rows], u[:rows]

    # ---- QP
    @staticmethod
    def settings(parity: bool = True, **kw) -> OrcSettings:
        st = OrcSettings()
        (lib().orc_parity_settings if parity else lib().orc_default_settings)(C.byref(st))
        for k, v in kw.items():
            setattr(st, k, v)
        return st

    def qp_solve(self, H, g, Cm, l, u, st: Optional[OrcSettings] = None, x0=None, y0=None, z0=None):
        st = st or self.settings()
        nv, m = H.shape[0], Cm.shape[0]
        H, g, Cm, l, u = (np.ascontiguousarray(a, dtype=np.float64) for a in (H, g, Cm, l, u))
        x = np.zeros(nv) if x0 is None else np.array(x0, dtype=np.float64)
        y = np.zeros(max(m, 1)) if y0 is None else np.array(y0, dtype=np.float64)
        z = np.zeros(max(m, 1)) if z0 is None else np.array(z0, dtype=np.float64)
        info = OrcInfo()
        self.L.orc_qp_solve(nv, m, _dp(H), _dp(g), _dp(Cm), _dp(l), _dp(u), C.byref(st), _dp(x), _dp(y), _dp(z), C.byref(info))
        return x, y[:m], z[:m], info

    def qp_kkt(self, H, g, Cm, l, u, x, y) -> float:
        H, g, Cm, l, u, x, y = (np.ascontiguousarray(a, dtype=np.float64) for a in (H, g, Cm, l, u, x, y))
        return float(self.L.orc_qp_kkt(H.shape[0], Cm.shape[0], _dp(H), _dp(g), _dp(Cm), _dp(l), _dp(u), _dp(x), _dp(y)))

    # ---- integration
    def integrate(self, p, q, s, nu, dt: float):
        p, q, s = (np.array(a, dtype=np.float64) for a in (p, q, s))
        nu = np.ascontiguousarray(nu, dtype=np.float64)
        self.L.orc_integrate(self.n, _dp(p), _dp(q), _dp(s), _dp(nu), C.c_double(dt))
        return p, q, s

    # ---- whole tick over a batch (state arrays are advanced in place when dt_integrate > 0)
    def tick_batch(self, base_pos, base_quat, s, setpoints, st: Optional[OrcSettings] = None, warm=None,
                   dt_integrate: float = 0.0, nticks: int = 1, nthreads: int = 1):
        st = st or self.settings()
        N = base_pos.shape[0]
        for a in (base_pos, base_quat, s):
            assert a.dtype == np.float64 and a.flags.c_contiguous
        sps = [None if sp is None else np.ascontiguousarray(sp, dtype=np.float64) for sp in setpoints]
        ptrs = (c_double_p * max(1, len(sps)))(*[_dp(sp) for sp in sps])
        nu = np.zeros((N, self.nv)); status = np.zeros(N, dtype=np.int32); iters = np.zeros(N, dtype=np.int32)
        wx = wy = wz = None
        if warm is not None:
            wx, wy, wz = warm
        self.L.orc_tick_batch_mt(C.byref(self.cm), self.ctasks, len(self.tasks), ptrs, C.byref(st), C.c_long(N),
                                 _dp(base_pos), _dp(base_quat), _dp(s), _dp(nu), _dp(wx), _dp(wy), _dp(wz),
                                 _ip(status), _ip(iters), C.c_double(dt_integrate), int(nticks), int(nthreads))
        return nu, status, iters

    def new_warm(self, N: int):
        m = max(self.m, 1)
        return np.zeros((N, self.nv)), np.zeros((N, m)), np.zeros((N, m))
