"""The benchmark / parity workloads of BASELINE.json (configs C1..C5) and their
synthetic inputs.

Inputs are generated for the *global* batch from one counter-based generator
(Philox, seed 20261106 + config id) and sliced by instance range, so that the
data of instance i do not depend on how the batch is sharded across GPUs.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import models as M
from . import tasks as T

SEED = 20261106


@dataclasses.dataclass
class Workload:
    name: str
    model: M.KinematicModel
    tasks: List[T.TaskSpec]
    dt: float
    description: str


@dataclasses.dataclass
class Inputs:
    base_pos: np.ndarray            # [N,3]
    base_quat: np.ndarray           # [N,4] (x,y,z,w)
    s: np.ndarray                   # [N,n]
    setpoints: List[Optional[np.ndarray]]  # per task [N,len] or None

    def slice(self, a: int, b: int) -> "Inputs":
        return Inputs(self.base_pos[a:b].copy(), self.base_quat[a:b].copy(), self.s[a:b].copy(),
                      [None if sp is None else sp[a:b].copy() for sp in self.setpoints])

    @property
    def N(self) -> int:
        return self.base_pos.shape[0]


# ---------------------------------------------------------------- numpy SO(3)
def quat_mul(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw,
                     aw * bw - ax * bx - ay * by - az * bz], axis=-1)


def so3_exp(phi: np.ndarray) -> np.ndarray:
    th = np.linalg.norm(phi, axis=-1, keepdims=True)
    k = np.where(th > 1e-8, np.sin(0.5 * th) / np.where(th > 1e-8, th, 1.0), 0.5 - th * th / 48.0)
    return np.concatenate([k * phi, np.cos(0.5 * th)], axis=-1)


def quat_to_rot(q: np.ndarray) -> np.ndarray:
    x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    R = np.empty(q.shape[:-1] + (3, 3))
    R[..., 0, 0] = 1 - 2 * (y * y + z * z); R[..., 0, 1] = 2 * (x * y - z * w); R[..., 0, 2] = 2 * (x * z + y * w)
    R[..., 1, 0] = 2 * (x * y + z * w); R[..., 1, 1] = 1 - 2 * (x * x + z * z); R[..., 1, 2] = 2 * (y * z - x * w)
    R[..., 2, 0] = 2 * (x * z - y * w); R[..., 2, 1] = 2 * (y * z + x * w); R[..., 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def rot_to_quat(R: np.ndarray) -> np.ndarray:
    """Batched rotation matrix -> unit quaternion (x,y,z,w), w >= 0 branch-free enough for inputs."""
    R = np.asarray(R)
    q = np.empty(R.shape[:-2] + (4,))
    t = np.trace(R, axis1=-2, axis2=-1)
    q[..., 3] = np.sqrt(np.maximum(0.0, 1 + t)) / 2
    q[..., 0] = np.sqrt(np.maximum(0.0, 1 + R[..., 0, 0] - R[..., 1, 1] - R[..., 2, 2])) / 2
    q[..., 1] = np.sqrt(np.maximum(0.0, 1 - R[..., 0, 0] + R[..., 1, 1] - R[..., 2, 2])) / 2
    q[..., 2] = np.sqrt(np.maximum(0.0, 1 - R[..., 0, 0] - R[..., 1, 1] + R[..., 2, 2])) / 2
    q[..., 0] = np.copysign(q[..., 0], R[..., 2, 1] - R[..., 1, 2])
    q[..., 1] = np.copysign(q[..., 1], R[..., 0, 2] - R[..., 2, 0])
    q[..., 2] = np.copysign(q[..., 2], R[..., 1, 0] - R[..., 0, 1])
    return q / np.linalg.norm(q, axis=-1, keepdims=True)


def fk_batch(model: M.KinematicModel, base_pos, base_quat, s) -> Tuple[np.ndarray, np.ndarray]:
    """Vectorised host FK used only to *place set-points* relative to the current
    pose; returns (link_R [N,n+1,3,3], link_p [N,n+1,3])."""
    N = base_pos.shape[0]
    n = model.n
    R = np.empty((N, n + 1, 3, 3))
    p = np.empty((N, n + 1, 3))
    R[:, 0] = quat_to_rot(base_quat)
    p[:, 0] = base_pos
    for k in range(1, n + 1):
        par = model.parent[k]
        a = model.axis[k]
        th = s[:, model.dof[k]]
        S = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        Rj = np.eye(3)[None] + np.sin(th)[:, None, None] * S[None] + (1 - np.cos(th))[:, None, None] * (S @ S)[None]
        R[:, k] = R[:, par] @ model.Xr[k][None] @ Rj
        p[:, k] = p[:, par] + R[:, par] @ model.Xp[k]
    return R, p


def frame_pose_batch(model, link_R, link_p, frame: int):
    L = model.frame_link[frame]
    return link_R[:, L] @ model.Fr[frame][None], link_p[:, L] + link_R[:, L] @ model.Fp[frame]


def com_batch(model, link_R, link_p):
    cw = link_p + np.einsum("nkij,kj->nki", link_R, model.com)
    return np.einsum("k,nki->ni", model.mass, cw) / model.mass.sum()


# ---------------------------------------------------------------- workloads
# gains/weights of utilities/balancing-position-control/config/robots/ergoCubGazeboV1/
#   blf_balancing_position_control/ik.ini:11-66
_ERGOCUB_REG_WEIGHT = np.array([1.0] * 3 + [2.0] * 8 + [1.0] * 12)


def _ergocub_core(m: M.KinematicModel) -> List[T.TaskSpec]:
    return [T.com_task(m, "COM_TASK", 2.0, priority=0),
            T.se3_task(m, "RIGHT_FOOT_TASK", "r_sole", 7.0, 5.0, priority=0),
            T.se3_task(m, "LEFT_FOOT_TASK", "l_sole", 7.0, 5.0, priority=0)]


def config(cid: int) -> Workload:
    """C1..C5 of BASELINE.json (C1/C2 on the iCub topology, C3..C5 on ergoCub)."""
    if cid in (1, 2):
        m = M.icub32()
        # tasks and gains of bindings/python/IK/tests/test_QP_inverse_kinematics.py:329-383,
        # extended with the second foot (BASELINE.json configs[0]: "SE3 feet + CoM + joint regularization")
        tasks = [T.com_task(m, "Com_task", 10.0, priority=0),
                 T.se3_task(m, "r_sole_task", "r_sole", 10.0, 10.0, priority=0),
                 T.se3_task(m, "l_sole_task", "l_sole", 10.0, 10.0, priority=0),
                 T.joint_tracking_task(m, "joint_tracking_task", [5.0] * m.n, priority=1, weight=[1.0] * m.n)]
        return Workload(f"C{cid}", m, tasks, 0.01,
                        "iCub-32 topology: 2xSE3 feet + CoM (priority 0, m=15) + JointTracking (priority 1)")
    m = M.ergocub23()
    if cid in (3, 5):
        dt = 0.01 if cid == 3 else 0.001
        tasks = _ergocub_core(m) + [
            T.joint_limits_task(m, "JOINT_LIMITS_TASK", [0.5] * m.n, dt, use_model_limits=True),
            T.joint_velocity_limits_task(m, "JOINT_VELOCITY_LIMITS_TASK", [2.0] * m.n, [-2.0] * m.n),
            T.so3_task(m, "TORSO_TASK", "chest", 5.0, priority=1, weight=[5.0, 5.0, 5.0]),
            T.joint_tracking_task(m, "JOINT_REGULARIZATION_TASK", [5.0] * m.n, priority=1, weight=_ERGOCUB_REG_WEIGHT)]
        return Workload(f"C{cid}", m, tasks, dt,
                        "ergoCub-23: CoM + 2xSE3 feet + JointLimits + JointVelocityLimits (priority 0, m=107), "
                        "SO3 chest + JointTracking (priority 1)")
    if cid == 4:
        tasks = _ergocub_core(m) + [
            T.se3_task(m, "LEFT_HAND_TASK", "l_hand_palm", 5.0, 5.0, priority=1, weight=[10.0] * 6),
            T.se3_task(m, "RIGHT_HAND_TASK", "r_hand_palm", 5.0, 5.0, priority=1, weight=[10.0] * 6),
            T.so3_task(m, "TORSO_TASK", "chest", 5.0, priority=1, weight=[5.0, 5.0, 5.0]),
            T.joint_tracking_task(m, "JOINT_REGULARIZATION_TASK", [5.0] * m.n, priority=1, weight=_ERGOCUB_REG_WEIGHT)]
        return Workload("C4", m, tasks, 0.01,
                        "ergoCub-23: feet + CoM priority 0 (m=15); hands 2xSE3 + torso SO3 weighted + JointTracking "
                        "(priority 1); no AngularMomentumTask")
    raise ValueError("config id must be 1..5")


def synth_inputs(w: Workload, N: int, cid: int = 0, seed: int = SEED, jr: float = 0.25, pr: float = 0.03,
                 ar: float = 0.06, cr: float = 0.02) -> Inputs:
    """Random joint configurations and SE(3)/CoM targets, SURVEY.md 8d."""
    m = w.model
    rng = np.random.Generator(np.random.Philox(key=seed + cid))
    d = m.q_upper - m.q_lower
    # joints: around the nominal pose, well inside the limits (most limit rows stay inactive,
    # a tail of instances sits close enough to a limit to activate it)
    s = m.q_nominal[None] + rng.uniform(-jr, jr, size=(N, m.n))
    near = rng.uniform(size=(N, m.n)) < 0.03
    s = np.where(near, np.where(rng.uniform(size=(N, m.n)) < 0.5, m.q_upper - 0.004 * d, m.q_lower + 0.004 * d)[...], s)
    s = np.clip(s, m.q_lower + 0.001 * d, m.q_upper - 0.001 * d)
    p = np.stack([rng.uniform(-0.5, 0.5, N), rng.uniform(-0.5, 0.5, N), rng.uniform(0.6, 0.9, N)], axis=1)
    q = so3_exp(rng.uniform(-0.3, 0.3, size=(N, 3)))
    link_R, link_p = fk_batch(m, p, q, s)
    sps: List[Optional[np.ndarray]] = []
    for t in w.tasks:
        if t.type in (T.TASK_SE3, T.TASK_SO3, T.TASK_R3):
            Rf, pf = frame_pose_batch(m, link_R, link_p, t.frame)
            dq = so3_exp(rng.uniform(-ar, ar, size=(N, 3)))
            dp = rng.uniform(-pr, pr, size=(N, 3))
            qd = quat_mul(rot_to_quat(Rf), dq)           # current pose o delta (right composition)
            pd = pf + np.einsum("nij,nj->ni", Rf, dp)
            if t.type == T.TASK_SE3:
                sps.append(np.concatenate([pd, qd, np.zeros((N, 6))], axis=1))
            elif t.type == T.TASK_SO3:
                sps.append(np.concatenate([qd, np.zeros((N, 3))], axis=1))
            else:
                sps.append(np.concatenate([pd, np.zeros((N, 3))], axis=1))
        elif t.type == T.TASK_COM:
            c = com_batch(m, link_R, link_p)
            sps.append(np.concatenate([c + rng.uniform(-cr, cr, size=(N, 3)), np.zeros((N, 3))], axis=1))
        elif t.type == T.TASK_JOINT_TRACKING:
            sps.append(np.concatenate([np.broadcast_to(m.q_nominal, (N, m.n)), np.zeros((N, m.n))], axis=1))
        else:
            sps.append(None)
    return Inputs(np.ascontiguousarray(p), np.ascontiguousarray(q), np.ascontiguousarray(s),
                  [None if sp is None else np.ascontiguousarray(sp) for sp in sps])


def synth_inputs_survey(w: Workload, N: int, cid: int = 0, seed: int = SEED) -> Inputs:
    """The distribution of SURVEY.md 8d as written: s ~ U(s- + 0.1 d, s+ - 0.1 d) per joint, frame targets = current pose o
    (dp ~ U(-5 cm, 5 cm)^3, dR = exp(U(-0.1, 0.1)^3)), CoM target = current CoM + U(-3 cm, 3 cm)^3.  Harder than synth_inputs: a
    larger share of the instances sits at a velocity limit or is infeasible under them."""
    m = w.model
    rng = np.random.Generator(np.random.Philox(key=seed + 100 + cid))
    d = m.q_upper - m.q_lower
    s = m.q_lower + (0.1 + 0.8 * rng.uniform(size=(N, m.n))) * d
    base = synth_inputs(w, N, cid, seed=seed, jr=0.0, pr=0.05, ar=0.1, cr=0.03)   # draws p, q and the deltas; the targets are rebuilt below
    p, q = base.base_pos, base.base_quat
    link_R, link_p = fk_batch(m, p, q, s)
    sps: List[Optional[np.ndarray]] = []
    for t in w.tasks:
        if t.type in (T.TASK_SE3, T.TASK_SO3, T.TASK_R3):
            Rf, pf = frame_pose_batch(m, link_R, link_p, t.frame)
            dq = so3_exp(rng.uniform(-0.1, 0.1, size=(N, 3)))
            dp = rng.uniform(-0.05, 0.05, size=(N, 3))
            qd = quat_mul(rot_to_quat(Rf), dq)
            pd = pf + np.einsum("nij,nj->ni", Rf, dp)
            if t.type == T.TASK_SE3:
                sps.append(np.concatenate([pd, qd, np.zeros((N, 6))], axis=1))
            elif t.type == T.TASK_SO3:
                sps.append(np.concatenate([qd, np.zeros((N, 3))], axis=1))
            else:
                sps.append(np.concatenate([pd, np.zeros((N, 3))], axis=1))
        elif t.type == T.TASK_COM:
            c = com_batch(m, link_R, link_p)
            sps.append(np.concatenate([c + rng.uniform(-0.03, 0.03, size=(N, 3)), np.zeros((N, 3))], axis=1))
        elif t.type == T.TASK_JOINT_TRACKING:
            sps.append(np.concatenate([np.broadcast_to(m.q_nominal, (N, m.n)), np.zeros((N, m.n))], axis=1))
        else:
            sps.append(None)
    return Inputs(np.ascontiguousarray(p), np.ascontiguousarray(q), np.ascontiguousarray(s),
                  [None if sp is None else np.ascontiguousarray(sp) for sp in sps])


def shard_range(N: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block of instance indices owned by ``rank``."""
    return (N * rank) // world, (N * (rank + 1)) // world
