"""The closed-loop scenarios of the reference's solver tests, rebuilt on random revolute trees
(src/IK/tests/QPInverseKinematicsTest.cpp): same sizes (n in {20, 35}, 10 extra frames), gains (5.0 on SE3/CoM,
1.0 on the regularisation, klim 0.5), limits (initial joints +- 0.5; zero width on joints 0 and 3 in the
"strict limits" case), dT = 10 ms, 30 ticks, Gaussian joint perturbation sigma 0.1 with seed 42."""
import numpy as np

import blf_b200  # noqa: F401
from blf_b200 import models as M, tasks as T, workloads as W

DT = 0.01       # QPInverseKinematicsTest.cpp:47
TICKS = 30      # :453
GAIN = 5.0      # :87


class Scenario:
    pass


def qp_ik_scenario(n: int, strict_limits: bool = False, unconstrained: bool = False, seed: int = 2) -> Scenario:
    m = M.random_tree(n, seed=seed, extra_frames=10)
    rng = np.random.default_rng(seed)
    # getDesiredReference (:221-243): random base transform, joints uniform in [0, 1)
    p_des = rng.uniform(-1, 1, 3)
    q_des = W.so3_exp(rng.uniform(-1.5, 1.5, 3)[None])[0]
    s_des = rng.uniform(0, 1, n)
    R, P = W.fk_batch(m, p_des[None], q_des[None], s_des[None])
    ee = m.frame_index(f"link{n}")                      # "model.getFrameName(numberOfJoints)" (:236)
    Rf, pf = W.frame_pose_batch(m, R, P, ee)
    com = W.com_batch(m, R, P)[0]
    # getSystem (:281-299): perturbed joints and base rotation angle
    g42 = np.random.default_rng(42)
    s0 = s_des + g42.normal(0, 0.1, n)
    phi = 2 * np.arctan2(np.linalg.norm(q_des[:3]), q_des[3])
    axis = q_des[:3] / max(np.linalg.norm(q_des[:3]), 1e-12)
    q0 = W.so3_exp((axis * (phi + g42.normal(0, 0.1)))[None])[0]
    half = np.full(n, 0.5)
    if strict_limits:
        half[0] = 0.0
        half[3] = 0.0                                   # :551-557
    prio = 1 if unconstrained else 0
    tasks = [T.se3_task(m, "se3_task", f"link{n}", GAIN, GAIN, priority=prio, weight=[10.0] * 6 if unconstrained else None),
             T.com_task(m, "com_task", GAIN, priority=prio, weight=[10.0] * 3 if unconstrained else None),
             T.joint_tracking_task(m, "regularization_task", np.ones(n), priority=1, weight=np.ones(n))]
    if not unconstrained:
        tasks.append(T.joint_limits_task(m, "joint_limits_task", np.full(n, 0.5), DT, use_model_limits=False,
                                         upper_limits=s0 + half, lower_limits=s0 - half))
    sc = Scenario()
    sc.model, sc.tasks, sc.ee = m, tasks, ee
    sc.p0, sc.q0, sc.s0 = p_des.copy(), q0, s0
    sc.com_des, sc.R_des, sc.p_ee_des = com, Rf[0], pf[0]
    sc.setpoints = [np.concatenate([pf[0], W.rot_to_quat(Rf)[0], np.zeros(6)]), np.concatenate([com, np.zeros(3)]),
                    np.concatenate([s_des, np.zeros(n)])] + ([] if unconstrained else [None])
    sc.strict = [0, 3] if strict_limits else []
    return sc


def tile_setpoints(sc: Scenario, N: int):
    return [None if sp is None else np.ascontiguousarray(np.tile(sp, (N, 1))) for sp in sc.setpoints]


def distance_gravity_scenario(n: int = 30, hard: bool = True, seed: int = 5) -> Scenario:
    """'QP-IK [Distance and Gravity tasks]' (QPInverseKinematicsTest.cpp:800-930: both at priority 0 beside the
    regularisation) and '... Unconstrained' (:932-1063: both at priority 1, weight 10).  Gains 5*5 and 5*10
    (:141, :152); the desired distance / body-frame gravity are those of the desired configuration (:265-271);
    the two frames are picked at least 5 indices away from the end effector and from each other (:253-263)."""
    m = M.random_tree(n, seed=seed, extra_frames=10)
    rng = np.random.default_rng(seed)
    p_des = rng.uniform(-1, 1, 3)
    q_des = W.so3_exp(rng.uniform(-1.5, 1.5, 3)[None])[0]
    s_des = rng.uniform(0, 1, n)
    R, P = W.fk_batch(m, p_des[None], q_des[None], s_des[None])
    while True:
        fg, fd = int(rng.integers(3, n - 5)), int(rng.integers(3, n - 5))
        if abs(fg - fd) >= 5:
            break
    Rd, pd = W.frame_pose_batch(m, R, P, m.frame_index(f"link{fd}"))
    Rg, pg = W.frame_pose_batch(m, R, P, m.frame_index(f"link{fg}"))
    g42 = np.random.default_rng(42)
    s0 = s_des + g42.normal(0, 0.1, n)
    phi = 2 * np.arctan2(np.linalg.norm(q_des[:3]), q_des[3])
    axis = q_des[:3] / max(np.linalg.norm(q_des[:3]), 1e-12)
    q0 = W.so3_exp((axis * (phi + g42.normal(0, 0.1)))[None])[0]
    prio = 0 if hard else 1
    tasks = [T.joint_tracking_task(m, "regularization_task", np.ones(n), priority=1, weight=np.ones(n)),
             T.distance_task(m, "distance_task", f"link{fd}", GAIN * 5, priority=prio, weight=None if hard else [10.0]),
             T.gravity_task(m, "gravity_task", f"link{fg}", GAIN * 10, priority=prio, weight=None if hard else [10.0] * 2)]
    sc = Scenario()
    sc.model, sc.tasks = m, tasks
    sc.frame_distance, sc.frame_gravity = m.frame_index(f"link{fd}"), m.frame_index(f"link{fg}")
    sc.p0, sc.q0, sc.s0 = p_des.copy(), q0, s0
    sc.distance_des, sc.gravity_des = float(np.linalg.norm(pd[0])), Rg[0][2].copy()
    sc.setpoints = [np.concatenate([s_des, np.zeros(n)]), np.array([sc.distance_des]), np.concatenate([sc.gravity_des, np.zeros(3)])]
    sc.strict = []
    return sc
