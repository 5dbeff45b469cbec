"""Oracle pins for K2 (tasks, assembly) and K3 (QP): the identities the reference's own unit tests check
(src/IK/tests/*TaskTest.cpp) re-expressed against independent numpy computations, KKT residuals, a scipy
cross-check and the reference's closed-loop convergence scenarios."""
import numpy as np
import pytest

import blf_b200  # noqa: F401
from blf_b200 import models as M, tasks as T, workloads as W
from oracle.oracle import Oracle
from helpers import random_state, rot_log, solve_qp_active_set
import scenarios as S


@pytest.fixture(scope="module")
def ergo():
    return M.ergocub23()


def _so3_ctrl(Rcur, qd, wff, kp):
    Rd = W.quat_to_rot(np.asarray(qd)[None])[0]
    return wff + kp * rot_log(Rd @ Rcur.T)        # left-trivialised error Log(R_d R^T)


def test_se3_task_A_is_jacobian_b_is_control_law(ergo):
    """SE3TaskTest.cpp:79-139: A == frame Jacobian, b == [R3 control; SO3 control]."""
    m = ergo
    rng = np.random.default_rng(0)
    t = T.se3_task(m, "t", "l_sole", [1.0, 2.0, 3.0], [4.0, 5.0, 6.0], priority=0)
    o = Oracle(m, [t])
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    sp = np.concatenate([rng.normal(size=3), W.so3_exp(rng.normal(size=(1, 3)))[0], rng.normal(size=6)])
    A, b = o.task_update(0, R, P, s, sp)
    J = o.frame_jacobian(R, P, t.frame)
    Rf, pf = o.frame_pose(R, P, t.frame)
    assert np.array_equal(A, J)
    assert np.allclose(b[:3], sp[7:10] + t.kp_linear * (sp[:3] - pf), atol=1e-13)
    assert np.allclose(b[3:], _so3_ctrl(Rf, sp[3:7], sp[10:13], t.kp_angular), atol=1e-12)


def test_se3_task_mask_and_disabled_controller(ergo):
    """SE3TaskTest.cpp:142-231 (mask: selected linear rows, then the 3 angular rows) and SE3Task.cpp:251-258."""
    m = ergo
    rng = np.random.default_rng(1)
    t = T.se3_task(m, "t", "r_sole", 2.0, 3.0, mask=(True, False, True))
    o = Oracle(m, [t])
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    sp = np.concatenate([rng.normal(size=3), W.so3_exp(rng.normal(size=(1, 3)))[0], rng.normal(size=6)])
    A, b = o.task_update(0, R, P, s, sp)
    J = o.frame_jacobian(R, P, t.frame)
    assert A.shape == (5, m.nv)
    assert np.array_equal(A, J[[0, 2, 3, 4, 5]])
    t.controller_enabled = False
    o.set_tasks([t])
    _, b2 = o.task_update(0, R, P, s, sp)
    assert np.allclose(b2, np.concatenate([sp[[7, 9]], sp[10:13]]))   # feed-forward only


def test_com_so3_r3_tasks(ergo):
    """CoMTaskTest.cpp:82-130, SO3TaskTest.cpp:83-131 (A = J.bottomRows<3>()), R3TaskTest."""
    m = ergo
    rng = np.random.default_rng(2)
    tasks = [T.com_task(m, "c", [1.0, 2.0, 3.0]), T.so3_task(m, "o", "chest", 5.0, priority=1, weight=[1, 1, 1]),
             T.r3_task(m, "r", "l_hand_palm", 4.0, mask=(False, True, True))]
    o = Oracle(m, tasks)
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    c, Jc = o.com(R, P)
    spc = rng.normal(size=6)
    A, b = o.task_update(0, R, P, s, spc)
    assert np.array_equal(A, Jc) and np.allclose(b, spc[3:] + tasks[0].kp_linear * (spc[:3] - c))
    spo = np.concatenate([W.so3_exp(rng.normal(size=(1, 3)))[0], rng.normal(size=3)])
    A, b = o.task_update(1, R, P, s, spo)
    Rf, pf = o.frame_pose(R, P, tasks[1].frame)
    assert np.array_equal(A, o.frame_jacobian(R, P, tasks[1].frame)[3:])
    assert np.allclose(b, _so3_ctrl(Rf, spo[:4], spo[4:], tasks[1].kp_angular), atol=1e-12)
    spr = rng.normal(size=6)
    A, b = o.task_update(2, R, P, s, spr)
    Rf, pf = o.frame_pose(R, P, tasks[2].frame)
    assert np.array_equal(A, o.frame_jacobian(R, P, tasks[2].frame)[[1, 2]])
    assert np.allclose(b, (spr[3:] + 4.0 * (spr[:3] - pf))[[1, 2]])


def test_joint_tasks(ergo):
    """JointTrackingTaskTest.cpp:84-114 (A = [0 I], b = -Kp s for a zero set-point), JointLimitsTaskTest.cpp:98-138
    (A = [I; -I] on the joint columns, b = klim (limit - s)/dt on both halves), JointVelocityLimitsTaskTest.cpp:88-130."""
    m = ergo
    n = m.n
    rng = np.random.default_rng(3)
    kp, klim = rng.uniform(1, 5, n), rng.uniform(0, 1, n)
    up, lo = m.q_upper, m.q_lower
    tasks = [T.joint_tracking_task(m, "jt", kp, priority=1, weight=np.ones(n)),
             T.joint_limits_task(m, "jl", klim, 0.01), T.joint_velocity_limits_task(m, "jv", np.full(n, 2.0), np.full(n, -3.0))]
    o = Oracle(m, tasks)
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    A, b = o.task_update(0, R, P, s, np.zeros(2 * n))
    assert np.array_equal(A, np.hstack([np.zeros((n, 6)), np.eye(n)])) and np.allclose(b, -kp * s)
    A, b = o.task_update(1, R, P, s, None)
    assert np.array_equal(A, np.vstack([np.hstack([np.zeros((n, 6)), np.eye(n)]), np.hstack([np.zeros((n, 6)), -np.eye(n)])]))
    assert np.allclose(b[:n], klim * (up - s) / 0.01) and np.allclose(b[n:], klim * (s - lo) / 0.01)
    A, b = o.task_update(2, R, P, s, None)
    assert np.allclose(b[:n], 2.0) and np.allclose(b[n:], 3.0)


@pytest.mark.parametrize("cid", [1, 3, 4])
def test_assembly_follows_advance(cid):
    """QPInverseKinematics.cpp:442-491: H = sum A'WA, g = -sum A'Wb, C/l/u stacked in task order, -1e30 lower bounds."""
    w = W.config(cid)
    inp = W.synth_inputs(w, 4, cid)
    o = Oracle(w.model, w.tasks)
    n, nv = w.model.n, w.model.nv
    for i in range(4):
        sps = [None if sp is None else sp[i] for sp in inp.setpoints]
        H, g, C, l, u = o.assemble(inp.base_pos[i], inp.base_quat[i], inp.s[i], sps)
        R, P = o.fk(inp.base_pos[i], inp.base_quat[i], inp.s[i])
        H2, g2, rows = np.zeros((nv, nv)), np.zeros(nv), []
        for ti, t in enumerate(w.tasks):
            A, b = o.task_update(ti, R, P, inp.s[i], sps[ti])
            if t.priority == 1:
                H2 += A.T @ np.diag(t.weight) @ A
                g2 -= A.T @ (t.weight * b)
            else:
                rows.append((A, b, t.is_inequality))
        assert np.allclose(H, H2, atol=1e-12) and np.allclose(g, g2, atol=1e-12)
        assert np.array_equal(C, np.vstack([r[0] for r in rows]))
        assert np.array_equal(u, np.concatenate([r[1] for r in rows]))
        assert np.array_equal(l, np.concatenate([np.full(len(r[1]), -1e30) if r[2] else r[1] for r in rows]))
        assert C.shape[0] == {1: 15, 3: 107, 4: 15}[cid]


@pytest.mark.parametrize("cid", [1, 3, 4])
def test_qp_kkt_and_scipy_crosscheck(cid):
    """The oracle's optimum satisfies the KKT conditions to <= 1e-9 and agrees with scipy's SLSQP."""
    w = W.config(cid)
    inp = W.synth_inputs(w, 6, cid)
    o = Oracle(w.model, w.tasks)
    for i in range(6):
        sps = [None if sp is None else sp[i] for sp in inp.setpoints]
        H, g, C, l, u = o.assemble(inp.base_pos[i], inp.base_quat[i], inp.s[i], sps)
        x, y, z, info = o.qp_solve(H, g, C, l, u)
        assert info.status == 0
        assert o.qp_kkt(H, g, C, l, u, x, y) <= 1e-9
        if i < 3:
            xs = solve_qp_active_set(H, g, C, l, u)
            assert np.abs(xs - x).max() < 2e-6


def test_qp_warm_start_and_default_settings():
    """OSQP defaults (what the reference actually runs, QPInverseKinematics.cpp:339) land within 1e-2 of the optimum;
    warm-starting from the solution converges at the first termination check."""
    w = W.config(3)
    inp = W.synth_inputs(w, 4, 3)
    o = Oracle(w.model, w.tasks)
    for i in range(4):
        sps = [None if sp is None else sp[i] for sp in inp.setpoints]
        H, g, C, l, u = o.assemble(inp.base_pos[i], inp.base_quat[i], inp.s[i], sps)
        x, y, z, info = o.qp_solve(H, g, C, l, u)
        xd, yd, zd, infod = o.qp_solve(H, g, C, l, u, o.settings(parity=False))
        assert infod.status == 0 and np.abs(xd - x).max() < 2e-2
        x2, _, _, info2 = o.qp_solve(H, g, C, l, u, o.settings(), x0=x, y0=y, z0=z)
        assert info2.iters <= 25 and np.abs(x2 - x).max() < 1e-8


@pytest.mark.parametrize("n", [20, 35])
@pytest.mark.parametrize("strict", [False, True])
def test_closed_loop_convergence_scenarios(n, strict):
    """'QP-IK' (QPInverseKinematicsTest.cpp:357-500, tolerance 2e-1) and 'With strict limits' (:502-640:
    |sdot_0|, |sdot_3| < 1e-2 on every tick, positions within 1e-2 of the initial ones)."""
    sc = S.qp_ik_scenario(n, strict_limits=strict)
    o = Oracle(sc.model, sc.tasks)
    p, q, s = sc.p0[None].copy(), sc.q0[None].copy(), sc.s0[None].copy()
    warm = o.new_warm(1)
    for _ in range(S.TICKS):
        nu, status, _ = o.tick_batch(p, q, s, S.tile_setpoints(sc, 1), warm=warm, dt_integrate=S.DT)
        assert status[0] == 0
        for j in sc.strict:
            assert abs(nu[0, 6 + j]) < 1e-2
    for j in sc.strict:
        assert abs(s[0, j] - sc.s0[j]) < 1e-2
    R, P = o.fk(p[0], q[0], s[0])
    com, _ = o.com(R, P)
    Rf, pf = o.frame_pose(R, P, sc.ee)
    if not strict:   # the strict case of the reference only checks the limits
        assert np.linalg.norm(com - sc.com_des) <= 2e-1 * min(np.linalg.norm(com), np.linalg.norm(sc.com_des))
        assert np.abs(sc.R_des.T @ (pf - sc.p_ee_des)).max() < 2e-1 and np.abs(rot_log(sc.R_des.T @ Rf)).max() < 2e-1


def test_unconstrained_scenario():
    """'...Unconstrained' (QPInverseKinematicsTest.cpp:932-1063): every task at priority 1."""
    sc = S.qp_ik_scenario(20, unconstrained=True)
    o = Oracle(sc.model, sc.tasks)
    assert o.m == 0
    p, q, s = sc.p0[None].copy(), sc.q0[None].copy(), sc.s0[None].copy()
    for _ in range(S.TICKS):
        nu, status, _ = o.tick_batch(p, q, s, S.tile_setpoints(sc, 1), dt_integrate=S.DT)
        assert status[0] == 0
    R, P = o.fk(p[0], q[0], s[0])
    com, _ = o.com(R, P)
    assert np.linalg.norm(com - sc.com_des) < 2e-1


def test_icub_python_test_vector():
    """The one concrete humanoid input of the reference (bindings/python/IK/tests/test_QP_inverse_kinematics.py:293-425):
    identity base, its 32 joint values, CoM/SE3(r_sole)/JointTracking tasks with its gains; its assertions are
    advance() == True and 'second solution differs from the first'."""
    m = M.icub32()
    s = np.array([0.4147046680648114, 0.020109925966605414, -0.008920202688283316, -0.9417931898020863, -0.5243497613177115,
                  -0.021969353979888327, 0.41490863370740766, 0.019729971064546152, -0.00875663754461344, -0.942238731363784,
                  -0.5245880299799013, -0.021615580177534933, 0.10368807143683698, -1.7513061346176159e-06, -3.762644846952276e-05,
                  -0.1599473872698949, 0.4341604471687158, 0.18350557928473807, 0.5390720970077302, 6.0766272665334824e-05,
                  -0.0010549820549572227, -6.457409077219657e-05, 0.0002255313408076292, 2.3204341497780863e-05, 5.920527253092593e-07,
                  -0.1591984430641039, 0.4340643739146118, 0.18311274358002663, 0.5398797698508592, 5.367856260643534e-05,
                  -0.0010455524092736824, 6.414366731688693e-05])
    tasks = [T.com_task(m, "Com_task", 10.0), T.se3_task(m, "se3_task", "r_sole", 10.0, 10.0),
             T.joint_tracking_task(m, "joint_tracking_task", [5.0] * 32, priority=1, weight=[1.0] * 32)]
    o = Oracle(m, tasks)
    p, q = np.zeros((1, 3)), np.array([[0.0, 0, 0, 1]])
    rng = np.random.default_rng(0)
    outs = []
    for com_z, y in ((0.53, 0.1), (0.5, 0.2)):
        sps = [np.array([[0, 0, com_z, 0, 0, 0]]), np.array([[-0.02, y, 0.0, 0, 0, 0, 1] + [0.0] * 6]),
               np.concatenate([s + rng.uniform(-5e-3, 5e-3, 32), np.zeros(32)])[None]]
        nu, status, _ = o.tick_batch(p.copy(), q.copy(), s[None].copy(), sps)
        assert status[0] == 0
        outs.append(nu[0])
    assert np.abs(outs[0] - outs[1]).max() > 1e-3


def test_distance_and_gravity_task_rows(ergo):
    """DistanceTaskTest.cpp / GravityTaskTest.cpp check A and b against the formulas of DistanceTask.cpp:196-243 and
    GravityTask.cpp:160-199; here additionally A is checked as the derivative of the controlled quantity."""
    m = ergo
    rng = np.random.default_rng(9)
    tasks = [T.distance_task(m, "d0", "l_hand_palm", 3.0, weight=[1.0]),
             T.distance_task(m, "d1", "l_hand_palm", 2.0, weight=[1.0], reference_frame_name="r_sole"),
             T.gravity_task(m, "g", "chest", 4.0, weight=[1.0, 1.0])]
    o = Oracle(m, tasks)
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    Rt, pt = o.frame_pose(R, P, tasks[0].frame)
    Rr, pr = o.frame_pose(R, P, tasks[1].ref_frame)
    Jt, Jr = o.frame_jacobian(R, P, tasks[0].frame), o.frame_jacobian(R, P, tasks[1].ref_frame)
    A, b = o.task_update(0, R, P, s, np.array([0.7]))
    assert A.shape == (1, 6 + m.n)
    assert np.allclose(A[0], pt @ Jt[:3] / np.linalg.norm(pt), atol=1e-13) and np.isclose(b[0], 3.0 * (0.7 - np.linalg.norm(pt)))
    A, b = o.task_update(1, R, P, s, np.array([0.2]))
    d = pt - pr
    assert np.allclose(A[0], d @ (Jt[:3] - Jr[:3]) / np.linalg.norm(d), atol=1e-13) and np.isclose(b[0], 2.0 * (0.2 - np.linalg.norm(d)))
    # d|d|/dt = A nu  (finite differences along a random velocity)
    from helpers import perturb_state
    nu = rng.normal(size=6 + m.n)
    eps = 1e-6
    p2, q2, s2 = perturb_state(p, q, s, nu, eps)
    R2, P2 = o.fk(p2, q2, s2)
    d2 = o.frame_pose(R2, P2, tasks[0].frame)[1] - o.frame_pose(R2, P2, tasks[1].ref_frame)[1]
    assert abs((np.linalg.norm(d2) - np.linalg.norm(d)) / eps - A[0] @ nu) < 1e-5
    # gravity: A = first two angular rows, b = (R ff)[:2] - kp [0 -1 0; 1 0 0] R z_des, z_des normalised
    Rc, _ = o.frame_pose(R, P, tasks[2].frame)
    z = np.array([0.3, -0.2, 2.0])
    ff = rng.normal(size=3)
    A, b = o.task_update(2, R, P, s, np.concatenate([z, ff]))
    assert np.array_equal(A, o.frame_jacobian(R, P, tasks[2].frame)[3:5])
    zn = Rc @ (z / np.linalg.norm(z))
    assert np.allclose(b, (Rc @ ff)[:2] - 4.0 * np.array([-zn[1], zn[0]]), atol=1e-13)
    # at the desired attitude (R z_des = e_z) the feedback vanishes
    A, b = o.task_update(2, R, P, s, np.concatenate([Rc[2], np.zeros(3)]))
    assert np.abs(b).max() < 1e-13
    # below the numeric threshold the Jacobian is divided by the threshold (DistanceTask.cpp:232)
    tasks[0].threshold = 10.0
    o2 = Oracle(m, tasks)
    A2, _ = o2.task_update(0, R, P, s, np.array([0.7]))
    assert np.allclose(A2[0], pt @ Jt[:3] / 10.0, atol=1e-13)


@pytest.mark.parametrize("hard", [True, False])
def test_distance_gravity_closed_loop(hard):
    """'QP-IK [Distance and Gravity tasks]' and '... Unconstrained' (QPInverseKinematicsTest.cpp:800-1063, tolerance 2e-2)."""
    sc = S.distance_gravity_scenario(30, hard=hard)
    o = Oracle(sc.model, sc.tasks)
    p, q, s = sc.p0[None].copy(), sc.q0[None].copy(), sc.s0[None].copy()
    for _ in range(S.TICKS):
        nu, status, _ = o.tick_batch(p, q, s, S.tile_setpoints(sc, 1), dt_integrate=S.DT)
        assert status[0] == 0
    R, P = o.fk(p[0], q[0], s[0])
    _, pd = o.frame_pose(R, P, sc.frame_distance)
    Rg, _ = o.frame_pose(R, P, sc.frame_gravity)
    assert abs(np.linalg.norm(pd) - sc.distance_des) <= 2e-2 * sc.distance_des
    assert np.abs(Rg[2] - sc.gravity_des).max() < 2e-2


def _angular_momentum_numeric(o, m, p, q, s, nu, eps=1e-6):
    """L about the CoM from first principles: link angular velocities and link-CoM velocities by central differences of the
    forward kinematics along nu, L = sum_k R_k I_k R_k' w_k + m_k (c_k - c) x cdot_k."""
    from helpers import perturb_state

    def link_state(pp, qq, ss):
        R, P = o.fk(pp, qq, ss)
        R = np.asarray(R).reshape(-1, 3, 3); P = np.asarray(P).reshape(-1, 3)
        return R, P + np.einsum('kij,kj->ki', R, m.com)
    Rp, cp = link_state(*perturb_state(p, q, s, nu, eps))
    Rm, cm = link_state(*perturb_state(p, q, s, nu, -eps))
    R0, c0 = link_state(p, q, s)
    com = (m.mass[:, None] * c0).sum(0) / m.mass.sum()
    L = np.zeros(3)
    for k in range(m.n + 1):
        w = rot_log(Rp[k] @ Rm[k].T) / (2 * eps)          # world-frame angular velocity of link k
        cdot = (cp[k] - cm[k]) / (2 * eps)
        L += R0[k] @ m.inertia[k] @ R0[k].T @ w + m.mass[k] * np.cross(c0[k] - com, cdot)
    return L


def test_angular_momentum_task(ergo):
    """AngularMomentumTask.cpp:137-210: A = rows 3..5 of the centroidal momentum matrix (masked), b = the desired momentum.
    The matrix is checked against the momentum computed from first principles, its base-linear block is zero (momentum about
    the CoM does not see a pure translation), and it is invariant under lumping (a model reduced to fewer joints and evaluated
    with the removed joints at zero carries the same momentum for the same motion)."""
    m = ergo
    rng = np.random.default_rng(21)
    tasks = [T.angular_momentum_task(m, "L", priority=0), T.angular_momentum_task(m, "Lxz", priority=1, weight=[2.0, 3.0], mask=(True, False, True))]
    o = Oracle(m, tasks)
    assert np.linalg.eigvalsh(m.inertia).min() > 0
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    A = o.angular_momentum_matrix(R, P)
    assert np.abs(A[:, :3]).max() < 1e-13
    for _ in range(3):
        nu = rng.normal(size=6 + m.n)
        assert np.abs(A @ nu - _angular_momentum_numeric(o, m, p, q, s, nu)).max() < 1e-6
    sp = rng.normal(size=3)
    A0, b0 = o.task_update(0, R, P, s, sp)
    assert np.array_equal(A0, A) and np.array_equal(b0, sp)
    A1, b1 = o.task_update(1, R, P, s, sp)
    assert np.array_equal(A1, A[[0, 2]]) and np.array_equal(b1, sp[[0, 2]])
    # lumping invariance: the 32-joint and the 26-joint iCub are the same robot when the six wrist joints rest at zero
    m32, m26 = M.icub32(), M.icub26()
    o32, o26 = Oracle(m32, [T.angular_momentum_task(m32, "L", priority=0)]), Oracle(m26, [T.angular_momentum_task(m26, "L", priority=0)])
    s26 = rng.uniform(-0.4, 0.4, m26.n)
    s32, nu26 = np.zeros(m32.n), rng.normal(size=6 + m26.n)
    nu32 = np.zeros(6 + m32.n)
    nu32[:6] = nu26[:6]
    for j26, name in enumerate(m26.joint_names):
        j32 = m32.joint_names.index(name)
        s32[j32], nu32[6 + j32] = s26[j26], nu26[6 + j26]
    R26, P26 = o26.fk(p, q, s26)
    R32, P32 = o32.fk(p, q, s32)
    assert np.abs(o26.angular_momentum_matrix(R26, P26) @ nu26 - o32.angular_momentum_matrix(R32, P32) @ nu32).max() < 1e-12


def test_exogenous_feedback_changes_b_only(ergo):
    """SE3Task.cpp:261-268 / :338-366 and CoMTask.cpp:192-195 / :262-273: with use_*_exogenous_feedback the controllers compare
    the set-point with the pose the caller measured (record tail), the Jacobian stays the model's."""
    m = ergo
    rng = np.random.default_rng(31)
    plain = [T.se3_task(m, "f", "l_sole", [1.0, 2.0, 3.0], 4.0), T.com_task(m, "c", 2.0)]
    exo = [T.se3_task(m, "f", "l_sole", [1.0, 2.0, 3.0], 4.0, use_position_exogenous_feedback=True, use_orientation_exogenous_feedback=True),
           T.se3_task(m, "g", "l_sole", [1.0, 2.0, 3.0], 4.0, use_position_exogenous_feedback=True),
           T.com_task(m, "c", 2.0, use_exogenous_feedback=True)]
    assert exo[0].record_len(m.n) == 20 and exo[1].record_len(m.n) == 20 and exo[2].record_len(m.n) == 9 and plain[0].record_len(m.n) == 13
    o, oe = Oracle(m, plain), Oracle(m, exo)
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    sp = np.concatenate([rng.normal(size=3), W.so3_exp(rng.normal(size=(1, 3)))[0], rng.normal(size=6)])
    p_fb, q_fb = rng.normal(size=3), W.so3_exp(rng.normal(size=(1, 3)))[0]
    A0, b0 = o.task_update(0, R, P, s, sp)
    A1, b1 = oe.task_update(0, R, P, s, np.concatenate([sp, p_fb, q_fb]))
    assert np.array_equal(A0, A1)
    assert np.allclose(b1[:3], sp[7:10] + np.array([1.0, 2.0, 3.0]) * (sp[:3] - p_fb))
    assert np.allclose(b1[3:], _so3_ctrl(W.quat_to_rot(q_fb[None])[0], sp[3:7], sp[10:13], np.full(3, 4.0)), atol=1e-12)
    # position only: the angular part is the model's
    A2, b2 = oe.task_update(1, R, P, s, np.concatenate([sp, p_fb, q_fb]))
    assert np.allclose(b2[:3], b1[:3]) and np.allclose(b2[3:], b0[3:])
    # feedback equal to the model's pose reproduces the plain task
    Rf, pf = o.frame_pose(R, P, plain[0].frame)
    _, b3 = oe.task_update(0, R, P, s, np.concatenate([sp, pf, W.rot_to_quat(Rf[None])[0]]))
    assert np.allclose(b3, b0, atol=1e-12)
    spc, c_fb = rng.normal(size=6), rng.normal(size=3)
    Ac0, _ = o.task_update(1, R, P, s, spc)
    Ac1, bc1 = oe.task_update(2, R, P, s, np.concatenate([spc, c_fb]))
    assert np.array_equal(Ac0, Ac1) and np.allclose(bc1, spc[3:] + 2.0 * (spc[:3] - c_fb))
