"""Oracle pins for K1 (FK / Jacobians / CoM) and K4 (integration) by mathematically independent means
(the reference holds no golden vectors for these; "parity unpinned")."""
import numpy as np
import pytest

import blf_b200  # noqa: F401
from blf_b200 import models as M, workloads as W
from oracle.oracle import Oracle
from helpers import perturb_state, random_state, rodrigues, rot_log, skew

MODELS = [M.ergocub23, M.icub32, M.icub26, lambda: M.random_tree(20, seed=2), lambda: M.random_tree(35, seed=3)]


@pytest.mark.parametrize("mk", MODELS)
def test_fk_matches_independent_numpy_fk(mk):
    m = mk()
    o = Oracle(m)
    rng = np.random.default_rng(0)
    for _ in range(5):
        p, q, s = random_state(m, rng)
        R, P = o.fk(p, q, s)
        R2, P2 = W.fk_batch(m, p[None], q[None], s[None])
        assert np.abs(R - R2[0]).max() < 1e-13 and np.abs(P - P2[0]).max() < 1e-13
        for k in range(m.n + 1):
            assert np.abs(R[k] @ R[k].T - np.eye(3)).max() < 1e-13


@pytest.mark.parametrize("mk", MODELS)
def test_frame_jacobian_vs_central_differences(mk):
    """J * nu must equal d/dt of the frame pose along nu (MIXED: world linear velocity of the frame
    origin, world angular velocity); central differences, tolerance 1e-7."""
    m = mk()
    o = Oracle(m)
    rng = np.random.default_rng(1)
    frames = list(range(0, len(m.frame_names), max(1, len(m.frame_names) // 6)))
    eps = 1e-6
    for _ in range(3):
        p, q, s = random_state(m, rng)
        R, P = o.fk(p, q, s)
        for f in frames:
            J = o.frame_jacobian(R, P, f)
            Rf, pf = o.frame_pose(R, P, f)
            for _ in range(3):
                dnu = rng.normal(size=m.nv)
                Rp, Pp = o.fk(*perturb_state(p, q, s, dnu, eps))
                Rm, Pm = o.fk(*perturb_state(p, q, s, dnu, -eps))
                Rfp, pfp = o.frame_pose(Rp, Pp, f)
                Rfm, pfm = o.frame_pose(Rm, Pm, f)
                v_fd = (pfp - pfm) / (2 * eps)
                w_fd = rot_log(Rfp @ Rfm.T) / (2 * eps)
                pred = J @ dnu
                assert np.abs(pred[:3] - v_fd).max() < 1e-7 * (1 + np.abs(v_fd).max())
                assert np.abs(pred[3:] - w_fd).max() < 1e-7 * (1 + np.abs(w_fd).max())


@pytest.mark.parametrize("mk", MODELS)
def test_jacobian_structure(mk):
    """Base block [I, -S(p_F - p_B); 0, I]; columns of joints off the base->frame path are zero."""
    m = mk()
    o = Oracle(m)
    rng = np.random.default_rng(2)
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    for f in range(0, len(m.frame_names), 3):
        J = o.frame_jacobian(R, P, f)
        _, pf = o.frame_pose(R, P, f)
        assert np.allclose(J[:3, :3], np.eye(3)) and np.allclose(J[3:, 3:6], np.eye(3)) and np.allclose(J[3:, :3], 0)
        assert np.allclose(J[:3, 3:6], -skew(pf - p), atol=1e-14)
        on_path = set()
        k = m.frame_link[f]
        while k > 0:
            on_path.add(int(m.dof[k]))
            k = m.parent[k]
        for j in range(m.n):
            if j not in on_path:
                assert np.all(J[:, 6 + j] == 0)
            else:
                assert np.isclose(np.linalg.norm(J[3:, 6 + j]), 1.0)


@pytest.mark.parametrize("mk", MODELS)
def test_com_and_com_jacobian(mk):
    m = mk()
    o = Oracle(m)
    rng = np.random.default_rng(3)
    eps = 1e-6
    p, q, s = random_state(m, rng)
    R, P = o.fk(p, q, s)
    c, Jc = o.com(R, P)
    c2 = W.com_batch(m, *W.fk_batch(m, p[None], q[None], s[None]))[0]
    assert np.abs(c - c2).max() < 1e-13
    for _ in range(4):
        dnu = rng.normal(size=m.nv)
        cp, _ = o.com(*o.fk(*perturb_state(p, q, s, dnu, eps)))
        cm, _ = o.com(*o.fk(*perturb_state(p, q, s, dnu, -eps)))
        fd = (cp - cm) / (2 * eps)
        assert np.abs(Jc @ dnu - fd).max() < 1e-7 * (1 + np.abs(fd).max())


def test_reduced_model_lumping_keeps_kinematics():
    """Freezing joints at zero and lumping (loadReducedModelFromFile semantics) must not move frames or the CoM."""
    full, red = M.icub32(), M.icub26()
    of, orr = Oracle(full), Oracle(red)
    rng = np.random.default_rng(4)
    p, q, s26 = random_state(red, rng)
    s32 = np.zeros(full.n)
    for j, name in enumerate(red.joint_names):
        s32[full.joint_names.index(name)] = s26[j]
    Rf, Pf = of.fk(p, q, s32)
    Rr, Pr = orr.fk(p, q, s26)
    assert abs(full.mass.sum() - red.mass.sum()) < 1e-12
    assert np.abs(of.com(Rf, Pf)[0] - orr.com(Rr, Pr)[0]).max() < 1e-13
    for name in ("l_sole", "r_sole", "l_hand_palm", "head_imu", "chest"):
        Ra, pa = of.frame_pose(Rf, Pf, full.frame_index(name))
        Rb, pb = orr.frame_pose(Rr, Pr, red.frame_index(name))
        assert np.abs(Ra - Rb).max() < 1e-13 and np.abs(pa - pb).max() < 1e-13


def test_model_text_roundtrip():
    m = M.ergocub23()
    m2 = M.model_from_text(m.to_text())
    assert m2.joint_names == m.joint_names and m2.frame_names == m.frame_names
    for a in ("parent", "dof", "Xr", "Xp", "axis", "mass", "com", "Fr", "Fp", "q_lower", "q_upper"):
        assert np.array_equal(getattr(m, a), getattr(m2, a))


# ---- SO(3) and the integrator -------------------------------------------------------------
def test_so3_exp_log_roundtrip_and_conversion():
    o = Oracle(M.ergocub23())
    import ctypes as C
    rng = np.random.default_rng(5)
    L = o.L
    for _ in range(200):
        phi = rng.normal(size=3)
        phi *= rng.uniform(0, 3.1) / np.linalg.norm(phi)
        q = np.empty(4); back = np.empty(3); Rm = np.empty((3, 3)); q2 = np.empty(4)
        L.orc_so3_exp(phi.ctypes.data_as(C.POINTER(C.c_double)), q.ctypes.data_as(C.POINTER(C.c_double)))
        L.orc_so3_log(q.ctypes.data_as(C.POINTER(C.c_double)), back.ctypes.data_as(C.POINTER(C.c_double)))
        assert np.abs(back - phi).max() < 1e-12
        L.orc_quat_to_rot(q.ctypes.data_as(C.POINTER(C.c_double)), Rm.ctypes.data_as(C.POINTER(C.c_double)))
        assert np.abs(Rm - rodrigues(phi)).max() < 1e-13
        L.orc_rot_to_quat(Rm.ctypes.data_as(C.POINTER(C.c_double)), q2.ctypes.data_as(C.POINTER(C.c_double)))
        assert min(np.abs(q2 - q).max(), np.abs(q2 + q).max()) < 1e-12


def test_integrator_closed_form_and_left_plus():
    """Reference integrator test (src/ContinuousDynamicalSystem/tests/IntegratorFloatingBaseSystemVelocityKinematics.cpp:25-88):
    constant twist / joint velocity, dT = 100 us for 0.5 s, closed form p0 + t v, AngleAxis(|w| t) R0, s0 + t sdot at 1e-3;
    with a NON-identity R0 so that left-plus (exp(w dt) R, the Python twin's `SO3Tangent(w t) + R0`,
    bindings/python/ContinuousDynamicalSystem/tests/test_floating_base_system_kinematics.py:9-12) is distinguished
    from right-plus."""
    m = M.ergocub23()
    o = Oracle(m)
    rng = np.random.default_rng(6)
    nu = rng.normal(size=m.nv)
    p0, q0, s0 = random_state(m, rng)
    dT, T = 1e-4, 0.5
    p, q, s = p0.copy(), q0.copy(), s0.copy()
    for _ in range(int(round(T / dT))):
        p, q, s = o.integrate(p, q, s, nu, dT)
    R_expected = rodrigues(nu[3:6] * T) @ W.quat_to_rot(q0[None])[0]
    R_wrong = W.quat_to_rot(q0[None])[0] @ rodrigues(nu[3:6] * T)
    R = W.quat_to_rot(q[None])[0]
    assert np.abs(p - (p0 + T * nu[:3])).max() < 1e-10
    assert np.abs(s - (s0 + T * nu[6:])).max() < 1e-10
    assert np.abs(R - R_expected).max() < 1e-9          # constant w: exponentials commute, exact up to round-off
    assert np.abs(R - R_wrong).max() > 1e-2              # right-plus would be wrong
    assert abs(np.linalg.norm(q) - 1) < 1e-14
