"""First pieces of the TSID path through the C ABI against oracle/dynamics.py: the acceleration-level
integrator (FloatingBaseSystemAccelerationKinematics + ForwardEuler) and the free-floating mass matrix / gravity forces."""
import numpy as np
import pytest

import blf_b200  # noqa: F401
from blf_b200 import capi, models as M, tasks as T, workloads as W
from oracle import dynamics as D

pytestmark = pytest.mark.gpu


def test_acceleration_kinematics_forward_euler_vs_oracle_and_closed_form():
    rng = np.random.default_rng(4)
    N, n, dt, K = 37, 23, 0.002, 25
    p, q = rng.normal(size=(N, 3)), W.so3_exp(rng.uniform(-1, 1, (N, 3)))
    s, dp, om, ds = rng.normal(size=(N, n)), rng.normal(size=(N, 3)), rng.normal(size=(N, 3)), rng.normal(size=(N, n))
    acc = rng.normal(size=(N, n + 6))
    out = capi.integrate_acceleration_state(p, q, s, dp, om, ds, acc, dt, K)
    ref = D.integrate_acceleration(p, q, s, dp, om, ds, acc, dt, K)
    for a, b in zip(out, ref):
        assert np.abs(a - b).max() <= 1e-12
    # closed form of explicit Euler under a constant acceleration: v_K = v_0 + K dt a, x_K = x_0 + K dt v_0 + dt^2 a K (K - 1) / 2
    assert np.abs(out[3] - (dp + K * dt * acc[:, :3])).max() <= 1e-12
    assert np.abs(out[0] - (p + K * dt * dp + dt * dt * acc[:, :3] * K * (K - 1) / 2)).max() <= 1e-12
    assert np.abs(out[2] - (s + K * dt * ds + dt * dt * acc[:, 6:] * K * (K - 1) / 2)).max() <= 1e-12
    assert np.abs(np.linalg.norm(out[1], axis=1) - 1.0).max() <= 1e-14
    with pytest.raises(capi.BlfikError):
        capi.integrate_acceleration_state(p, q, s, dp, om, ds, acc, 0.0, 1)


@pytest.mark.parametrize("mk", [M.ergocub23, M.icub32, lambda: M.random_tree(17, 5)])
def test_mass_matrix_and_gravity_forces_vs_finite_difference_oracle(mk):
    m = mk()
    N = 5
    rng = np.random.default_rng(9)
    p, q = rng.uniform(-1, 1, (N, 3)), W.so3_exp(rng.uniform(-1.2, 1.2, (N, 3)))
    s = m.q_nominal + rng.uniform(-0.6, 0.6, (N, m.n))
    tasks = [T.joint_tracking_task(m, "j", np.ones(m.n), priority=1, weight=np.ones(m.n))]
    prob = capi.Problem(capi.Model(m), tasks, N)
    prob.set_state(p, q, s)
    Mg, Gg = prob.dynamics(9.81)
    nv = m.n + 6
    for i in range(N):
        Mo, Go = D.mass_matrix_fd(m, p[i], q[i], s[i]), D.gravity_forces_fd(m, p[i], q[i], s[i], 9.81)
        assert np.abs(Mg[i] - Mo).max() <= 1e-7 * (1 + np.abs(Mo).max()), np.abs(Mg[i] - Mo).max()
        assert np.abs(Gg[i] - Go).max() <= 1e-6 * (1 + np.abs(Go).max())
        assert np.abs(Mg[i] - Mg[i].T).max() <= 1e-12 * np.abs(Mg[i]).max()
        assert np.linalg.eigvalsh(0.5 * (Mg[i] + Mg[i].T)).min() > 0.0
        assert np.abs(Mg[i][:3, :3] - m.mass.sum() * np.eye(3)).max() <= 1e-11 * m.mass.sum()   # the whole robot translating
        assert abs(Gg[i][2] - 9.81 * m.mass.sum()) <= 1e-10 * m.mass.sum() and np.abs(Gg[i][:2]).max() <= 1e-12
        nu = rng.normal(size=nv)                       # kinetic energy of a random velocity, from link velocities
        v, w = D.link_velocities_fd(m, p[i], q[i], s[i])
        vk, wk = np.einsum("i,ikc->kc", nu, v), np.einsum("i,ikc->kc", nu, w)
        R, _ = D._links(m, p[i], q[i], s[i])
        IW = np.einsum("kij,kjl,kml->kim", R, m.inertia, R)
        E = 0.5 * float(m.mass @ (vk * vk).sum(axis=1)) + 0.5 * float(np.einsum("kc,kcd,kd->", wk, IW, wk))
        assert abs(0.5 * nu @ Mg[i] @ nu - E) <= 1e-7 * (1 + E)
    prob.close()


@pytest.mark.parametrize("mk", [M.ergocub23, M.icub32, lambda: M.random_tree(17, 5)])
def test_bias_forces_and_bias_accelerations(mk):
    """blfik_bias_forces: h = C(q, nu) nu + g(q), Jdot nu of frames and of the CoM, MIXED representation.
    (a) against the projected Newton-Euler restatement (oracle/dynamics.py: momenta from Jacobians written down from the
    forward kinematics, differentiated in time by finite differences; those Jacobians are themselves checked against pure
    finite differences of the link poses);  (b) against Lagrange's equations formed from the GPU's own mass matrix -- for the
    base translation and the joints, which are true coordinates, h_i - g_i = (Mdot nu)_i - 1/2 nu' (dM/dq_i) nu -- which uses no
    Jacobian at all;  (c) the power identity nu'(h - g) = 1/2 nu' Mdot nu on every row including the base rotation;
    (d) frame bias accelerations against finite differences in time of the oracle's frame Jacobians;  (e) nu = 0 gives the
    gravity forces, and the bias terms are quadratic in nu."""
    from oracle.oracle import Oracle
    m = mk()
    n, nv = m.n, m.n + 6
    rng = np.random.default_rng(12)
    N = 3
    p, q = rng.uniform(-1, 1, (N, 3)), W.so3_exp(rng.uniform(-1.2, 1.2, (N, 3)))
    s = m.q_nominal + rng.uniform(-0.6, 0.6, (N, n))
    nu = rng.uniform(-1.5, 1.5, (N, nv))
    g = 9.81
    tasks = [T.joint_tracking_task(m, "j", np.ones(n), priority=1, weight=np.ones(n))]
    frames = list(rng.choice(len(m.frame_names), size=min(4, len(m.frame_names)), replace=False))
    prob = capi.Problem(capi.Model(m), tasks, N)
    prob.set_state(p, q, s)
    h, fb, cb = prob.bias_forces(nu, frames, g)
    Mg, Gg = prob.dynamics(g)
    kin = prob.kinematics(frames)
    assert np.abs(prob.last_frame_velocities - np.einsum("nfrc,nc->nfr", kin["frame_jac"], nu)).max() <= 1e-12     # J nu of the frames
    o = Oracle(m, tasks)
    d = 1e-5
    for i in range(N):
        # the oracle's Jacobians against finite differences of the poses
        Jv, Jw, _, _ = D.link_jacobians(m, p[i], q[i], s[i])
        vfd, wfd = D.link_velocities_fd(m, p[i], q[i], s[i])
        assert np.abs(Jv - vfd.transpose(1, 2, 0)).max() <= 1e-7 and np.abs(Jw - wfd.transpose(1, 2, 0)).max() <= 1e-7
        # (a)
        ho = D.bias_forces_fd(m, p[i], q[i], s[i], nu[i], g, d)
        assert np.abs(h[i] - ho).max() <= 2e-6 * (1 + np.abs(ho).max()), (np.abs(h[i] - ho).max(), np.abs(ho).max())
        assert np.abs(cb[i] - D.com_bias_acc_fd(m, p[i], q[i], s[i], nu[i], d)).max() <= 1e-6
        # (d)
        def jac_nu(state):
            R, P = o.fk(*state)
            return np.stack([o.frame_jacobian(R, P, int(f)) @ nu[i] for f in frames])
        fbo = (jac_nu(D.moved(p[i], q[i], s[i], nu[i], d)) - jac_nu(D.moved(p[i], q[i], s[i], nu[i], -d))) / (2 * d)
        assert np.abs(fb[i] - fbo).max() <= 2e-6 * (1 + np.abs(fbo).max())
    # (b), (c): mass matrices of the moved states, one batch: [robot][+-][along nu | along variable j]
    dd = 1e-4
    for i in range(N):
        states = [D.moved(p[i], q[i], s[i], nu[i], sg * dd) for sg in (1, -1)]
        coords = [c for c in range(nv) if not 3 <= c < 6]
        for c in coords:
            e = np.zeros(nv)
            e[c] = 1.0
            states += [D.moved(p[i], q[i], s[i], e, sg * dd) for sg in (1, -1)]
        pb = capi.Problem(capi.Model(m), tasks, len(states))
        pb.set_state(np.stack([x[0] for x in states]), np.stack([x[1] for x in states]), np.stack([x[2] for x in states]))
        Ms, _ = pb.dynamics(g)
        pb.close()
        Mdot = (Ms[0] - Ms[1]) / (2 * dd)
        cor = h[i] - Gg[i]
        assert abs(nu[i] @ cor - 0.5 * nu[i] @ Mdot @ nu[i]) <= 1e-5 * (1 + abs(nu[i] @ cor))
        for k, c in enumerate(coords):
            dM = (Ms[2 + 2 * k] - Ms[3 + 2 * k]) / (2 * dd)
            lag = (Mdot @ nu[i])[c] - 0.5 * nu[i] @ dM @ nu[i]
            assert abs(cor[c] - lag) <= 1e-5 * (1 + np.abs(cor).max()), (c, cor[c], lag)
    # (e)
    h0, fb0, cb0 = prob.bias_forces(np.zeros((N, nv)), frames, g)
    assert np.abs(h0 - Gg).max() <= 1e-10 * (1 + np.abs(Gg).max()) and np.abs(fb0).max() == 0.0 and np.abs(cb0).max() == 0.0
    h2, fb2, cb2 = prob.bias_forces(2.0 * nu, frames, g)
    assert np.abs((h2 - Gg) - 4.0 * (h - Gg)).max() <= 1e-9 * (1 + np.abs(h).max())
    assert np.abs(fb2 - 4.0 * fb).max() <= 1e-10 * (1 + np.abs(fb).max()) and np.abs(cb2 - 4.0 * cb).max() <= 1e-10
    with pytest.raises(capi.BlfikError):
        prob.bias_forces(nu, [len(m.frame_names)], g)
    prob.close()
