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
