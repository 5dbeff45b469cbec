"""dynamics.py -- TEST INFRASTRUCTURE (only tests/ may import it).

CPU restatements, in numpy, of the first TSID-side pieces:

* `integrate_acceleration`: FloatingBaseSystemAccelerationKinematics under ForwardEuler
  (src/ContinuousDynamicalSystem/src/FloatingBaseSystemAccelerationKinematics.cpp:20-75,
  include/.../ForwardEuler.h:72-124): x_{k+1} = x_k + dt f(x_k, u) component by component, the rotation by the left-plus
  Exp(omega dt) R (`tangent * dt + group`).
* `mass_matrix_fd` / `gravity_forces_fd`: the free-floating mass matrix and the generalised gravity forces in the MIXED
  representation (iDynTree::KinDynComputations::getFreeFloatingMassMatrix / generalizedGravityForces, external; read by
  src/TSID/src/BaseDynamicsTask.cpp and JointDynamicsTask.cpp) WITHOUT any Jacobian formula: link velocities come from central
  finite differences of the forward kinematics along each velocity variable, M_ij = sum_k m_k v_k(e_i).v_k(e_j) +
  w_k(e_i)' I_k^W w_k(e_j), G_i = dU/d(e_i), U = sum_k m_k g z_k.  Parity unpinned (iDynTree is not in this image); pinned here by
  physics: kinetic and potential energy.
"""
from __future__ import annotations

import numpy as np

import blf_b200  # noqa: F401
from blf_b200 import workloads as W


def _exp_quat(phi):
    th = np.linalg.norm(phi)
    k = np.sin(0.5 * th) / th if th > 1e-12 else 0.5
    return np.array([k * phi[0], k * phi[1], k * phi[2], np.cos(0.5 * th)])


def _quat_mul(a, b):
    return np.array([a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1], a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0],
                     a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3], a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2]])


def integrate_acceleration(p, q, s, dp, omega, ds, acc, dt, nsteps=1):
    p, q, s, dp, omega, ds = (np.array(a, dtype=np.float64) for a in (p, q, s, dp, omega, ds))
    for i in range(p.shape[0]):
        for _ in range(nsteps):
            qn = _quat_mul(_exp_quat(omega[i] * dt), q[i])
            p[i] = p[i] + dp[i] * dt
            s[i] = s[i] + ds[i] * dt
            q[i] = qn / np.linalg.norm(qn)
            dp[i] = dp[i] + acc[i, :3] * dt
            omega[i] = omega[i] + acc[i, 3:6] * dt
            ds[i] = ds[i] + acc[i, 6:] * dt
    return p, q, s, dp, omega, ds


def _perturbed(p, q, s, i, eps):
    """state moved by eps along velocity variable i (MIXED: base linear in the world, base angular in the world about the base
    origin, joint i - 6)"""
    p2, q2, s2 = p.copy(), q.copy(), s.copy()
    if i < 3:
        p2[i] += eps
    elif i < 6:
        phi = np.zeros(3)
        phi[i - 3] = eps
        q2 = _quat_mul(_exp_quat(phi), q)
    else:
        s2[i - 6] += eps
    return p2, q2, s2


def _links(model, p, q, s):
    R, P = W.fk_batch(model, p[None], q[None], s[None])
    R, P = R[0], P[0]
    c = P + np.einsum("kij,kj->ki", R, model.com)
    return R, c


def link_velocities_fd(model, p, q, s, eps=1e-6):
    """v[i, k] (CoM velocity of link k) and w[i, k] (its angular velocity, world frame) for a unit velocity along variable i"""
    nv, nl = model.n + 6, model.n + 1
    v, w = np.zeros((nv, nl, 3)), np.zeros((nv, nl, 3))
    for i in range(nv):
        Rp, cp = _links(model, *_perturbed(p, q, s, i, eps))
        Rm, cm = _links(model, *_perturbed(p, q, s, i, -eps))
        v[i] = (cp - cm) / (2 * eps)
        dR = np.einsum("kij,klj->kil", Rp, Rm)          # R+ R-^T = Exp(2 eps w)
        w[i] = np.stack([dR[:, 2, 1] - dR[:, 1, 2], dR[:, 0, 2] - dR[:, 2, 0], dR[:, 1, 0] - dR[:, 0, 1]], axis=1) / (4 * eps)
    return v, w


def mass_matrix_fd(model, p, q, s, eps=1e-6):
    v, w = link_velocities_fd(model, p, q, s, eps)
    R, _ = _links(model, p, q, s)
    IW = np.einsum("kij,kjl,kml->kim", R, model.inertia, R)
    M = np.einsum("k,ikc,jkc->ij", model.mass, v, v) + np.einsum("ikc,kcd,jkd->ij", w, IW, w)
    return M


def gravity_forces_fd(model, p, q, s, g=9.80665, eps=1e-6):
    nv = model.n + 6
    G = np.zeros(nv)
    for i in range(nv):
        _, cp = _links(model, *_perturbed(p, q, s, i, eps))
        _, cm = _links(model, *_perturbed(p, q, s, i, -eps))
        G[i] = g * float(model.mass @ (cp[:, 2] - cm[:, 2])) / (2 * eps)
    return G


# ---- velocity-dependent terms: generalised bias forces and bias accelerations (kinDyn->generalizedBiasForces, getFrameBiasAcc,
# getCenterOfMassBiasAcc; read by src/TSID/src/BaseDynamicsTask.cpp:178, JointDynamicsTask.cpp:233, SE3Task.cpp:254, CoMTask.cpp:195).
# Restated as the projected Newton-Euler equations with NO recursion over velocities: per link, linear momentum m Jv nu and
# angular momentum I^W Jw nu from Jacobians written down directly from the forward kinematics; their time derivative along the
# motion with nu held constant by central finite differences of the state (MIXED: p + d v_B, Exp(d w_B) R, s + d sd);
#     h = sum_k Jv_k' (d/dt (m_k v_ck) + m_k g z) + Jw_k' d/dt (I_k^W w_k).

def _ancestors(model):
    nl = model.n + 1
    anc = [[] for _ in range(nl)]          # links on the path base -> k with a joint (k itself included, the base excluded)
    for k in range(1, nl):
        anc[k] = anc[int(model.parent[k])] + [k]
    return anc


def link_jacobians(model, p, q, s):
    """Jv[k] (3 x nv, CoM of link k), Jw[k] (3 x nv), c[k], R[k]: MIXED representation, straight from the link poses"""
    R, P = W.fk_batch(model, p[None], q[None], s[None])
    R, P = R[0], P[0]
    nl, nv = model.n + 1, model.n + 6
    c = P + np.einsum("kij,kj->ki", R, model.com)
    Jv, Jw = np.zeros((nl, 3, nv)), np.zeros((nl, 3, nv))
    anc = _ancestors(model)
    for k in range(nl):
        Jv[k, :, :3] = np.eye(3)
        Jw[k, :, 3:6] = np.eye(3)
        r = c[k] - P[0]
        Jv[k, :, 3:6] = -np.array([[0, -r[2], r[1]], [r[2], 0, -r[0]], [-r[1], r[0], 0]])
        for a in anc[k]:
            j = int(model.dof[a])
            ax = R[a] @ model.axis[a]
            Jw[k, :, 6 + j] = ax
            Jv[k, :, 6 + j] = np.cross(ax, c[k] - P[a])
    return Jv, Jw, c, R


def moved(p, q, s, nu, d):
    """the state after moving for a time d with the generalised velocity nu held constant (MIXED)"""
    return p + d * nu[:3], _quat_mul(_exp_quat(d * nu[3:6]), q), s + d * nu[6:]


def bias_forces_fd(model, p, q, s, nu, g=9.80665, d=1e-5):
    def momenta(state):
        Jv, Jw, _, R = link_jacobians(model, *state)
        IW = np.einsum("kij,kjl,kml->kim", R, model.inertia, R)
        return model.mass[:, None] * np.einsum("kci,i->kc", Jv, nu), np.einsum("kcd,kd->kc", IW, np.einsum("kci,i->kc", Jw, nu))
    Pp, Lp = momenta(moved(p, q, s, nu, d))
    Pm, Lm = momenta(moved(p, q, s, nu, -d))
    dP, dL = (Pp - Pm) / (2 * d), (Lp - Lm) / (2 * d)
    Jv, Jw, _, _ = link_jacobians(model, p, q, s)
    dP[:, 2] += model.mass * g
    return np.einsum("kci,kc->i", Jv, dP) + np.einsum("kci,kc->i", Jw, dL)


def com_bias_acc_fd(model, p, q, s, nu, d=1e-5):
    def vcom(state):
        Jv, _, _, _ = link_jacobians(model, *state)
        return np.einsum("k,kci,i->c", model.mass, Jv, nu) / model.mass.sum()
    return (vcom(moved(p, q, s, nu, d)) - vcom(moved(p, q, s, nu, -d))) / (2 * d)
