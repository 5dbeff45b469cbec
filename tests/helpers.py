"""Shared helpers of the test-suite (pure numpy; independent of both the oracle and the CUDA path)."""
import numpy as np

import blf_b200  # noqa: F401
from blf_b200 import workloads as W


def skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])


def rodrigues(phi):
    th = np.linalg.norm(phi)
    if th < 1e-12:
        return np.eye(3) + skew(phi)
    k = phi / th
    K = skew(k)
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K


def random_state(model, rng, spread=0.4):
    p = rng.uniform(-0.5, 0.5, 3)
    q = W.so3_exp(rng.uniform(-0.8, 0.8, 3)[None])[0]
    s = model.q_nominal + rng.uniform(-spread, spread, model.n)
    return p, q, s


def perturb_state(p, q, s, dnu, eps):
    """(p, R, s) moved along the MIXED velocity dnu for a time eps (left-plus on SO(3))."""
    q2 = W.quat_mul(W.so3_exp((dnu[3:6] * eps)[None]), q[None])[0]
    return p + dnu[:3] * eps, q2, s + dnu[6:] * eps


def rot_log(R):
    c = np.clip((np.trace(R) - 1) / 2, -1, 1)
    th = np.arccos(c)
    if th < 1e-9:
        return np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]]) / 2
    return th / (2 * np.sin(th)) * np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])


def solve_qp_active_set(H, g, C, l, u, tol=1e-10, max_iter=200):
    """Independent dense primal active-set-free cross-check: solve the QP  min 1/2x'Hx+g'x, l<=Cx<=u
    with scipy's SLSQP started from a feasible least-squares point.  Returns x."""
    from scipy.optimize import minimize
    nv = H.shape[0]
    eq = l == u
    cons = []
    if eq.any():
        cons.append({"type": "eq", "fun": lambda x: C[eq] @ x - u[eq], "jac": lambda x: C[eq]})
    fin_u = (~eq) & (u < 1e29)
    fin_l = (~eq) & (l > -1e29)
    if fin_u.any():
        cons.append({"type": "ineq", "fun": lambda x: u[fin_u] - C[fin_u] @ x, "jac": lambda x: -C[fin_u]})
    if fin_l.any():
        cons.append({"type": "ineq", "fun": lambda x: C[fin_l] @ x - l[fin_l], "jac": lambda x: C[fin_l]})
    x0 = np.linalg.lstsq(C[eq], u[eq], rcond=None)[0] if eq.any() else np.zeros(nv)
    r = minimize(lambda x: 0.5 * x @ H @ x + g @ x, x0, jac=lambda x: H @ x + g, constraints=cons, method="SLSQP",
                 options={"ftol": 1e-15, "maxiter": 500})
    return r.x


def referee(H, g, C, l, u):
    """Exact verdict on one instance of the reference's QP (min 1/2 x'Hx + g'x, l <= Cx <= u), independent of both solvers:
    feasibility by scipy's HiGHS linear programme, the optimum by SLSQP.  Returns (feasible, x or None)."""
    from scipy.optimize import linprog
    nv = H.shape[0]
    eq = l == u
    fin_u, fin_l = (~eq) & (u < 1e29), (~eq) & (l > -1e29)
    A_ub = np.vstack([C[fin_u], -C[fin_l]]) if (fin_u.any() or fin_l.any()) else None
    b_ub = np.concatenate([u[fin_u], -l[fin_l]]) if A_ub is not None else None
    r = linprog(np.zeros(nv), A_ub=A_ub, b_ub=b_ub, A_eq=C[eq] if eq.any() else None, b_eq=u[eq] if eq.any() else None,
                bounds=[(None, None)] * nv, method="highs")
    if r.status == 2:
        return False, None
    return True, solve_qp_active_set(H, g, C, l, u)


def eps_optimal(H, g, C, l, u, x, x_exact, eps=1e-7):
    """What a status "solved" of the reference's solver means at its accuracy setting (OSQP, eps_abs = eps_rel = eps): the point
    violates no row by more than eps (1 + |Cx|_inf) and its cost is not above the optimum's.  On an ill-conditioned instance
    (multipliers of 1e4 and more: a violation of 5e-8 moves the cost by 4e-4 and the point by 3e-5) two such points differ by
    more than the 1e-5 the well-conditioned instances are held to; this is the test such an instance gets instead."""
    Cx = C @ x
    viol = max(np.maximum(Cx - u, 0.0).max(), np.maximum(l - Cx, 0.0).max())
    f = lambda v: 0.5 * v @ H @ v + g @ v
    return viol <= eps * (1.0 + np.abs(Cx).max()) and f(x) <= f(x_exact) + 1e-12 * (1.0 + abs(f(x_exact)))
