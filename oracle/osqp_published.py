"""osqp_published.py -- TEST INFRASTRUCTURE (only tests/ and tools/ may import it; the product never does).

A second, independently written restatement of the QP solver behind `QPInverseKinematics::advance()`
(src/IK/src/QPInverseKinematics.cpp:61-146 sets the problem up through osqp-eigen v0.7.0, :494-528 solves it with OSQP v0.6.3):
the published OSQP algorithm -- Stellato, Banjac, Goulart, Bemporad, Boyd, "OSQP: an operator splitting solver for quadratic
programs", Math. Prog. Comp. 12 (2020), Algorithm 1 and sections 5.1 (modified Ruiz equilibration), 5.2 (per-constraint and
adaptive rho), 3.4 (termination) -- INCLUDING the scaling and the residual un-scaling that `oracle/blfik_oracle.c` leaves out, and
run at the settings the UNMODIFIED reference runs with (OSQP defaults: rho 0.1, sigma 1e-6, alpha 1.6, eps_abs = eps_rel = 1e-3,
10 scaling iterations, adaptive rho, no polish; QPInverseKinematics.cpp:339 only sets the verbosity).

Purpose: bound how far the answer of the reference at its stock accuracy sits from the optimum that the C
oracle (eps 1e-7 + polish) and the CUDA path return; `tests/test_oracle_osqp_defaults.py` reports the distribution on C3.

Parity unpinned: OSQP itself is not in this image.  One deliberate deviation: OSQP's default adaptive-rho interval is derived
from the wall-clock time of the first factorisation (non-deterministic); here rho is reconsidered every 25 iterations, the interval
its termination checks use and the value its timing rule is rounded to.
"""
from __future__ import annotations

import numpy as np

OSQP_INFTY = 1e30
RHO_MIN, RHO_MAX, RHO_EQ_FACTOR, RHO_TOL = 1e-6, 1e6, 1e3, 1e-4
MIN_SCALING, MAX_SCALING = 1e-4, 1e4


def _limit(v):
    v = np.where(v < MIN_SCALING, 1.0, v)
    return np.minimum(v, MAX_SCALING)


def ruiz_scale(P, q, A, l, u, iters=10):
    """Modified Ruiz equilibration of the KKT matrix [[P, A'], [A, 0]] with cost scaling (paper section 5.1, scaling.c)."""
    n, m = P.shape[0], A.shape[0]
    D, E, c = np.ones(n), np.ones(m), 1.0
    P, q, A = P.copy(), q.copy(), A.copy()
    for _ in range(iters):
        dn = np.maximum(np.abs(P).max(axis=0), np.abs(A).max(axis=0) if m else 0.0)
        en = np.abs(A).max(axis=1) if m else np.zeros(0)
        d = 1.0 / np.sqrt(_limit(dn))
        e = 1.0 / np.sqrt(_limit(en))
        P = d[:, None] * P * d[None, :]
        A = e[:, None] * A * d[None, :]
        q = d * q
        D, E = D * d, E * e
        # cost scaling: mean column norm of P against the norm of q
        cn = np.abs(P).max(axis=0).mean()
        qn = np.abs(q).max()
        cn = 1.0 if cn < MIN_SCALING else min(cn, MAX_SCALING)
        qn = 1.0 if qn < MIN_SCALING else min(qn, MAX_SCALING)
        gamma = 1.0 / max(cn, qn)
        P, q, c = gamma * P, gamma * q, c * gamma
    return P, q, A, l * E, u * E, D, E, c


def solve(P, q, A, l, u, rho=0.1, sigma=1e-6, alpha=1.6, eps_abs=1e-3, eps_rel=1e-3, max_iter=4000, check_every=25,
          adaptive_rho=True, scaling=10, eps_prim_inf=1e-4, eps_dual_inf=1e-4):
    """OSQP Algorithm 1 on min 1/2 x'Px + q'x s.t. l <= Ax <= u.  Returns (x, y, status, iterations); status 0 solved,
    1 maximum iterations reached, 2 primal infeasible, 3 dual infeasible (paper section 3.4: the certificates on the increments
    of y and x between two iterations, checked with the termination criteria; QPInverseKinematics::advance() fails on anything
    but solved / solved inaccurate, QPInverseKinematics.cpp:518-523)."""
    n, m = P.shape[0], A.shape[0]
    Ps, qs, As, ls, us, D, E, c = ruiz_scale(P, q, A, l, u, scaling) if scaling else (P, q, A, l, u, np.ones(n), np.ones(m), 1.0)
    Dinv, Einv = 1.0 / D, 1.0 / E
    eq = (us - ls) < RHO_TOL
    loose = (ls < -OSQP_INFTY * MIN_SCALING) & (us > OSQP_INFTY * MIN_SCALING)

    def rho_vec(r):
        v = np.full(m, r)
        v[eq] = RHO_EQ_FACTOR * r
        v[loose] = RHO_MIN
        return v

    def factor(rv):
        K = np.block([[Ps + sigma * np.eye(n), As.T], [As, -np.diag(1.0 / rv)]])
        return np.linalg.inv(K)   # dense, small: the "factorisation"

    rv = rho_vec(rho)
    Kinv = factor(rv)
    x, z, y = np.zeros(n), np.zeros(m), np.zeros(m)
    status, it = 1, 0
    for it in range(1, max_iter + 1):
        x_prev, y_prev = x, y
        rhs = np.concatenate([sigma * x - qs, z - y / rv])
        sol = Kinv @ rhs
        xt, nu = sol[:n], sol[n:]
        zt = z + (nu - y) / rv
        xn = alpha * xt + (1 - alpha) * x
        zr = alpha * zt + (1 - alpha) * z
        zn = np.clip(zr + y / rv, ls, us)
        y = y + rv * (zr - zn)
        x, z = xn, zn
        if it % check_every and it != max_iter:
            continue
        # residuals and tolerances in the UNSCALED problem (paper section 5.1, last paragraph; util.c compute_*_res)
        Ax, Px, Aty = As @ x, Ps @ x, As.T @ y
        r_prim = np.abs(Einv * (Ax - z)).max() if m else 0.0
        r_dual = np.abs(Dinv * (Px + qs + Aty)).max() / c
        e_prim = eps_abs + eps_rel * max(np.abs(Einv * Ax).max() if m else 0.0, np.abs(Einv * z).max() if m else 0.0)
        e_dual = eps_abs + eps_rel * max(np.abs(Dinv * Px).max(), np.abs(Dinv * Aty).max(), np.abs(Dinv * qs).max()) / c
        if r_prim <= e_prim and r_dual <= e_dual:
            status = 0
            break
        # certificates, in the unscaled problem (util.c is_primal_infeasible / is_dual_infeasible)
        dy = E * (y - y_prev)
        ndy = np.abs(dy).max() if m else 0.0
        if ndy > 1e-30:
            finite_u, finite_l = u < OSQP_INFTY * MIN_SCALING, l > -OSQP_INFTY * MIN_SCALING
            support = float(np.sum(np.where(finite_u, u, 0.0) * np.maximum(dy, 0.0)) + np.sum(np.where(finite_l, l, 0.0) * np.minimum(dy, 0.0)))
            unbounded_dir = np.any((dy > eps_prim_inf * ndy) & ~finite_u) or np.any((dy < -eps_prim_inf * ndy) & ~finite_l)
            if not unbounded_dir and support <= -eps_prim_inf * ndy and np.abs(Dinv * (As.T @ (y - y_prev))).max() <= eps_prim_inf * ndy:
                status = 2
                break
        dx = D * (x - x_prev)
        ndx = np.abs(dx).max()
        if ndx > 1e-30 and q @ dx <= -eps_dual_inf * ndx and np.abs(Dinv * (Ps @ (x - x_prev))).max() / c <= eps_dual_inf * ndx:
            Adx = Einv * (As @ (x - x_prev))
            ok_u = np.all((Adx <= eps_dual_inf * ndx) | (u >= OSQP_INFTY * MIN_SCALING))
            ok_l = np.all((Adx >= -eps_dual_inf * ndx) | (l <= -OSQP_INFTY * MIN_SCALING))
            if ok_u and ok_l:
                status = 3
                break
        if adaptive_rho:   # section 5.2: rho <- rho sqrt(normalised primal residual / normalised dual residual)
            pn = np.abs(Ax - z).max() / max(np.abs(Ax).max(), np.abs(z).max(), 1e-10) if m else 0.0
            dnm = np.abs(Px + qs + Aty).max() / max(np.abs(Px).max(), np.abs(Aty).max(), np.abs(qs).max(), 1e-10)
            new = float(np.clip(rho * np.sqrt(pn / max(dnm, 1e-10)), RHO_MIN, RHO_MAX))
            if new > 5.0 * rho or new < rho / 5.0:   # adaptive_rho_tolerance
                rho = new
                rv = rho_vec(rho)
                Kinv = factor(rv)
    return D * x, (E * y) / c, status, it
