import numpy as np

def hermite_coeffs(T, x0, v0, a0, x1, v1, a1):
    d = x1 - x0
    c3 = (20 * d - (8 * v1 + 12 * v0) * T - (3 * a0 - a1) * T * T) / (2 * T ** 3)
    c4 = (-30 * d + (14 * v1 + 16 * v0) * T + (3 * a0 - 2 * a1) * T * T) / (2 * T ** 4)
    c5 = (12 * d - 6 * (v1 + v0) * T - (a0 - a1) * T * T) / (2 * T ** 5)
    return np.stack([x0, v0, a0 / 2, c3, c4, c5])

def blocks(times):
    T = np.diff(times); n = len(times); m = n - 2
    L = np.zeros((m, 2, 2)); D = np.zeros((m, 2, 2)); U = np.zeros((m, 2, 2))
    for i in range(m):
        Tp, Tn = T[i], T[i + 1]
        L[i] = [[-24 / Tp**2, -3 / Tp], [-168 / Tp**3, -24 / Tp**2]]
        D[i] = [[-36 / Tp**2 + 36 / Tn**2, 9 / Tp + 9 / Tn], [-192 / Tp**3 - 192 / Tn**3, 36 / Tp**2 - 36 / Tn**2]]
        U[i] = [[24 / Tn**2, -3 / Tn], [-168 / Tn**3, 24 / Tn**2]]
    return T, L, D, U

def rhs(times, x):
    T = np.diff(times); m = len(times) - 2
    r = np.zeros((m, 2) + x.shape[1:])
    for i in range(m):
        Tp, Tn = T[i], T[i + 1]
        dp, dn = x[i + 1] - x[i], x[i + 2] - x[i + 1]
        r[i, 0] = -60 * dp / Tp**3 + 60 * dn / Tn**3
        r[i, 1] = -360 * dp / Tp**4 - 360 * dn / Tn**4
    return r

def solve_dense(times, x, v0, a0, vf, af):
    T, L, D, U = blocks(times); m = len(times) - 2
    r = rhs(times, x)
    r[0] -= np.einsum('ab,b...->a...', L[0], np.stack([v0, a0]))
    r[-1] -= np.einsum('ab,b...->a...', U[-1], np.stack([vf, af]))
    A = np.zeros((2 * m, 2 * m))
    for i in range(m):
        A[2*i:2*i+2, 2*i:2*i+2] = D[i]
        if i > 0: A[2*i:2*i+2, 2*i-2:2*i] = L[i]
        if i < m - 1: A[2*i:2*i+2, 2*i+2:2*i+4] = U[i]
    sol = np.linalg.solve(A, r.reshape(2 * m, -1)).reshape((m, 2) + x.shape[1:])
    return sol, np.linalg.cond(A)

def solve_thomas(times, x, v0, a0, vf, af, scaled=False):
    T, L, D, U = blocks(times); m = len(times) - 2
    r = rhs(times, x)
    r[0] -= np.einsum('ab,b...->a...', L[0], np.stack([v0, a0]))
    r[-1] -= np.einsum('ab,b...->a...', U[-1], np.stack([vf, af]))
    Dp = D.copy(); rp = r.copy()
    for i in range(1, m):
        M = L[i] @ np.linalg.inv(Dp[i - 1])
        Dp[i] = D[i] - M @ U[i - 1]
        rp[i] = r[i] - np.einsum('ab,b...->a...', M, rp[i - 1])
    sol = np.zeros_like(r)
    sol[-1] = np.einsum('ab,b...->a...', np.linalg.inv(Dp[-1]), rp[-1])
    for i in range(m - 2, -1, -1):
        sol[i] = np.einsum('ab,b...->a...', np.linalg.inv(Dp[i]), rp[i] - np.einsum('ab,b...->a...', U[i], sol[i + 1]))
    return sol

rng = np.random.default_rng(0)
for (t0, t1, n) in [(0.32, 642.0, 100), (0.32, 2.64, 100), (0.0, 1.0, 5), (0.0, 3.0, 64)]:
    for uneven in (False, True):
        c = rng.uniform(-1, 1, (6, 4))
        times = np.linspace(t0, t1, n)
        if uneven:
            times = np.sort(np.concatenate([[t0, t1], rng.uniform(t0, t1, n - 2)]))
        P = lambda t, k=0: sum(c[i] * np.prod(np.arange(i, i - k, -1)) * t ** (i - k) for i in range(k, 6))
        x = np.stack([P(t) for t in times])
        v0, a0, vf, af = P(times[0], 1), P(times[0], 2), P(times[-1], 1), P(times[-1], 2)
        sd, cond = solve_dense(times, x, v0, a0, vf, af)
        st = solve_thomas(times, x, v0, a0, vf, af)
        ve = np.stack([P(t, 1) for t in times[1:-1]]); ae = np.stack([P(t, 2) for t in times[1:-1]])
        rel = lambda a, b: np.abs(a - b).max() / np.abs(b).max()
        print(f"{t0}-{t1} n={n} uneven={uneven} cond={cond:.2e} min dT {np.diff(times).min():.2e} dense v {rel(sd[:,0],ve):.1e} a {rel(sd[:,1],ae):.1e}   thomas v {rel(st[:,0],ve):.1e} a {rel(st[:,1],ae):.1e}")
