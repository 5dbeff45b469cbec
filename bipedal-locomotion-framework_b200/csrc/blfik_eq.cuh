// blfik_eq.cuh -- the tick for models too wide for the register-resident kernel: 32 < nv <= 64 or more than 32 dense rows.
// BASELINE.json configs[0] and [1] are of this class (iCubGazeboV2_5 with 32 joints: SE3 feet + CoM as constraints, joint
// regularisation as cost).  One warp per robot, everything in that warp's shared memory, the QP in factored form.
//
// Without inequality rows (no JointLimitsTask / JointVelocityLimitsTask, no priority-0 JointTrackingTask) the QP of
// QPInverseKinematics::advance (QPInverseKinematics.cpp:442-546) is one linear solve (eq_solve_fixed with nothing fixed):
//     min 1/2 x'(D0 + Bc'Bc) x + (g0 - Bc'bc)'x   s.t.  Be x = be
// D0 the diagonal cost (joint regularisation), Bc the dense cost rows scaled by sqrt(weight), Be the equality rows.  It is
// never formed as H: with B = [Be; Bc] stacked in task order (K rows), E = diag(0 on equality rows, 1 on cost rows) and the
// proximal term 1/2 (x - x_k)' R (x - x_k), R = max(D0, sigma) - D0 (sigma on the base columns, where D0 is zero, nothing on a
// variable whose own cost already exceeds sigma), that makes D = D0 + R invertible, one step is
//     G z = B D^-1 q - [be; bc],   G = E + B D^-1 B'  (K x K, SPD),   x_{k+1} = D^-1 (q - B'z),   q = R x_k - g0
// (z = the multipliers of the equality rows and the residuals Bc x - bc of the cost rows).  G is factored once per tick and
// inverted (K <= 64); the steps repeat until x stops moving -- R |x_{k+1} - x_k| is the stationarity residual of
// the original problem, and Be x = be holds at every step.  Dependent equality rows are dropped by the factorisation as in
// the generic kernel and the equality residual is checked at the end (status 2 if they contradict each other).
#pragma once
#include "blfik_device.cuh"
#include "blfik_fast.cuh"   // dmma884

namespace blfik {

// One solve with the variables marked in w.y (0 free; +1 / -1 / 2: fixed at the upper / lower bound / at lo = hi) held at
// their bounds: they leave D^-1 (w.gp = 0), keep their value through w.r, and enter the right-hand side of the rows like any
// other known term.  Leaves x in w.x and z in w.lam; false if a pivot was not finite.
__device__ __forceinline__ int eq_prepare(const TickParams& P, const WM& w, int lane)
{
    const Layout& L = P.L;
    const int nv = L.nv;
    const double sigma = P.st.sigma;
    // Proximal weight per free variable, R_i = max(D0_i, sigma) - D0_i: only where the diagonal cost is missing (the base columns)
    // or below sigma.  D^-1 q = rho x_k + pc with rho = R / D (w.z) and the constant pc = -D^-1 g0 (w.r), so a step moves nothing
    // but the few variables F with rho > 0 (listed in w.act): the steps run on those and z, and the other variables are
    // evaluated once at the end.
    int nF = 0;
    for (int base = 0; base < nv; base += 32) {
        const int i = base + lane;
        bool f = false;
        if (i < nv) {
            const double code = L.box ? w.y[i] : 0.0;
            if (code != 0.0) {
                const double v = code > 0.0 && code < 2.0 ? w.hi[i] : w.lo[i];
                w.gp[i] = 0.0, w.z[i] = 0.0, w.r[i] = v, w.x[i] = v;
            } else {
                const double d0 = w.Hs[i], d = fmax(d0, sigma), dinv = 1.0 / d;
                w.gp[i] = dinv;
                w.z[i] = (d - d0) * dinv;
                w.r[i] = -dinv * w.g[i];
                f = d > d0;
            }
        }
        const unsigned m = __ballot_sync(kFull, f);
        if (f) w.act[nF + __popc(m & ((1u << lane) - 1u))] = i;
        nF += __popc(m);
    }
    __syncwarp();
    return nF;
}

__device__ __forceinline__ bool eq_solve_generic(const TickParams& P, const WM& w, int lane, int nF, int& steps, bool& converged)
{
    const Layout& L = P.L;
    const int nv = L.nv, K = L.K, LD = L.LD, LDE = L.LDE;
    // G = E + (B D^-1) B' as 8 x 8 tiles on the FP64 tensor path (DMMA m8n8k4, as in the register-resident kernel), tiles on
    // and below the diagonal only: one instruction per 256 multiply-adds where the lane-per-entry loop spent 1,500 per robot
    double gmax = 0.0;
    {
        const int gr = lane >> 2, tc = lane & 3;
        const int RT = (K + 7) >> 3, KS = (nv + 3) >> 2;
        for (int ti = 0; ti < RT; ti++) {
            const int row = 8 * ti + gr;
            const double* ba = w.Cs + (row < K ? row : 0) * LD;
            for (int tj = 0; tj <= ti; tj++) {
                const int rowb = 8 * tj + gr;
                const double* bb = w.Cs + (rowb < K ? rowb : 0) * LD;
                double c0 = 0.0, c1 = 0.0;
                for (int s = 0; s < KS; s++) {
                    const int col = 4 * s + tc;
                    const bool in = col < nv;
                    const double a = (in && row < K) ? ba[col] * w.gp[col] : 0.0;
                    const double b = (in && rowb < K) ? bb[col] : 0.0;
                    dmma884(c0, c1, a, b);
                }
                const int col = 8 * tj + 2 * tc;
                if (row < K) {
                    if (col < K) {
                        if (row == col) c0 += w.bd[row], gmax = fmax(gmax, c0);
                        w.Ss[row * LDE + col] = c0;
                    }
                    if (col + 1 < K) {
                        if (row == col + 1) c1 += w.bd[row], gmax = fmax(gmax, c1);
                        w.Ss[row * LDE + col + 1] = c1;
                    }
                }
            }
        }
    }
    gmax = warp_max(gmax);
    __syncwarp();
    // G = L L', right-looking, row i by lane; a pivot below the floor marks a row that depends on the previous ones: it is
    // left out (zero column, zero diagonal) exactly as chol_inplace(..., drop = true) does
    bool ok = true;
    {
        const double floor_ = 1e-11 * gmax + 1e-300;
        for (int k = 0; k < K; k++) {
            const double piv = w.Ss[k * LDE + k];
            const bool dropped = !(piv > floor_);
            if (dropped && !(piv == piv)) ok = false;
            const double inv = dropped ? 0.0 : rsqrt(piv), d = piv * inv;   // one reciprocal square root instead of sqrt and a division
            __syncwarp();
            for (int i = k + 1 + lane; i < K; i += 32) w.Ss[i * LDE + k] *= inv;
            if (lane == 0) w.Ss[k * LDE + k] = d;
            __syncwarp();
            for (int i = k + 1 + lane; i < K; i += 32) {
                const double lik = w.Ss[i * LDE + k];
                for (int j = k + 1; j <= i; j++) w.Ss[i * LDE + j] -= lik * w.Ss[j * LDE + k];
            }
            __syncwarp();
        }
    }
    // Linv = L^-1 (lower), column c by lane; a dropped row (zero diagonal) leaves its row and column zero
    for (int c = lane; c < K; c += 32) {
        for (int i = 0; i < K; i++) {
            double v = (i == c) ? 1.0 : 0.0;
            for (int j = c; j < i; j++) v -= w.Ss[i * LDE + j] * w.Ym[j * LDE + c];
            const double dg = w.Ss[i * LDE + i];
            w.Ym[i * LDE + c] = (i >= c && dg != 0.0) ? v / dg : 0.0;
        }
    }
    __syncwarp();
    for (int k = lane; k < K; k += 32) {   // constant part of the right-hand side: B pc - [be; bc]
        const double* bk = w.Cs + k * LD;
        double v = -w.b[k];
        for (int c = 0; c < nv; c++) v += bk[c] * w.r[c];
        w.wb[k] = v;
    }
    __syncwarp();
    const int max_steps = 60;
    converged = false;
    int it = 0;
    for (; it < max_steps && !converged; it++) {
        for (int k = lane; k < K; k += 32) {
            const double* bk = w.Cs + k * LD;
            double v = w.wb[k];
            for (int f = 0; f < nF; f++) {
                const int c = w.act[f];
                v += bk[c] * w.z[c] * w.x[c];
            }
            w.bp[k] = v;
        }
        __syncwarp();
        for (int k = lane; k < K; k += 32) {
            double v = 0.0;
            for (int j = 0; j <= k; j++) v += w.Ym[k * LDE + j] * w.bp[j];
            w.t[k] = v;
        }
        __syncwarp();
        for (int k = lane; k < K; k += 32) {
            double v = 0.0;
            for (int j = k; j < K; j++) v += w.Ym[j * LDE + k] * w.t[j];
            w.lam[k] = v;
        }
        __syncwarp();
        bool settled = true, finite = true;
        for (int f = lane; f < nF; f += 32) {
            const int c = w.act[f];
            double bz = 0.0;
            for (int k = 0; k < K; k++) bz += w.Cs[k * LD + c] * w.lam[k];
            const double v = w.z[c] * w.x[c] + w.r[c] - w.gp[c] * bz;
            settled = settled && (fabs(v - w.x[c]) <= 1e-10 * (1.0 + fabs(v)));
            finite = finite && (v == v);
            w.x[c] = v;
        }
        converged = __all_sync(kFull, settled);
        if (!__all_sync(kFull, finite)) break;
    }
    // every variable from the last z (one more step for those of F)
    for (int c = lane; c < nv; c += 32) {
        double bz = 0.0;
        for (int k = 0; k < K; k++) bz += w.Cs[k * LD + c] * w.lam[k];
        w.x[c] = w.z[c] * w.x[c] + w.r[c] - w.gp[c] * bz;
    }
    __syncwarp();
    steps += it;
    return ok;
}

// The same solve for K <= KP <= 24 dense rows and at most min(8, 31 - KP) proximal variables (C1 / C2: 15 rows, the six base
// columns), on the pattern of solve_setup (blfik_split.cuh):
//  * the Cholesky of G runs on rotating register rows, lane r = row r, and everything that needs L^-1 applied rides along in
//    the lanes that hold no row: a lane whose row is used up carries the forward substitution of its unit vector (so the
//    inverse factor costs no second pass), lane KP carries the constant right-hand side wb, lanes KP + 1 ... carry the columns
//    of B that belong to the proximal variables F;
//  * with t0 = L^-1 wb and V = L^-1 B_F the step x_F <- rho x_F + pc_F - D^-1_F B_F' G^-1 (wb + B_F rho x_F) is the affine map
//    x_F <- y + pc_F - D^-1_F (h + S y), y = rho x_F, S = V'V (|F| x |F|), h = V't0: a step costs |F| multiply-adds per lane
//    instead of two triangular products with L^-1; z and the other variables are evaluated once, after the last step.
// A pivot below the floor drops its row exactly as the generic path does (a zero reciprocal pivot zeroes the column of L,
// the row and the column of L^-1 and the rider entries).
template <int KP>
__device__ __forceinline__ bool eq_solve_rider(const TickParams& P, const WM& w, int lane, int nF, int& steps, bool& converged)
{
    static_assert(KP % 8 == 0 && KP <= 24, "rows by lanes, with spare lanes for the riders");
    const Layout& L = P.L;
    const int nv = L.nv, K = L.K, LD = L.LD;
    constexpr int LDG = ld2odd(KP), LDV = KP + 1, RT = KP / 8;
    double* Gs = w.Ss;            // G by columns, then row c: (L^-1)_{c, 0..c} followed by L_{c+1.., c}
    double* Vs = w.Wm;            // [1 + nF][LDV]: t0, V
    double* Sm = w.Wm + 9 * LDV;  // [9][9]: products of the columns of Vs
    for (int k = lane; k < K; k += 32) {   // constant part of the right-hand side: B pc - [be; bc]
        const double* bk = w.Cs + k * LD;
        double v0 = -w.b[k], v1 = 0.0;
        int c = 0;
        for (; c + 1 < nv; c += 2) v0 = fma(bk[c], w.r[c], v0), v1 = fma(bk[c + 1], w.r[c + 1], v1);
        if (c < nv) v0 = fma(bk[c], w.r[c], v0);
        w.wb[k] = v0 + v1;
    }
    double gmax = 0.0;
    {
        const int gr = lane >> 2, tc = lane & 3;
        const int KS = (nv + 3) >> 2;
#pragma unroll
        for (int ti = 0; ti < RT; ti++) {
            const int row = 8 * ti + gr;
            const double* ba = w.Cs + (row < K ? row : 0) * LD;
#pragma unroll
            for (int tj = 0; tj <= ti; tj++) {
                const int rowb = 8 * tj + gr;
                const double* bb = w.Cs + (rowb < K ? rowb : 0) * LD;
                double c0 = 0.0, c1 = 0.0;
                for (int s = 0; s < KS; s++) {
                    const int col = 4 * s + tc;
                    const bool in = col < nv;
                    const double a = (in && row < K) ? ba[col] * w.gp[col] : 0.0;
                    const double b = (in && rowb < K) ? bb[col] : 0.0;
                    dmma884(c0, c1, a, b);
                }
                const int col = 8 * tj + 2 * tc;
                const double er = row < K ? w.bd[row] : 1.0;   // padding rows: identity
                if (row == col) c0 += er, gmax = fmax(gmax, c0);
                if (row == col + 1) c1 += er, gmax = fmax(gmax, c1);
                Gs[col * LDG + row] = c0, Gs[(col + 1) * LDG + row] = c1;
                if (tj < ti) Gs[row * LDG + col] = c0, Gs[row * LDG + col + 1] = c1;
            }
        }
    }
    gmax = warp_max(gmax);
    __syncwarp();
    bool ok = true;
    {
        const double floor_ = 1e-11 * gmax + 1e-300;
        double gA[KP], gB[KP];
        const int myr = lane < KP ? lane : 0;
        const int rider = lane - KP;   // 0: wb, 1 + f: column act[f] of B
        const int rc = (rider >= 1 && rider <= nF) ? w.act[rider - 1] : 0;
#pragma unroll
        for (int j = 0; j < KP; j++) {
            double v = Gs[j * LDG + myr];
            if (rider >= 0) v = (j < K && rider <= nF) ? (rider == 0 ? w.wb[j] : w.Cs[j * LD + rc]) : 0.0;
            gA[j] = v;
        }
        __syncwarp();
        double* vout = Vs + (rider >= 0 && rider <= nF ? rider : 0) * LDV;
        const bool rides = rider >= 0 && rider <= nF;
#pragma unroll 1
        for (int k = 0; k < KP; k += 2) {
            {
                const double piv = __shfl_sync(kFull, gA[0], k);
                const bool dropped = !(piv > floor_);
                if (dropped && !(piv == piv)) ok = false;
                const double rinv = dropped ? 0.0 : rsqrt(piv);
                const double f = gA[0] * rinv;
                if (lane < KP) Gs[k * LDG + lane] = (lane == k) ? rinv : f;
                if (rides) vout[k] = f;
                __syncwarp();
                rot_update<KP, 0>(gA, gB, lane == k ? f + rinv : f, Gs + k * LDG + k);
            }
            {
                const int k1 = k + 1;
                const double piv = __shfl_sync(kFull, gB[0], k1);
                const bool dropped = !(piv > floor_);
                if (dropped && !(piv == piv)) ok = false;
                const double rinv = dropped ? 0.0 : rsqrt(piv);
                const double f = gB[0] * rinv;
                if (lane < KP) Gs[k1 * LDG + lane] = (lane == k1) ? rinv : f;
                if (rides) vout[k1] = f;
                __syncwarp();
                rot_update<KP, 1>(gB, gA, lane == k1 ? f + rinv : f, Gs + k1 * LDG + k1);
            }
        }
    }
    __syncwarp();
    // products of the columns (t0, V_1 ... V_nF): entry (a, b) by lane over the 9 x 8 slots
    for (int e = lane; e < 72; e += 32) {
        const int a = e >> 3, b = (e & 7) + 1;
        if (a <= nF && b <= nF) {
            const double *va = Vs + a * LDV, *vb = Vs + b * LDV;
            double s0 = 0.0, s1 = 0.0;
#pragma unroll
            for (int k = 0; k < KP; k += 2) s0 = fma(va[k], vb[k], s0), s1 = fma(va[k + 1], vb[k + 1], s1);
            Sm[a * 9 + b] = s0 + s1;
        }
    }
    __syncwarp();
    const bool inF = lane < nF;
    const int cF = inF ? w.act[lane] : 0;
    double srow[8];
#pragma unroll
    for (int g = 0; g < 8; g++) srow[g] = (inF && g < nF) ? Sm[(lane + 1) * 9 + g + 1] : 0.0;
    const double hf = inF ? Sm[lane + 1] : 0.0;
    const double zf = inF ? w.z[cF] : 0.0, rf = inF ? w.r[cF] : 0.0, gpf = inF ? w.gp[cF] : 0.0;
    double xf = inF ? w.x[cF] : 0.0;
    const int max_steps = 60;
    converged = false;
    int it = 0;
    for (; it < max_steps && !converged; it++) {
        const double y = zf * xf;
        double acc = hf;
#pragma unroll
        for (int g = 0; g < 8; g++) acc = fma(srow[g], __shfl_sync(kFull, y, g), acc);
        const double v = y + rf - gpf * acc;
        const bool settled = !inF || fabs(v - xf) <= 1e-10 * (1.0 + fabs(v));
        const bool finite = v == v;
        xf = v;
        converged = __all_sync(kFull, settled);
        if (!__all_sync(kFull, finite)) break;
    }
    // z = L^-T (t0 + V y) from the last iterate, then every variable (one more step for those of F)
    {
        const double y = zf * xf;
        double t = lane < KP ? Vs[lane] : 0.0;
#pragma unroll
        for (int g = 0; g < 8; g++) {
            const double yg = __shfl_sync(kFull, y, g);
            if (g < nF && lane < KP) t = fma(Vs[(g + 1) * LDV + lane], yg, t);
        }
        if (lane < KP) w.t[lane] = t;
        if (inF) w.x[cF] = xf;
        __syncwarp();
        if (lane < K) {
            double v0 = 0.0, v1 = 0.0;
#pragma unroll
            for (int j = 0; j < KP; j += 2) {
                v0 = fma(j >= lane ? Gs[j * LDG + lane] : 0.0, w.t[j], v0);
                v1 = fma(j + 1 >= lane ? Gs[(j + 1) * LDG + lane] : 0.0, w.t[j + 1], v1);
            }
            w.lam[lane] = v0 + v1;
        }
        __syncwarp();
    }
    for (int c = lane; c < nv; c += 32) {
        const double* bc = w.Cs + c;
        double b0 = 0.0, b1 = 0.0;
        int k = 0;
        for (; k + 1 < K; k += 2) b0 = fma(bc[k * LD], w.lam[k], b0), b1 = fma(bc[(k + 1) * LD], w.lam[k + 1], b1);
        if (k < K) b0 = fma(bc[k * LD], w.lam[k], b0);
        w.x[c] = w.z[c] * w.x[c] + w.r[c] - w.gp[c] * (b0 + b1);
    }
    __syncwarp();
    steps += it;
    return ok;
}

template <int KP>
__device__ __forceinline__ bool eq_solve_fixed(const TickParams& P, const WM& w, int lane, int& steps, bool& converged)
{
    const int nF = eq_prepare(P, w, lane);
    if constexpr (KP > 0) {
        constexpr int cap = 31 - KP < 8 ? 31 - KP : 8;
        if (nF <= cap) return eq_solve_rider<KP>(P, w, lane, nF, steps, converged);
    }
    return eq_solve_generic(P, w, lane, nF, steps, converged);
}

// The QP of one tick.  Without a box (L.box == 0) it is the single solve above.  With one (JointLimitsTask,
// JointVelocityLimitsTask, a priority-0 JointTrackingTask) the active set is found by block principal pivoting: every solve is
// followed by the test of the complementarity conditions -- a free joint outside its bounds is fixed there, a fixed joint whose
// multiplier has the wrong sign is released -- all at once while the number of violations keeps falling (the primal-dual
// active-set step of the register-resident kernel), one joint at a time, lowest index first, when it stalls (Murty's least-index
// rule, the safeguard of Judice and Pires that makes the pivoting finite on a strictly convex problem).  No violation left is
// the KKT certificate: primal feasible, dual feasible, stationary to the tolerance of the steps.
template <int KP>
__device__ __forceinline__ QpResult k3_solve_eq(const TickParams& P, const WM& w, bool warm, long long robot, int lane)
{
    const Layout& L = P.L;
    const int nv = L.nv, K = L.K, LD = L.LD, n = nv - 6;
    QpResult res = {0, 0};
    bool infeasible_box = false;
    double gs = 0.0;
    for (int i = lane; i < nv; i += 32) {
        w.x[i] = warm ? P.wx[(size_t)nv * robot + i] : 0.0;
        gs = fmax(gs, fabs(w.g[i]));
        if (L.box) {
            double code = 0.0;
            if (i >= 6) {
                if (w.lo[i] > w.hi[i]) infeasible_box = true;
                if (w.lo[i] >= w.hi[i]) code = 2.0;
                else if (warm) {
                    const double c = P.wy[(size_t)n * robot + i - 6];
                    code = (c == 1.0 && w.hi[i] < kInfHalf) ? 1.0 : ((c == -1.0 && w.lo[i] > -kInfHalf) ? -1.0 : 0.0);
                }
            }
            w.y[i] = code;
        }
    }
    gs = warp_max(gs);
    if (__any_sync(kFull, infeasible_box)) {
        res.status = 2;
        return res;
    }
    __syncwarp();
    bool ok = true, converged = false, certified = false;
    const int max_pivots = L.box ? 32 : 1;   // a feasible problem needs a handful; the bound is what an infeasible one costs
    int best = 1 << 30, stall = 0;
    for (int piv = 0; piv < max_pivots; piv++) {
        ok = eq_solve_fixed<KP>(P, w, lane, res.iters, converged) && ok;
        if (!L.box) {
            // the equality rows hold at every step by construction; what the step cap can leave unfinished is the cost
            // (a cost weight far below sigma contracts slowly), and then the tick must not claim SOLVED
            certified = converged;
            break;
        }
        // complementarity: new code per joint (also from steps that have not settled: an over-full set on the way makes the
        // rows nearly dependent and the steps slow; only the certificate needs them converged)
        int ninf = 0, first = nv;
        for (int base = 0; base < nv; base += 32) {
            const int i = base + lane;
            bool flip = false;
            double ncode = 0.0;
            if (i >= 6 && i < nv) {
                const double code = w.y[i];
                ncode = code;
                if (code == 0.0) {
                    const double v = w.x[i];
                    if (v > w.hi[i] + 1e-9 * (1.0 + fabs(w.hi[i]))) ncode = 1.0;
                    else if (v < w.lo[i] - 1e-9 * (1.0 + fabs(w.lo[i]))) ncode = -1.0;
                } else if (code != 2.0) {
                    double rg = w.Hs[i] * w.x[i] + w.g[i];     // gradient of the Lagrangian without the bound's own multiplier
                    for (int k = 0; k < K; k++) rg += w.Cs[k * LD + i] * w.lam[k];
                    const double tol = 1e-9 * (1.0 + gs);
                    if ((code > 0.0 && rg > tol) || (code < 0.0 && rg < -tol)) ncode = 0.0;
                }
                flip = ncode != code;
            }
            const unsigned m = __ballot_sync(kFull, flip);
            if (m && first == nv) first = base + __ffs(m) - 1;
            ninf += __popc(m);
            if (i < nv) w.xb[i] = ncode;   // parked until the rule below decides which flips happen
        }
        __syncwarp();
        if (ninf == 0) {
            certified = converged;
            break;
        }
        if (ninf < best) best = ninf, stall = 0;
        else stall++;
        if (stall <= 2) {
            for (int i = lane; i < nv; i += 32) w.y[i] = w.xb[i];
        } else if (lane == 0) {
            w.y[first] = w.xb[first];
        }
        __syncwarp();
    }
    // equality residual: exact at every step unless rows were dropped as dependent and contradict the kept ones
    bool bad = false;
    for (int k = lane; k < K; k += 32) {
        if (w.bd[k] != 0.0) continue;
        const double* bk = w.Cs + k * LD;
        double v = -w.b[k];
        for (int i = 0; i < nv; i++) v += bk[i] * w.x[i];
        if (!(fabs(v) <= 1e-6 * (1.0 + fabs(w.b[k])))) bad = true;
    }
    bad = __any_sync(kFull, bad);
    // contradicting rows count only under a certified set: a set abandoned on the way may hold joints that no solution fixes
    res.status = !ok ? 2 : (certified ? (bad ? 2 : 0) : 1);
    return res;
}

// Same contract as blfik_tick_kernel (K1 -> K2 -> K3 -> K4, `nticks` times per launch); P.L is the row layout built by
// make_layout_eq on the host.  KP: 16 / 24 = the dense rows fit that many lanes (eq_solve_rider), 0 = any K <= 64.
template <int KP>
__global__ void __launch_bounds__(256, 2) blfik_tick_eq_kernel(const TickParams P)
{
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const WM w = carve(smem + (size_t)wib * P.L.total, P.L);
    const int n = P.m.n, nv = P.L.nv;
    for (long long robot = P.first + (long long)blockIdx.x * wpb + wib; robot < P.N; robot += (long long)gridDim.x * wpb) {
        double* bp = P.base_pos + 3 * robot;
        double* bq = P.base_quat + 4 * robot;
        double* s = P.s + (size_t)n * robot;
        if (P.packed && !(P.flags & BLFIK_TICK_KEEP_STATE)) unpack_state(P, robot, bp, bq, s, n, lane);
        bool warm = (P.flags & BLFIK_TICK_WARM_START) != 0;
        QpResult res = {2, 0};
        int iters = 0, worst = 0;
        for (int tick = 0; tick < P.nticks; tick++) {
            k1_fk(P.m, w, bp, bq, s, lane);
            k2_assemble<true>(P, w, robot, s, lane, P.t0 + tick * P.dt);
            res = k3_solve_eq<KP>(P, w, warm, robot, lane);
            if (res.status == 1 && (P.flags & BLFIK_TICK_STRICT)) res.status = 2;
            iters += res.iters;
            worst = res.status > worst ? res.status : worst;
            if (res.status == 2) {   // no velocity to report (the reference's advance() fails): hold still
                for (int i = lane; i < nv; i += 32) w.x[i] = 0.0;
                __syncwarp();
            }
            for (int i = lane; i < nv; i += 32) {
                P.nu[(size_t)nv * robot + i] = w.x[i];
                P.wx[(size_t)nv * robot + i] = w.x[i];
                if (P.nu_host) P.nu_host[(size_t)nv * robot + i] = w.x[i];
                // the active set of this tick seeds the next one, also inside a multi-tick launch (w.y does not survive K1 / K2)
                if (P.L.box && i >= 6) P.wy[(size_t)n * robot + i - 6] = w.y[i];
            }
            __syncwarp();
            if (res.status == 2) break;   // a failed tick ends this robot's rollout: its state stays where the failure happened
            if (P.flags & BLFIK_TICK_INTEGRATE) k4_integrate(n, bp, bq, s, w.x, P.dt, lane);
        }
        for (int i = lane; i < n; i += 32) {   // warm start of the next launch: the joint part of x (wz); the active set is in wy
            P.wz[(size_t)n * robot + i] = w.x[6 + i];
            if (!P.L.box) P.wy[(size_t)n * robot + i] = 0.0;
        }
        if (lane == 0) {
            P.status[robot] = worst;   // the worst tick of the launch, not the last one
            if (P.status_host) P.status_host[robot] = worst;
            P.iters[robot] = iters;
        }
        __syncwarp();
    }
}

}  // namespace blfik
