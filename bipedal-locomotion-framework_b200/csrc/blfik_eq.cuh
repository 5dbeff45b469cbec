// blfik_eq.cuh -- the tick for problems WITHOUT inequality rows (no JointLimitsTask / JointVelocityLimitsTask, no priority-0
// JointTrackingTask) on models too wide for the register-resident kernel: 32 < nv <= 64 or more than 32 dense rows.
// BASELINE.json configs[0] and [1] are of this class (iCubGazeboV2_5 with 32 joints: SE3 feet + CoM as constraints, joint
// regularisation as cost).  One warp per robot, everything in that warp's shared memory.
//
// With no inequality the QP of QPInverseKinematics::advance (QPInverseKinematics.cpp:442-546) is one linear solve:
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

namespace blfik {

__device__ __forceinline__ QpResult k3_solve_eq(const TickParams& P, const WM& w, bool warm, long long robot, int lane)
{
    const Layout& L = P.L;
    const int nv = L.nv, K = L.K, LD = L.LD, LDE = L.LDE;
    const double sigma = P.st.sigma;
    QpResult res = {0, 0};
    // proximal weight per variable: only where the diagonal cost is missing (the base columns) or below sigma; w.Hs becomes it
    for (int i = lane; i < nv; i += 32) {
        const double d0 = w.Hs[i], d = fmax(d0, sigma);
        w.gp[i] = 1.0 / d;
        w.Hs[i] = d - d0;
        w.x[i] = warm ? P.wx[(size_t)nv * robot + i] : 0.0;
    }
    __syncwarp();
    // G, lower triangle, one entry per lane and pass
    const int ntri = K * (K + 1) / 2;
    double gmax = 0.0;
    for (int p = lane; p < ntri; p += 32) {
        int k = (int)((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
        while (k * (k + 1) / 2 > p) k--;
        while ((k + 1) * (k + 2) / 2 <= p) k++;
        const int l = p - k * (k + 1) / 2;
        const double* bk = w.Cs + k * LD;
        const double* bl = w.Cs + l * LD;
        double v = 0.0;
        for (int i = 0; i < nv; i++) v += bk[i] * bl[i] * w.gp[i];
        if (k == l) {
            v += w.bd[k];
            gmax = fmax(gmax, v);
        }
        w.Ss[k * LDE + l] = v;
    }
    gmax = warp_max(gmax);
    __syncwarp();
    bool ok = chol_inplace(w.Ss, K, LDE, 1e-11 * gmax + 1e-300, lane, true);
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
    const int max_steps = 60;
    bool converged = false;
    int it = 0;
    for (; it < max_steps && !converged; it++) {
        for (int i = lane; i < nv; i += 32) w.r[i] = w.gp[i] * (w.Hs[i] * w.x[i] - w.g[i]);   // D^-1 q
        __syncwarp();
        for (int k = lane; k < K; k += 32) {
            const double* bk = w.Cs + k * LD;
            double v = -w.b[k];
            for (int i = 0; i < nv; i++) v += bk[i] * w.r[i];
            w.wb[k] = v;
        }
        __syncwarp();
        for (int k = lane; k < K; k += 32) {
            double v = 0.0;
            for (int j = 0; j <= k; j++) v += w.Ym[k * LDE + j] * w.wb[j];
            w.bp[k] = v;
        }
        __syncwarp();
        for (int k = lane; k < K; k += 32) {
            double v = 0.0;
            for (int j = k; j < K; j++) v += w.Ym[j * LDE + k] * w.bp[j];
            w.lam[k] = v;
        }
        __syncwarp();
        double step = 0.0, xmax = 0.0;
        for (int i = lane; i < nv; i += 32) {
            double bz = 0.0;
            for (int k = 0; k < K; k++) bz += w.Cs[k * LD + i] * w.lam[k];
            const double v = w.r[i] - w.gp[i] * bz;
            step = fmax(step, fabs(v - w.x[i]));
            xmax = fmax(xmax, fabs(v));
            w.x[i] = v;
        }
        step = warp_max(step), xmax = warp_max(xmax);
        __syncwarp();
        if (!(step == step)) break;   // NaN
        converged = step <= 1e-10 * (1.0 + xmax);
    }
    res.iters = it;
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
    res.status = (!ok || bad) ? 2 : (converged ? 0 : 1);
    return res;
}

// Same contract as blfik_tick_kernel (K1 -> K2 -> K3 -> K4, `nticks` times per launch); P.L is the row layout built by
// make_layout_eq on the host.
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
        if (P.packed) unpack_state(P, robot, bp, bq, s, n, lane);
        bool warm = (P.flags & BLFIK_TICK_WARM_START) != 0;
        QpResult res = {2, 0};
        int iters = 0;
        for (int tick = 0; tick < P.nticks; tick++) {
            k1_fk(P.m, w, bp, bq, s, lane);
            k2_assemble<true>(P, w, robot, s, lane, P.t0 + tick * P.dt);
            res = k3_solve_eq(P, w, warm, robot, lane);
            iters += res.iters;
            for (int i = lane; i < nv; i += 32) {
                P.nu[(size_t)nv * robot + i] = w.x[i];
                P.wx[(size_t)nv * robot + i] = w.x[i];
                if (P.nu_host) P.nu_host[(size_t)nv * robot + i] = w.x[i];
            }
            __syncwarp();
            if (P.flags & BLFIK_TICK_INTEGRATE) k4_integrate(n, bp, bq, s, w.x, P.dt, lane);
        }
        for (int i = lane; i < n; i += 32) {   // the inequality part of the warm start does not exist here
            P.wz[(size_t)n * robot + i] = w.x[6 + i];
            P.wy[(size_t)n * robot + i] = 0.0;
        }
        if (lane == 0) {
            P.status[robot] = res.status;
            if (P.status_host) P.status_host[robot] = res.status;
            P.iters[robot] = iters;
        }
        __syncwarp();
    }
}

}  // namespace blfik
