// blfik_capi.cu -- kernels + the C ABI of include/blfik.h (libblfik_cuda.so).
// No torch types, no CPU fallback: every compute entry point needs a CUDA device.
#include "../../include/blfik.h"
#include "blfik_device.cuh"
#include "blfik_fast.cuh"
#include "blfik_eq.cuh"
#include "blfik_split.cuh"
#include "blfik_assemble.cuh"
#include "blfik_inst.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <new>
#include <string>
#include <cstdint>
#include <vector>

using namespace blfik;

// =================================================================== kernels
// The tick: K1 -> K2 -> K3 -> (K4), one warp per robot, `nticks` times per launch.
__global__ void __launch_bounds__(256) blfik_tick_kernel(const TickParams P)
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
        if (warm) {
            for (int i = lane; i < nv; i += 32) {
                w.x[i] = P.wx[(size_t)nv * robot + i];
                w.z[i] = i >= 6 ? P.wz[(size_t)n * robot + i - 6] : 0.0;
                w.y[i] = i >= 6 ? P.wy[(size_t)n * robot + i - 6] : 0.0;
            }
        }
        QpResult res = {2, 0};
        int iters = 0, worst = 0;
        for (int tick = 0; tick < P.nticks; tick++) {
            k1_fk(P.m, w, bp, bq, s, lane);
            k2_assemble(P, w, robot, s, lane, P.t0 + tick * P.dt);
            if (P.dbg.H && tick == 0) {
                const int me = P.L.me, LD = P.L.LD;
                for (int i = lane; i < nv; i += 32) {
                    for (int j = 0; j < nv; j++) P.dbg.H[((size_t)robot * nv + i) * nv + j] = w.Hs[i * LD + j];
                    P.dbg.g[(size_t)robot * nv + i] = w.g[i];
                    for (int r = 0; r < me; r++) P.dbg.Ceq[((size_t)robot * me + r) * nv + i] = w.Cs[r * LD + i];
                    if (i >= 6) {
                        P.dbg.lo[(size_t)robot * n + i - 6] = w.lo[i];
                        P.dbg.hi[(size_t)robot * n + i - 6] = w.hi[i];
                    }
                }
                for (int r = lane; r < me; r += 32) P.dbg.beq[(size_t)robot * me + r] = w.b[r];
            }
            res = k3_solve(P, w, warm, lane);
            if (res.status == 1 && (P.flags & BLFIK_TICK_STRICT)) res.status = 2;
            iters += res.iters;
            worst = res.status > worst ? res.status : worst;
            if (res.status == 2) {   // no velocity to report (the reference's advance() fails): hold still
                for (int i = lane; i < nv; i += 32) w.x[i] = 0.0;
                __syncwarp();
            }
            warm = (P.flags & BLFIK_TICK_WARM_START) != 0;
            for (int i = lane; i < nv; i += 32) {
                P.nu[(size_t)nv * robot + i] = w.x[i];
                if (P.nu_host) P.nu_host[(size_t)nv * robot + i] = w.x[i];
            }
            __syncwarp();
            if (res.status == 2) break;   // a failed tick ends this robot's rollout: its state stays where the failure happened
            if (P.flags & BLFIK_TICK_INTEGRATE) k4_integrate(n, bp, bq, s, w.x, P.dt, lane);
        }
        for (int i = lane; i < nv; i += 32) {
            P.wx[(size_t)nv * robot + i] = w.x[i];
            if (i >= 6) {
                P.wz[(size_t)n * robot + i - 6] = w.z[i];
                P.wy[(size_t)n * robot + i - 6] = w.y[i];
            }
        }
        if (lane == 0) {
            P.status[robot] = worst;   // the worst tick of the launch, not the last one
            if (P.status_host) P.status_host[robot] = worst;
            P.iters[robot] = iters;
        }
        __syncwarp();
    }
}

// K1 stand-alone: frame poses, frame Jacobians, CoM and CoM Jacobian of the current state.
__global__ void __launch_bounds__(256) blfik_kinematics_kernel(const KinParams P)
{
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const WM w = carve(smem + (size_t)wib * P.L.total, P.L);
    const int n = P.m.n, nv = n + 6, LD = P.L.LD;
    for (long long robot = (long long)blockIdx.x * wpb + wib; robot < P.N; robot += (long long)gridDim.x * wpb) {
        k1_fk(P.m, w, P.base_pos + 3 * robot, P.base_quat + 4 * robot, P.s + (size_t)n * robot, lane);
        for (int f = 0; f < P.nframes; f++) {
            double R[9], pf[3];
            const int fr = P.frames[f];
            k1_frame_pose(P.m, w, fr, R, pf);
            const size_t o = (size_t)robot * P.nframes + f;
            if (P.frame_pos && lane < 3) P.frame_pos[o * 3 + lane] = pf[lane];
            if (P.frame_rot && lane < 9) P.frame_rot[o * 9 + lane] = R[lane];
            if (P.frame_jac) {
                k1_frame_jac(P.m, w, P.m.frame_link[fr], pf, w.Jt, LD, lane);
                __syncwarp();
                for (int r = 0; r < 6; r++)
                    for (int c = lane; c < nv; c += 32) P.frame_jac[(o * 6 + r) * nv + c] = w.Jt[r * LD + c];
                __syncwarp();
            }
        }
        if (P.com || P.com_jac) {
            double com[3];
            k1_com(P.m, w, com, P.com_jac ? w.Jt : nullptr, LD, lane);
            __syncwarp();
            if (P.com && lane < 3) P.com[3 * robot + lane] = com[lane];
            if (P.com_jac)
                for (int r = 0; r < 3; r++)
                    for (int c = lane; c < nv; c += 32) P.com_jac[((size_t)robot * 3 + r) * nv + c] = w.Jt[r * LD + c];
        }
        __syncwarp();
    }
}

// K4 stand-alone: nsteps ForwardEuler steps with a held control input.
__global__ void __launch_bounds__(256) blfik_integrate_kernel(int n, long long N, double* base_pos, double* base_quat,
                                                              double* s, const double* nu, double dt, int nsteps)
{
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long robot = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; robot < N; robot += warps)
        for (int k = 0; k < nsteps; k++)
            k4_integrate(n, base_pos + 3 * robot, base_quat + 4 * robot, s + (size_t)n * robot,
                         nu + (size_t)(n + 6) * robot, dt, lane);
}

// K4 at the acceleration level: FloatingBaseSystemAccelerationKinematics (state p, R, s, dp, omega, ds; input [dtwist; dds],
// FloatingBaseSystemAccelerationKinematics.cpp:20-75) under ForwardEuler (ForwardEuler.h:72-124): every component moves by its
// OLD derivative times dt, the rotation by the left-plus Exp(omega dt) R.  One warp per robot, nsteps steps with the input held.
__global__ void __launch_bounds__(256) blfik_integrate_acceleration_kernel(int n, long long N, double* bp, double* bq, double* s, double* dp,
                                                                           double* omega, double* ds, const double* acc, double dt, int nsteps)
{
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long robot = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; robot < N; robot += warps) {
        const double* a = acc + (size_t)(n + 6) * robot;
        for (int k = 0; k < nsteps; k++) {
            const double w[3] = {omega[3 * robot] * dt, omega[3 * robot + 1] * dt, omega[3 * robot + 2] * dt};
            double dq[4], q[4] = {bq[4 * robot], bq[4 * robot + 1], bq[4 * robot + 2], bq[4 * robot + 3]}, qn[4];
            so3_exp(w, dq);
            quat_mul(dq, q, qn);
            const double inv = 1.0 / sqrt(qn[0] * qn[0] + qn[1] * qn[1] + qn[2] * qn[2] + qn[3] * qn[3]);
            __syncwarp();
            if (lane < 3) {
                bp[3 * robot + lane] += dp[3 * robot + lane] * dt;      // position by the old velocity ...
                dp[3 * robot + lane] += a[lane] * dt;                   // ... velocity by the acceleration
                omega[3 * robot + lane] += a[3 + lane] * dt;
            }
            if (lane < 4) bq[4 * robot + lane] = qn[lane] * inv;
            for (int j = lane; j < n; j += 32) {
                s[(size_t)n * robot + j] += ds[(size_t)n * robot + j] * dt;
                ds[(size_t)n * robot + j] += a[6 + j] * dt;
            }
            __syncwarp();
        }
    }
}

// Free-floating mass matrix and generalised gravity forces in the MIXED representation (what iDynTree's
// getFreeFloatingMassMatrix / generalizedGravityForces return to the TSID tasks: src/TSID/src/BaseDynamicsTask.cpp,
// JointDynamicsTask.cpp).  One warp per robot, lane i owns row i (and i + 32):
//   M = sum_k  m_k Jv_k' Jv_k + Jw_k' I_k^W Jw_k      (Jv_k: linear Jacobian of link k's CoM, Jw_k: its angular Jacobian)
//   G = M_total J_com' (0, 0, g)
__global__ void __launch_bounds__(256) blfik_dynamics_kernel(const KinParams P, double* mass_matrix, double* gravity, double g_acc)
{
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const WM w = carve(smem + (size_t)wib * P.L.total, P.L);
    const DevModel& m = P.m;
    const int n = m.n, nv = n + 6, LD = P.L.LD;
    for (long long robot = (long long)blockIdx.x * wpb + wib; robot < P.N; robot += (long long)gridDim.x * wpb) {
        k1_fk(m, w, P.base_pos + 3 * robot, P.base_quat + 4 * robot, P.s + (size_t)n * robot, lane);
        for (int i = lane; i < nv; i += 32)
            for (int j = 0; j < nv; j++) w.Hs[i * LD + j] = 0.0;
        __syncwarp();
        for (int k = 0; k < m.nl; k++) {
            const double mk = m.mass[k];
            if (!(mk > 0.0) && !m.inertia) continue;
            // world CoM of link k and its world inertia
            const double* R = w.linkR + 9 * k;
            double ck[3] = {w.linkP[3 * k], w.linkP[3 * k + 1], w.linkP[3 * k + 2]};
            if (mk > 0.0) ck[0] = w.mcw[3 * k] / mk, ck[1] = w.mcw[3 * k + 1] / mk, ck[2] = w.mcw[3 * k + 2] / mk;
            // columns of the CoM Jacobian of link k, one per lane, into Jt rows 0-2 (linear), 3-5 (angular), 6-8 (I^W angular)
            for (int c = lane; c < nv; c += 32) {
                double lin[3] = {0, 0, 0}, ang[3] = {0, 0, 0};
                if (c < 3) {
                    lin[c] = 1.0;
                } else if (c < 6) {
                    const double r[3] = {ck[0] - w.linkP[0], ck[1] - w.linkP[1], ck[2] - w.linkP[2]};
                    ang[c - 3] = 1.0;
                    cross3(ang, r, lin);
                } else if ((m.anc[k] >> (c - 6)) & 1ull) {
                    const int kj = m.link_of_dof[c - 6];
                    const double d[3] = {ck[0] - w.linkP[3 * kj], ck[1] - w.linkP[3 * kj + 1], ck[2] - w.linkP[3 * kj + 2]};
                    ang[0] = w.axisW[3 * kj], ang[1] = w.axisW[3 * kj + 1], ang[2] = w.axisW[3 * kj + 2];
                    cross3(ang, d, lin);
                }
                double iw[3] = {0, 0, 0};
                if (m.inertia) {
                    const double* I6 = m.inertia + 6 * k;
                    double al[3], il[3];
                    mat3T_vec(R, ang, al);
                    il[0] = I6[0] * al[0] + I6[1] * al[1] + I6[2] * al[2];
                    il[1] = I6[1] * al[0] + I6[3] * al[1] + I6[4] * al[2];
                    il[2] = I6[2] * al[0] + I6[4] * al[1] + I6[5] * al[2];
                    mat3_vec(R, il, iw);
                }
                for (int r3 = 0; r3 < 3; r3++) w.Jt[r3 * LD + c] = lin[r3], w.A[r3 * LD + c] = ang[r3], w.B[r3 * LD + c] = iw[r3];
            }
            __syncwarp();
            for (int i = lane; i < nv; i += 32) {
                const double li[3] = {mk * w.Jt[i], mk * w.Jt[LD + i], mk * w.Jt[2 * LD + i]};
                const double wi[3] = {w.B[i], w.B[LD + i], w.B[2 * LD + i]};
                for (int j = 0; j < nv; j++)
                    w.Hs[i * LD + j] += li[0] * w.Jt[j] + li[1] * w.Jt[LD + j] + li[2] * w.Jt[2 * LD + j] + wi[0] * w.A[j] + wi[1] * w.A[LD + j] + wi[2] * w.A[2 * LD + j];
            }
            __syncwarp();
        }
        if (mass_matrix)
            for (int i = 0; i < nv; i++)
                for (int j = lane; j < nv; j += 32) mass_matrix[((size_t)robot * nv + i) * nv + j] = w.Hs[i * LD + j];
        if (gravity) {
            double com[3];
            k1_com(m, w, com, w.Jt, LD, lane);
            __syncwarp();
            const double M = 1.0 / m.inv_total_mass;
            for (int c = lane; c < nv; c += 32) gravity[(size_t)robot * nv + c] = M * g_acc * w.Jt[2 * LD + c];
        }
        __syncwarp();
    }
}

// Velocity-dependent terms the TSID tasks read: the generalised bias forces h = C(q, nu) nu + g(q) of
// M nu' + h = S'tau + sum J'f (kinDyn->generalizedBiasForces, BaseDynamicsTask.cpp:178, JointDynamicsTask.cpp:233), the bias
// acceleration Jdot nu of frames (kinDyn->getFrameBiasAcc, TSID SE3Task.cpp:254, SO3Task.cpp:189, R3Task.cpp:219) and of the CoM
// (getCenterOfMassBiasAcc, TSID CoMTask.cpp:195), all in the MIXED representation, where nu' = 0 means that the base origin and
// the base angular velocity (both in world coordinates) stay constant.  One warp per robot.  Forward sweep by tree level
// (link by lane): w_k = w_p + a_k sd_k, w'_k = w'_p + (w_p x a_k) sd_k, origin velocity and acceleration carried by the
// parent; per link the force m (a_c + g z) and the moment I w' + w x I w about its CoM that this motion needs; column c of h is
// the projection of those on the motion of variable c (all links for the base, the distal ones for a joint).
__global__ void __launch_bounds__(256) blfik_bias_kernel(const KinParams P, const double* __restrict__ nu, double g_acc, double* __restrict__ bias,
                                                         double* __restrict__ frame_bias, double* __restrict__ frame_vel, double* __restrict__ com_bias)
{
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const WM w = carve(smem + (size_t)wib * P.L.total, P.L);
    const DevModel& m = P.m;
    const int n = m.n, nv = n + 6, nl = m.nl;
    double *omW = w.A, *vO = w.A + 3 * nl, *dOm = w.A + 6 * nl, *aO = w.A + 9 * nl, *Fk = w.A + 12 * nl, *Nk = w.A + 15 * nl;
    for (long long robot = (long long)blockIdx.x * wpb + wib; robot < P.N; robot += (long long)gridDim.x * wpb) {
        k1_fk(m, w, P.base_pos + 3 * robot, P.base_quat + 4 * robot, P.s + (size_t)n * robot, lane);
        const double* v = nu + (size_t)nv * robot;
        if (lane < 3) omW[lane] = v[3 + lane], vO[lane] = v[lane], dOm[lane] = 0.0, aO[lane] = 0.0;
        __syncwarp();
        for (int d = 1; d <= m.maxdepth; d++) {
            for (int k = lane; k < nl; k += 32) {
                if (m.depth[k] != d) continue;
                const int par = m.parent[k];
                const double sd = v[6 + m.dof[k]];
                const double a[3] = {w.axisW[3 * k], w.axisW[3 * k + 1], w.axisW[3 * k + 2]};
                const double wp[3] = {omW[3 * par], omW[3 * par + 1], omW[3 * par + 2]};
                const double dwp[3] = {dOm[3 * par], dOm[3 * par + 1], dOm[3 * par + 2]};
                const double r[3] = {w.linkP[3 * k] - w.linkP[3 * par], w.linkP[3 * k + 1] - w.linkP[3 * par + 1], w.linkP[3 * k + 2] - w.linkP[3 * par + 2]};
                double wxa[3], wxr[3], dwxr[3], wwr[3];
                cross3(wp, a, wxa), cross3(wp, r, wxr), cross3(dwp, r, dwxr), cross3(wp, wxr, wwr);
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    omW[3 * k + c] = wp[c] + a[c] * sd;
                    dOm[3 * k + c] = dwp[c] + wxa[c] * sd;
                    vO[3 * k + c] = vO[3 * par + c] + wxr[c];
                    aO[3 * k + c] = aO[3 * par + c] + dwxr[c] + wwr[c];
                }
            }
            __syncwarp();
        }
        double macc[3] = {0.0, 0.0, 0.0};
        for (int k = lane; k < nl; k += 32) {
            const double mk = m.mass[k];
            const double wk[3] = {omW[3 * k], omW[3 * k + 1], omW[3 * k + 2]}, dwk[3] = {dOm[3 * k], dOm[3 * k + 1], dOm[3 * k + 2]};
            double f[3] = {0.0, 0.0, 0.0}, mo[3] = {0.0, 0.0, 0.0};
            if (mk > 0.0) {
                const double rc[3] = {w.mcw[3 * k] / mk - w.linkP[3 * k], w.mcw[3 * k + 1] / mk - w.linkP[3 * k + 1], w.mcw[3 * k + 2] / mk - w.linkP[3 * k + 2]};
                double dwxr[3], wxr[3], wwr[3];
                cross3(dwk, rc, dwxr), cross3(wk, rc, wxr), cross3(wk, wxr, wwr);
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    const double ac = aO[3 * k + c] + dwxr[c] + wwr[c];
                    macc[c] += mk * ac;
                    f[c] = mk * (ac + (c == 2 ? g_acc : 0.0));
                }
            }
            if (m.inertia) {
                const double* R = w.linkR + 9 * k;
                const double* I6 = m.inertia + 6 * k;
                auto apply = [&](const double* x, double* y) {   // I^W x = R I R' x
                    double xl[3], il[3];
                    mat3T_vec(R, x, xl);
                    il[0] = I6[0] * xl[0] + I6[1] * xl[1] + I6[2] * xl[2];
                    il[1] = I6[1] * xl[0] + I6[3] * xl[1] + I6[4] * xl[2];
                    il[2] = I6[2] * xl[0] + I6[4] * xl[1] + I6[5] * xl[2];
                    mat3_vec(R, il, y);
                };
                double Idw[3], Iw[3], wxIw[3];
                apply(dwk, Idw), apply(wk, Iw);
                cross3(wk, Iw, wxIw);
#pragma unroll
                for (int c = 0; c < 3; c++) mo[c] = Idw[c] + wxIw[c];
            }
#pragma unroll
            for (int c = 0; c < 3; c++) Fk[3 * k + c] = f[c], Nk[3 * k + c] = mo[c];
        }
        if (com_bias) {
#pragma unroll
            for (int c = 0; c < 3; c++) macc[c] = warp_sum(macc[c]) * m.inv_total_mass;
            if (lane < 3) com_bias[3 * robot + lane] = lane == 0 ? macc[0] : (lane == 1 ? macc[1] : macc[2]);
        }
        __syncwarp();
        if (bias) {
            for (int c = lane; c < nv; c += 32) {
                double h = 0.0;
                if (c < 3) {
                    for (int k = 0; k < nl; k++) h += Fk[3 * k + c];
                } else {
                    const int j = c - 6;
                    const int kj = c < 6 ? 0 : m.link_of_dof[j];
                    double a[3] = {0.0, 0.0, 0.0};
                    if (c < 6) a[c - 3] = 1.0;
                    else a[0] = w.axisW[3 * kj], a[1] = w.axisW[3 * kj + 1], a[2] = w.axisW[3 * kj + 2];
                    const double o[3] = {w.linkP[3 * kj], w.linkP[3 * kj + 1], w.linkP[3 * kj + 2]};
                    double t[3] = {0.0, 0.0, 0.0};
                    for (int k = 0; k < nl; k++) {
                        if (c >= 6 && !((m.anc[k] >> j) & 1ull)) continue;
                        const double mk = m.mass[k];
                        if (mk > 0.0) {
                            const double r[3] = {w.mcw[3 * k] / mk - o[0], w.mcw[3 * k + 1] / mk - o[1], w.mcw[3 * k + 2] / mk - o[2]};
                            const double f[3] = {Fk[3 * k], Fk[3 * k + 1], Fk[3 * k + 2]};
                            double rxf[3];
                            cross3(r, f, rxf);
                            t[0] += rxf[0], t[1] += rxf[1], t[2] += rxf[2];
                        }
                        t[0] += Nk[3 * k], t[1] += Nk[3 * k + 1], t[2] += Nk[3 * k + 2];
                    }
                    h = a[0] * t[0] + a[1] * t[1] + a[2] * t[2];
                }
                bias[(size_t)robot * nv + c] = h;
            }
        }
        if (frame_bias || frame_vel) {
            for (int f = lane; f < P.nframes; f += 32) {
                const int fr = P.frames[f], L = m.frame_link[fr];
                double off[3];
                const double fp[3] = {m.Fp[3 * fr], m.Fp[3 * fr + 1], m.Fp[3 * fr + 2]};
                mat3_vec(w.linkR + 9 * L, fp, off);
                const double wk[3] = {omW[3 * L], omW[3 * L + 1], omW[3 * L + 2]}, dwk[3] = {dOm[3 * L], dOm[3 * L + 1], dOm[3 * L + 2]};
                double dwxr[3], wxr[3], wwr[3];
                cross3(dwk, off, dwxr), cross3(wk, off, wxr), cross3(wk, wxr, wwr);
                if (frame_bias) {
                    double* out = frame_bias + ((size_t)robot * P.nframes + f) * 6;
#pragma unroll
                    for (int c = 0; c < 3; c++) out[c] = aO[3 * L + c] + dwxr[c] + wwr[c], out[3 + c] = dwk[c];
                }
                if (frame_vel) {   // J nu: velocity of the frame origin and angular velocity of its link
                    double* out = frame_vel + ((size_t)robot * P.nframes + f) * 6;
#pragma unroll
                    for (int c = 0; c < 3; c++) out[c] = vO[3 * L + c] + wxr[c], out[3 + c] = wk[c];
                }
            }
        }
        __syncwarp();
    }
}

// Coefficients of the quintic splines of one task (Math::QuinticSpline, QuinticSpline.h:266-415), one thread per robot and
// coordinate.  The interior-knot system is block tridiagonal with 2 x 2 blocks that depend on the knot times only, so the
// host eliminates the blocks once (fac[i] = L_i, M_i = L_i D'_{i-1}^-1, D'_i^-1, U_i) and a thread runs the two sweeps on
// its own right-hand side.  The sweeps keep their intermediates in the coefficient slots the thread owns:
// coef[robot][interval][k][coordinate], interval j starting at knot j.
__global__ void __launch_bounds__(128) blfik_spline_coefficients_kernel(long long N, int nk, const double* __restrict__ times,
                                                                        const double* __restrict__ fac, const double* __restrict__ x,
                                                                        const double* __restrict__ bc, double* __restrict__ coef, int dim)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * dim) return;
    const long long robot = idx / dim;
    const int c = (int)(idx - robot * dim), S = 6 * dim;
    const double* xr = x + (size_t)robot * nk * dim + c;
    double* cf = coef + (size_t)robot * (nk - 1) * S + c;
    double v0 = 0.0, a0 = 0.0, vf = 0.0, af = 0.0;
    if (bc) {
        const double* b = bc + (size_t)robot * 4 * dim + c;
        v0 = b[0], a0 = b[dim], vf = b[2 * dim], af = b[3 * dim];
    }
    const int m = nk - 2;
    double p0 = 0.0, p1 = 0.0;
    for (int i = 0; i < m; i++) {   // forward sweep: r'_i = r_i - M_i r'_{i-1}
        const double Tp = times[i + 1] - times[i], Tn = times[i + 2] - times[i + 1];
        const double dp = xr[dim * (i + 1)] - xr[dim * i], dn = xr[dim * (i + 2)] - xr[dim * (i + 1)];
        const double ip = 1.0 / Tp, in = 1.0 / Tn;
        double r0 = 60.0 * (dn * in * in * in - dp * ip * ip * ip);
        double r1 = -360.0 * (dp * ip * ip * ip * ip + dn * in * in * in * in);
        const double* f = fac + 16 * i;
        if (i == 0) r0 -= f[0] * v0 + f[1] * a0, r1 -= f[2] * v0 + f[3] * a0;
        if (i == m - 1) r0 -= f[12] * vf + f[13] * af, r1 -= f[14] * vf + f[15] * af;
        if (i > 0) {
            const double t0 = r0 - (f[4] * p0 + f[5] * p1), t1 = r1 - (f[6] * p0 + f[7] * p1);
            r0 = t0, r1 = t1;
        }
        cf[S * (i + 1) + 3 * dim] = r0, cf[S * (i + 1) + 4 * dim] = r1;
        p0 = r0, p1 = r1;
    }
    double vn = vf, an = af;
    for (int i = m - 1; i >= 0; i--) {   // back substitution: (v, a)_i = D'_i^-1 (r'_i - U_i (v, a)_{i+1})
        const double* f = fac + 16 * i;
        double r0 = cf[S * (i + 1) + 3 * dim], r1 = cf[S * (i + 1) + 4 * dim];
        if (i < m - 1) r0 -= f[12] * vn + f[13] * an, r1 -= f[14] * vn + f[15] * an;
        vn = f[8] * r0 + f[9] * r1, an = f[10] * r0 + f[11] * r1;
        cf[S * (i + 1) + dim] = vn, cf[S * (i + 1) + 2 * dim] = 0.5 * an;
    }
    cf[dim] = v0, cf[2 * dim] = 0.5 * a0;
    for (int j = 0; j < nk - 1; j++) {   // the interval polynomials from their two end states
        const double T = times[j + 1] - times[j], iT = 1.0 / T;
        const double x0 = xr[dim * j], d = xr[dim * (j + 1)] - x0;
        const double vs = cf[S * j + dim], as = 2.0 * cf[S * j + 2 * dim];
        const double ve = j + 1 <= nk - 2 ? cf[S * (j + 1) + dim] : vf, ae = j + 1 <= nk - 2 ? 2.0 * cf[S * (j + 1) + 2 * dim] : af;
        const double iT3 = iT * iT * iT;
        cf[S * j] = x0;
        cf[S * j + 3 * dim] = 0.5 * iT3 * (20.0 * d - (8.0 * ve + 12.0 * vs) * T - (3.0 * as - ae) * T * T);
        cf[S * j + 4 * dim] = 0.5 * iT3 * iT * (-30.0 * d + (14.0 * ve + 16.0 * vs) * T + (3.0 * as - 2.0 * ae) * T * T);
        cf[S * j + 5 * dim] = 0.5 * iT3 * iT * iT * (12.0 * d - 6.0 * (ve + vs) * T - (as - ae) * T * T);
    }
}

// The same for Math::CubicSpline (CubicSpline.h:77-330): the knot velocities make the acceleration continuous, a scalar
// tridiagonal system per robot and coordinate; fac[i] = l_i, m_i = l_i / d'_{i-1}, 1 / d'_i, u_i at offsets 0, 4, 8, 12.  The
// cubic of every interval is stored as a quintic with c_4 = c_5 = 0, so the tick evaluates both kinds with the same code.
// bc: [N][2][3], velocity at the first and at the last knot.
__global__ void __launch_bounds__(128) blfik_cubic_spline_coefficients_kernel(long long N, int nk, const double* __restrict__ times,
                                                                              const double* __restrict__ fac, const double* __restrict__ x,
                                                                              const double* __restrict__ bc, double* __restrict__ coef, int dim)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * dim) return;
    const long long robot = idx / dim;
    const int c = (int)(idx - robot * dim), S = 6 * dim;
    const double* xr = x + (size_t)robot * nk * dim + c;
    double* cf = coef + (size_t)robot * (nk - 1) * S + c;
    double v0 = 0.0, vf = 0.0;
    if (bc) v0 = bc[(size_t)robot * 2 * dim + c], vf = bc[(size_t)robot * 2 * dim + dim + c];
    const int m = nk - 2;
    double prev = 0.0;
    for (int i = 0; i < m; i++) {
        const double ip = 1.0 / (times[i + 1] - times[i]), in = 1.0 / (times[i + 2] - times[i + 1]);
        const double dp = xr[dim * (i + 1)] - xr[dim * i], dn = xr[dim * (i + 2)] - xr[dim * (i + 1)];
        const double* f = fac + 16 * i;
        double r = 6.0 * (dp * ip * ip + dn * in * in);
        if (i == 0) r -= f[0] * v0;
        if (i == m - 1) r -= f[12] * vf;
        if (i > 0) r -= f[4] * prev;
        cf[S * (i + 1) + 3 * dim] = r;
        prev = r;
    }
    double vn = vf;
    for (int i = m - 1; i >= 0; i--) {
        const double* f = fac + 16 * i;
        double r = cf[S * (i + 1) + 3 * dim];
        if (i < m - 1) r -= f[12] * vn;
        vn = f[8] * r;
        cf[S * (i + 1) + dim] = vn;
    }
    cf[dim] = v0;
    for (int j = 0; j < nk - 1; j++) {
        const double T = times[j + 1] - times[j], iT = 1.0 / T;
        const double x0 = xr[dim * j], d = xr[dim * (j + 1)] - x0;
        const double vs = cf[S * j + dim], ve = j + 1 <= nk - 2 ? cf[S * (j + 1) + dim] : vf;
        cf[S * j] = x0;
        cf[S * j + 2 * dim] = (3.0 * d - (2.0 * vs + ve) * T) * iT * iT;
        cf[S * j + 3 * dim] = (-2.0 * d + (vs + ve) * T) * iT * iT * iT;
        cf[S * j + 4 * dim] = 0.0, cf[S * j + 5 * dim] = 0.0;
    }
}

// The set-point a tick at time t reads: the stored record with the attached generators applied, one thread per robot.
__global__ void __launch_bounds__(128) blfik_evaluate_setpoint_kernel(const DevTask* tasks, int t, long long N, double time, double* out)
{
    const long long robot = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (robot >= N) return;
    const DevTask& tk = tasks[t];
    const double* sp = tk.setpoint + (size_t)robot * tk.sp_stride;
    if (tk.type == T_JOINT_TRACKING) {   // [position (n); velocity (n)]
        const int n = tk.fb_off / 2;
        for (int j = 0; j < n; j++) joint_setpoint(tk, robot, time, sp, false, n, j, out[(size_t)robot * tk.fb_off + j], out[(size_t)robot * tk.fb_off + n + j]);
        return;
    }
    double spl[13];
    const double* use = traj_eval(tk, robot, time, sp, spl);
    for (int c = 0; c < tk.fb_off; c++) out[(size_t)robot * tk.fb_off + c] = use[c];
}

__global__ void blfik_count_unsolved_kernel(const int* status, long long N, unsigned long long* out)
{
    unsigned long long c = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (long long)gridDim.x * blockDim.x)
        c += status[i] != 0;
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(kFull, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// FP64 roofline probe: 8 independent FMA chains per thread, no memory traffic.
__global__ void __launch_bounds__(512) blfik_dfma_probe_kernel(double* sink, int iters, double a, double b)
{
    double c0 = threadIdx.x, c1 = c0 + 1, c2 = c0 + 2, c3 = c0 + 3, c4 = c0 + 4, c5 = c0 + 5, c6 = c0 + 6, c7 = c0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            c0 = fma(c0, a, b), c1 = fma(c1, a, b), c2 = fma(c2, a, b), c3 = fma(c3, a, b);
            c4 = fma(c4, a, b), c5 = fma(c5, a, b), c6 = fma(c6, a, b), c7 = fma(c7, a, b);
        }
    }
    const double r = ((c0 + c1) + (c2 + c3)) + ((c4 + c5) + (c6 + c7));
    if (r == 123.456) sink[0] = r;
}

// =================================================================== host side
static thread_local std::string g_err;

static int fail(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) return fail(BLFIK_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct blfik_model {
    int n = 0, nf = 0, maxdepth = 0;
    std::vector<int> parent, dof, depth, link_of_dof, frame_link;
    std::vector<unsigned long long> anc;
    std::vector<double> Xr, Xp, axis, mass, com, msub, Fr, Fp, inertia;   // inertia: [nl][6], empty = point masses
    double total_mass = 0;
};

struct HostTask {
    blfik_task_desc d;
    std::vector<double> kp_joint, weight, klim, upper, lower;
    int rows = 0, sp_len = 0, row_offset = 0, brow = 0, pk_off = 0;   // sp_len: doubles per robot of the record (set-point + feedback)
    int fb_len = 0;   // exogenous feedback behind the set-point: 7 (SE3: p, q), 3 (CoM: p) or 0
    int base_c = 0, pk_off_c = 0;   // compact packed layout: doubles of the pose part, offset inside the record
    double traj_amp[6] = {0, 0, 0, 0, 0, 0}, traj_omega = 0.0;
    double* d_traj_phase = nullptr;
    double* d_mj_target = nullptr;   // minimum-jerk generator: target set-points [N][set-point length]
    double mj_t0 = 0.0, mj_T = 0.0;  // mj_T == 0: off
    int sp_nk = 0, sp_cap = 0;       // quintic-spline generator: knots in force (0: off), knots the buffers hold
    double *d_sp_times = nullptr, *d_sp_fac = nullptr, *d_sp_coef = nullptr;
    bool setpoint_set = false;
    std::vector<double> sched_target;  // weight schedule (blfik_set_weight_schedule): first-order approach of `weight` to this vector ...
    double sched_a = 0.0;              // ... at rate a = 3 / settling_time per second; 0: no schedule
    bool shared = false;             // batch-shared set-point: one record for all robots, not part of the packed records
    double* d_setpoint = nullptr;
    double *d_kp_joint = nullptr, *d_weight = nullptr, *d_klim = nullptr, *d_upper = nullptr, *d_lower = nullptr;
};

constexpr int kMaxChunks = 16;
constexpr int kCounterSlots = kMaxChunks;
constexpr int kMaxPieces = 64;      // pieces of the streaming-input mode   // work counters of the register-resident kernel, one per launch that may be in flight

struct blfik_problem {
    const blfik_model* model = nullptr;
    int device = 0;
    long long N = 0;
    int n = 0, nv = 0, me = 0;
    std::vector<HostTask> tasks;
    Settings st;
    Layout L;
    DevModel dm;
    std::vector<void*> allocs;
    DevTask* d_tasks = nullptr;
    bool tasks_dirty = true;
    double *base_pos = nullptr, *base_quat = nullptr, *s = nullptr, *nu = nullptr, *wx = nullptr, *wz = nullptr, *wy = nullptr;
    int *status = nullptr, *iters = nullptr;
    int* work_counter = nullptr;   // dynamic hand-out of robots in the register-resident kernel
    // two-launch path, longest first (SolveParams::slow_*): two lists and two stamp arrays used alternately, their lengths, the
    // stamp of the next solve launch and the stamp for which the list written last is valid (-1: none)
    int* d_slow = nullptr;         // [2][N] lists, [2][N] stamps, [2] lengths, [N] consecutive unsolved ticks
    int* d_bins = nullptr;         // [64] robots per hardness bin, [64][N] their indices (hand-out order of the solve kernel)
    int* d_bins_chunk = nullptr;   // the same for the chunk launches of the host tick: [kCounterSlots][64] counts, then [64 N] indices (a chunk's lists at 64 x first)
    int solve_stamp = 0;
    int slow_list_for = -1;
    bool streaming_launch = false; // set around the one launch of the streaming-input mode
    bool streaming_off = false;    // the mode timed out once (serialised execution): chunk launches from then on
    int* d_arrived = nullptr;      // streaming-input mode of blfik_tick_packed: robots whose record is on the device; [1] = abort word
    int* h_stream_ctl = nullptr;   // pinned + mapped: [0..kMaxPieces) cumulative robot counts (sources of the flag copies), [kMaxPieces] = abort
    unsigned long long* d_count = nullptr;
    bool state_set = false, ticked = false;
    DebugOut dbg = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool debug = false;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int warps_per_block = 4, grid = 148;
    size_t smem_bytes = 0;
    // equality-only kernel (blfik_eq.cuh): problems without inequality rows that the register-resident kernel cannot hold
    bool eq_kernel = false;
    void (*eq_func)(const TickParams) = nullptr;   // blfik_tick_eq_kernel<KP> for this problem's row count
    Layout Leq;
    int eq_wpb = 8, eq_grid = 148;
    size_t eq_smem = 0;
    int num_sms = 148;
    long long launches = 0;
    // register-resident kernel (blfik_fast.cuh), used when nv <= 32 and K <= 32
    int K = 0;                       // dense task rows (cost + equality)
    int path = 0;                    // 0 auto, 1 force the generic shared-memory kernel
    void (*fast_kernel)(const FastParams) = nullptr;
    FastLayout F;
    size_t fast_smem = 0;
    int fast_grid = 0, fast_wpb = 12;
    int max_as_iters = 12, max_polish_iters = 40, max_total_iters = 200;
    double* d_model_pack = nullptr;
    int njump = 0, has_com = 0;
    // two-launch path (blfik_split.cuh): assemble (one thread per robot) -> QP records -> solve (one warp per robot)
    void (*solve_kernel)(const SolveParams) = nullptr;
    int rec_len = 0, rec_off_c = 0, rec_off_v = 0, rec_kp = 0;
    void (*asm_kernel)(const AsmParams) = nullptr;
    int solve_wpb = 1, solve_blocks = 16;
    size_t solve_smem = 0, asm_smem = 0;
    int pack_len = 0, asm_threads = 128, asm_D = 0, asm_nevents = 0;
    int* d_asm_events = nullptr;
    double* d_qp = nullptr;          // [N][rec_len], allocated by the first tick that takes this path
    bool timing = false;             // blfik_kernel_timing: events around the two kernels of the last full-batch tick
    cudaEvent_t ev_t[3] = {nullptr, nullptr, nullptr};
    // pipelined host-buffer tick (blfik_tick_host)
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
    cudaStream_t asm_stream[4] = {nullptr, nullptr, nullptr, nullptr};     // blfik_tick_packed on the two-launch path: high-priority streams of the assemble kernels
    cudaEvent_t ev_asm[16] = {};
    cudaStream_t chunk_stream[4] = {nullptr, nullptr, nullptr, nullptr};   // blfik_tick_packed: consecutive chunk kernels alternate, so one fills the SMs the other's tail leaves idle
    cudaEvent_t ev_in[16] = {}, ev_done[16] = {}, ev_start = nullptr;
    double time = 0.0;               // time of the next tick (advanced by integrating ticks; set-point trajectories)
    int packed_len = 0;              // doubles per packed input record: 7 + n + sum of set-point lengths
    int packed_len_c = 0;            // the same in the compact layout (BLFIK_TICK_NO_FEEDFORWARD)
    int packed_len_max = 0;          // the longest a record can be (no batch-shared set-point): what d_packed is sized for
    double* d_packed = nullptr;      // device copy of the packed records (allocated on first use)
};

// Block-shared constants of the register-resident kernel: per-link records [nl][20] = Xr(9) Xp(3) axis(3) com(3) mass pad,
// pointer-jumping tables [njump][32] (int), link_of_dof[32] + dof_of_link[32] (int), ancestor masks [nl] (u64), subtree masses [n].
static std::vector<double> pack_model(const blfik_model& m, FastLayout& F, int& njump)
{
    const int n = m.n, nl = n + 1;
    njump = 0;
    while ((1 << njump) < m.maxdepth + 1) njump++;
    std::vector<int> jump((size_t)(njump > 0 ? njump : 1) * 32, -1);
    {
        std::vector<int> cur(32, -1);
        for (int k = 0; k < nl; k++) cur[k] = m.parent[k];
        for (int st = 0; st < njump; st++) {
            for (int k = 0; k < 32; k++) jump[st * 32 + k] = cur[k];
            std::vector<int> nxt(32, -1);
            for (int k = 0; k < 32; k++) nxt[k] = cur[k] >= 0 ? cur[cur[k]] : -1;
            cur = nxt;
        }
    }
    int o = 20 * nl;
    o = (o + 1) & ~1;
    F.sh_jump = o, o += (njump > 0 ? njump : 1) * 16;
    F.sh_lod = o, o += 32;
    F.sh_anc = o, o += nl;
    F.sh_msub = o, o += n;
    o = (o + 1) & ~1;
    F.sh_model_len = o;
    std::vector<double> pack(o, 0.0);
    for (int k = 0; k < nl; k++) {
        double* r = pack.data() + 20 * k;
        for (int c = 0; c < 9; c++) r[c] = m.Xr[9 * k + c];
        for (int c = 0; c < 3; c++) r[9 + c] = m.Xp[3 * k + c], r[12 + c] = m.axis[3 * k + c], r[15 + c] = m.com[3 * k + c];
        r[18] = m.mass[k];
    }
    int* ji = reinterpret_cast<int*>(pack.data() + F.sh_jump);
    for (size_t i = 0; i < jump.size(); i++) ji[i] = jump[i];
    int* li = reinterpret_cast<int*>(pack.data() + F.sh_lod);
    for (int j = 0; j < 32; j++) li[j] = j < n ? m.link_of_dof[j] : 0;
    for (int k = 0; k < 32; k++) li[32 + k] = (k >= 1 && k < nl) ? m.dof[k] : 0;
    unsigned long long* an = reinterpret_cast<unsigned long long*>(pack.data() + F.sh_anc);
    for (int k = 0; k < nl; k++) an[k] = m.anc[k];
    for (int j = 0; j < n; j++) pack[F.sh_msub + j] = m.msub[j];
    return pack;
}

template <int NV, int KP>
static FastLayout make_fast_layout(int nl)
{
    constexpr int LDB = FastDims<NV, KP>::LDB, LDG = FastDims<NV, KP>::LDG;
    FastLayout F;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 1) & ~1; return r; };
    F.linkR = take(9 * nl), F.linkP = take(3 * nl), F.axisW = take(3 * nl), F.mcw = take(3 * nl);
    // the link arrays are dead after K2, L_G' is born in the setup: same memory
    // (+ KP: the rotating elimination reads up to KP slots past a column)
    F.LGt = F.linkR;
    if (o < KP * LDG + KP) o = (KP * LDG + KP + 1) & ~1;
    F.Bm = take(KP * LDB);
    // link arrays, LGt and Bm are dead once U is known: the N rows (NV x LDB) live on top of them
    if (o < NV * LDB) o = (NV * LDB + 1) & ~1;
    // G (by columns), then U rows; dead after N is formed: the active-set matrix (|A| x (|A||1)) lives there
    const int n = nl - 1;
    int r2 = (NV > KP ? NV : KP) * LDG;
    if (r2 < n * (n | 1)) r2 = n * (n | 1);
    F.Us = take(r2);
    F.cvec = take(KP), F.evec = take(KP), F.dinv = take(LDB), F.rvec = take(LDB), F.wv = take(LDG);
    F.dA = take(32), F.muA = take(32), F.idxA = take(16);
    F.Ns = F.linkR, F.MA = F.Us;
    F.ma_cap = n < 32 ? n : 32;
    F.total = o;
    return F;
}

template <int NV, int KP, bool AM = false>
static void pick_fast(blfik_problem* p)
{
    p->fast_kernel = blfik_tick_fast_kernel<NV, KP, AM>;
    p->fast_wpb = FastDims<NV, KP>::WARPS;   // the kernel's launch bound
    p->solve_kernel = blfik_solve_kernel<NV, KP>;
    p->rec_len = QpRec<NV, KP>::LEN, p->rec_off_c = QpRec<NV, KP>::cvec, p->rec_off_v = QpRec<NV, KP>::vec;
    p->rec_kp = KP;
    p->asm_kernel = blfik_assemble_kernel<KP>;
    p->solve_wpb = SolveDims<NV, KP>::WPB, p->solve_blocks = SolveDims<NV, KP>::BLOCKS;
    p->solve_smem = (size_t)SolveDims<NV, KP>::WPB * SolveLayout<NV, KP>::TOTAL * sizeof(double);
    p->F = make_fast_layout<NV, KP>(p->n + 1);
    // the control-law buffer ([ntasks][12]) lives in the U region during K2
    // (it always fits: ntasks * 12 <= 32 * 12 <= NV * LDG for every instantiation with NV >= 29, KP >= 16)
}

static int base_compact_of(int type, int n);
// Offsets of the tasks inside a packed input record, full and compact layout.  Tasks without a set-point and tasks whose
// set-point is batch-shared take no room (offset -1).
static void layout_packed(blfik_problem* p)
{
    const int n = p->n;
    p->packed_len = p->packed_len_c = 7 + n;
    for (HostTask& h : p->tasks) {
        h.base_c = base_compact_of(h.d.type, n);
        if (h.sp_len == 0 || h.shared) {
            h.pk_off = h.pk_off_c = -1;
            continue;
        }
        h.pk_off = p->packed_len;
        p->packed_len += h.sp_len;
        h.pk_off_c = p->packed_len_c;
        p->packed_len_c += h.base_c + h.fb_len;
    }
}

static int rows_of(const blfik_task_desc& d, int n)
{
    const int nm = (d.mask[0] != 0) + (d.mask[1] != 0) + (d.mask[2] != 0);
    switch (d.type) {
    case BLFIK_TASK_SE3: return nm + 3;
    case BLFIK_TASK_COM: return nm;
    case BLFIK_TASK_R3: return nm;
    case BLFIK_TASK_SO3: return 3;
    case BLFIK_TASK_DISTANCE: return 1;
    case BLFIK_TASK_GRAVITY: return 2;
    case BLFIK_TASK_ANGULAR_MOMENTUM: return nm;
    case BLFIK_TASK_JOINT_TRACKING: return n;
    case BLFIK_TASK_JOINT_LIMITS: return 2 * n;
    case BLFIK_TASK_JOINT_VELOCITY_LIMITS: return 2 * n;
    default: return -1;
    }
}
static int sp_len_of(int type, int n)
{
    switch (type) {
    case BLFIK_TASK_SE3: return 13;
    case BLFIK_TASK_COM: return 6;
    case BLFIK_TASK_R3: return 6;
    case BLFIK_TASK_SO3: return 7;
    case BLFIK_TASK_DISTANCE: return 1;
    case BLFIK_TASK_GRAVITY: return 6;
    case BLFIK_TASK_ANGULAR_MOMENTUM: return 3;
    case BLFIK_TASK_JOINT_TRACKING: return 2 * n;
    default: return 0;
    }
}
// pose part of a set-point (what is left when the feed-forward terms are dropped)
static int base_compact_of(int type, int n)
{
    switch (type) {
    case BLFIK_TASK_SE3: return 7;
    case BLFIK_TASK_COM: return 3;
    case BLFIK_TASK_R3: return 3;
    case BLFIK_TASK_SO3: return 4;
    case BLFIK_TASK_DISTANCE: return 1;
    case BLFIK_TASK_GRAVITY: return 3;
    case BLFIK_TASK_ANGULAR_MOMENTUM: return 3;
    case BLFIK_TASK_JOINT_TRACKING: return n;
    default: return 0;
    }
}
static bool is_dense(int type) { return is_dense_type(type); }
static bool has_frame(int type) { return type == BLFIK_TASK_SE3 || type == BLFIK_TASK_SO3 || type == BLFIK_TASK_R3 || type == BLFIK_TASK_DISTANCE || type == BLFIK_TASK_GRAVITY; }
static bool is_inequality(int type) { return type == BLFIK_TASK_JOINT_LIMITS || type == BLFIK_TASK_JOINT_VELOCITY_LIMITS; }

extern "C" const char* blfik_last_error(void) { return g_err.c_str(); }

extern "C" void blfik_default_settings(blfik_solver_settings* s)
{
    if (!s) return;
    s->rho = 1.0, s->sigma = 1e-2, s->alpha = 1.6, s->eps = 1e-7;
    s->admm_iters = 10, s->max_rounds = 8, s->polish_iters = 4, s->eq_iters = 5;
}

// ------------------------------------------------------------------- model
extern "C" int blfik_model_create(const blfik_model_desc* d, blfik_model** out)
{
    if (!d || !out) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] null argument");
    const int n = d->n_dofs, nl = n + 1, nf = d->n_frames;
    if (n < 1 || n > 58) return fail(BLFIK_ERR_UNSUPPORTED, "[blfik_model_create] n_dofs must be in [1, 58], got %d", n);
    if (nf < 0) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] negative n_frames");
    if (!d->parent || !d->dof || !d->X_rot || !d->X_pos || !d->axis || !d->mass || !d->com || (nf && (!d->frame_link || !d->frame_rot || !d->frame_pos)))
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] null array in the model description");
    blfik_model* m = new (std::nothrow) blfik_model;
    if (!m) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] out of host memory");
    m->n = n, m->nf = nf;
    m->parent.assign(d->parent, d->parent + nl), m->dof.assign(d->dof, d->dof + nl);
    m->Xr.assign(d->X_rot, d->X_rot + 9 * nl), m->Xp.assign(d->X_pos, d->X_pos + 3 * nl);
    m->axis.assign(d->axis, d->axis + 3 * nl), m->mass.assign(d->mass, d->mass + nl), m->com.assign(d->com, d->com + 3 * nl);
    m->frame_link.assign(d->frame_link, d->frame_link + nf);
    if (d->inertia) m->inertia.assign(d->inertia, d->inertia + 6 * nl);
    m->Fr.assign(d->frame_rot, d->frame_rot + 9 * nf), m->Fp.assign(d->frame_pos, d->frame_pos + 3 * nf);
    m->depth.assign(nl, 0), m->anc.assign(nl, 0ull), m->link_of_dof.assign(n, -1), m->msub.assign(n, 0.0);
    std::vector<char> seen(n, 0);
    if (m->parent[0] != -1) { delete m; return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] parent[0] must be -1"); }
    for (int k = 1; k < nl; k++) {
        const int p = m->parent[k], j = m->dof[k];
        if (p < 0 || p >= k) { delete m; return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] links must be topologically ordered (parent[%d]=%d)", k, p); }
        if (j < 0 || j >= n || seen[j]) { delete m; return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] dof[] must be a permutation of 0..n-1"); }
        seen[j] = 1;
        m->link_of_dof[j] = k;
        m->depth[k] = m->depth[p] + 1;
        if (m->depth[k] > m->maxdepth) m->maxdepth = m->depth[k];
        m->anc[k] = m->anc[p] | (1ull << j);
    }
    for (int k = 0; k < nl; k++) {
        m->total_mass += m->mass[k];
        for (int j = 0; j < n; j++)
            if ((m->anc[k] >> j) & 1ull) m->msub[j] += m->mass[k];
    }
    if (!(m->total_mass > 0)) { delete m; return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] total mass must be positive"); }
    for (int f = 0; f < nf; f++)
        if (m->frame_link[f] < 0 || m->frame_link[f] >= nl) { delete m; return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_model_create] frame %d is attached to link %d", f, m->frame_link[f]); }
    *out = m;
    return BLFIK_OK;
}

extern "C" void blfik_model_destroy(blfik_model* m) { delete m; }

// ------------------------------------------------------------------- problem
template <typename T>
static int dev_upload(blfik_problem* p, const std::vector<T>& v, T** out)
{
    *out = nullptr;
    if (v.empty()) return BLFIK_OK;
    void* ptr = nullptr;
    CU(cudaMalloc(&ptr, v.size() * sizeof(T)));
    p->allocs.push_back(ptr);
    CU(cudaMemcpy(ptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    *out = static_cast<T*>(ptr);
    return BLFIK_OK;
}
template <typename T>
static int dev_alloc(blfik_problem* p, size_t count, T** out, bool zero = true)
{
    void* ptr = nullptr;
    CU(cudaMalloc(&ptr, (count ? count : 1) * sizeof(T)));
    p->allocs.push_back(ptr);
    if (zero) CU(cudaMemset(ptr, 0, (count ? count : 1) * sizeof(T)));
    *out = static_cast<T*>(ptr);
    return BLFIK_OK;
}

// Row layout of the equality-only kernel (blfik_eq.cuh): B (K x LD) in Cs, the diagonal cost in Hs (a vector here), no
// nv x nv matrix at all.  What only K1 / K2 touch (link poses, the task Jacobian scratch Jt, the unused box, and A, where a
// DistanceTask with a reference frame builds that frame's Jacobian) shares its words with what only K3 touches (G and its
// Cholesky factor in Ss, the inverse factor in Ym, the iteration vectors): the two phases never overlap, and the smaller
// footprint is what decides how many robots an SM holds.
// lanes the dense rows of the factored-form kernel are padded to (blfik_tick_eq_kernel<KP>); 0: the generic instantiation
static int eq_rider_rows(int K) { return K <= 16 ? 16 : (K <= 24 ? 24 : 0); }

static Layout make_layout_eq(int n, int me, int K, bool needs_A, bool box)
{
    Layout L;
    const int nv = n + 6, nl = n + 1;
    L.nv = nv, L.me = me, L.K = K, L.box = box ? 1 : 0;
    L.LD = (nv & 1) ? nv : nv + 1;
    L.LDE = (K & 1) ? K : K + 1;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 1) & ~1; return r; };
    // both phases
    L.Cs = take(K * L.LD), L.Hs = take(nv), L.g = take(nv), L.b = take(K), L.bd = take(K);
    if (box) L.lo = take(nv), L.hi = take(nv);
    const int shared0 = o;
    // K1 / K2 only
    L.linkR = take(9 * nl), L.linkP = take(3 * nl), L.axisW = take(3 * nl), L.mcw = take(3 * nl);
    L.Jt = take(6 * L.LD);
    if (!box) L.lo = take(nv), L.hi = take(nv);
    L.A = needs_A ? take(6 * L.LD) : L.Jt;
    const int end12 = o;
    // K3 only, over the same words
    o = shared0;
    // eq_solve_rider<KP> (K <= 24): the factor region holds KP x ld2odd(KP) + the rotating elimination's overshoot, the riders'
    // outputs and their products go where the generic kernel has no use for Wm
    const int KP = eq_rider_rows(K);
    const int fac = KP ? KP * ld2odd(KP) + KP : 0;
    L.Ss = take(K * L.LDE), L.Ym = take(K * L.LDE);
    if (o - L.Ss < fac) o = L.Ss + ((fac + 1) & ~1);
    L.Wm = KP ? take(9 * (KP + 1) + 81) : 0;
    const int KV = K > KP ? K : KP;
    L.x = take(nv), L.r = take(nv), L.gp = take(nv), L.wb = take(KV), L.lam = take(KV), L.bp = take(KV), L.t = take(KV);
    L.act = take((nv + 1) / 2 + 1);
    L.z = take(nv), L.y = take(nv), L.xb = take(nv);   // rho, the active-set codes, the codes proposed by the pivoting rule
    const int end3 = o;
    L.B = 0;
    L.total = end12 > end3 ? end12 : end3;
    return L;
}

static Layout make_layout(int n, int me)
{
    Layout L;
    const int nv = n + 6, nl = n + 1;
    L.nv = nv, L.me = me, L.K = 0, L.box = 0;
    L.LD = (nv & 1) ? nv : nv + 1;
    L.LDE = (me & 1) ? me : me + 1;
    int o = 0;
    auto take = [&](int cnt) { int r = o; o += (cnt + 1) & ~1; return r; };
    L.linkR = take(9 * nl), L.linkP = take(3 * nl), L.axisW = take(3 * nl), L.mcw = take(3 * nl);
    L.Hs = take(nv * L.LD), L.A = take(nv * L.LD), L.B = take(nv * L.LD);
    L.Cs = take(me * L.LD), L.Ym = take(nv * L.LDE), L.Wm = take(nv * L.LDE), L.Ss = take(me * L.LDE), L.Jt = take(6 * L.LD);
    const int vmax = nv > me ? nv : me;
    L.g = take(nv), L.b = take(me), L.lo = take(nv), L.hi = take(nv), L.x = take(nv), L.z = take(nv), L.y = take(nv);
    L.r = take(nv), L.xb = take(nv), L.wb = take(me), L.lam = take(vmax), L.gp = take(nv), L.bp = take(vmax), L.bd = take(nv), L.t = take(nv);
    L.act = take((nv + 1) / 2 + 1);
    L.total = o;
    return L;
}

static int refresh_tasks(blfik_problem* p)
{
    if (!p->tasks_dirty) return BLFIK_OK;
    std::vector<DevTask> dt(p->tasks.size());
    for (size_t t = 0; t < p->tasks.size(); t++) {
        HostTask& h = p->tasks[t];
        DevTask& d = dt[t];
        d.type = h.d.type, d.priority = h.d.priority, d.frame = h.d.frame, d.rows = h.rows, d.enabled = h.d.controller_enabled;
        d.sp_len = h.sp_len, d.row_offset = h.row_offset, d.brow = h.brow;
        d.flink = 0, d.pk_off = h.pk_off, d.pk_off_c = h.pk_off_c, d.base_c = h.base_c;
        d.exo_pos = h.d.use_position_exogenous_feedback != 0, d.exo_ori = h.d.use_orientation_exogenous_feedback != 0;
        d.fb_off = h.sp_len - h.fb_len;
        d.sp_stride = h.shared ? 0 : h.sp_len;
        d.traj_phase = h.d_traj_phase, d.traj_omega = h.d_traj_phase ? h.traj_omega : 0.0;
        for (int c = 0; c < 6; c++) d.traj_amp[c] = h.traj_amp[c];
        d.mj_target = h.mj_T > 0.0 ? h.d_mj_target : nullptr, d.mj_t0 = h.mj_t0, d.mj_T = h.mj_T;
        d.sp_nk = h.sp_nk, d.sp_times = h.d_sp_times, d.sp_coef = h.d_sp_coef, d.sp_dim = h.d.type == BLFIK_TASK_JOINT_TRACKING ? p->model->n : 3;
        d.gen = (d.traj_omega != 0.0 ? 1 : 0) | (d.mj_target ? 2 : 0) | (h.sp_nk >= 2 ? 4 : 0);
        for (int c = 0; c < 9; c++) d.Fr[c] = (c % 4 == 0) ? 1.0 : 0.0;
        for (int c = 0; c < 3; c++) d.Fp[c] = 0.0;
        d.flink2 = -1, d.thr = h.d.distance_threshold > 0 ? h.d.distance_threshold : 0.001;
        for (int c = 0; c < 3; c++) d.Fp2[c] = 0.0;
        if (h.d.type == BLFIK_TASK_DISTANCE && h.d.ref_frame >= 0 && h.d.ref_frame < p->model->nf) {
            d.flink2 = p->model->frame_link[h.d.ref_frame];
            for (int c = 0; c < 3; c++) d.Fp2[c] = p->model->Fp[3 * h.d.ref_frame + c];
        }
        if (h.d.frame >= 0 && h.d.frame < p->model->nf) {
            d.flink = p->model->frame_link[h.d.frame];
            for (int c = 0; c < 9; c++) d.Fr[c] = p->model->Fr[9 * h.d.frame + c];
            for (int c = 0; c < 3; c++) d.Fp[c] = p->model->Fp[3 * h.d.frame + c];
        }
        for (int c = 0; c < 3; c++) d.mask[c] = h.d.mask[c] != 0, d.kp_lin[c] = h.d.kp_linear[c], d.kp_ang[c] = h.d.kp_angular[c];
        auto up = [&](const std::vector<double>& v, double* dst) -> int {
            if (!v.empty()) CU(cudaMemcpyAsync(dst, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, p->stream));
            return BLFIK_OK;
        };
        int rc;
        if ((rc = up(h.kp_joint, h.d_kp_joint)) || (rc = up(h.weight, h.d_weight)) || (rc = up(h.klim, h.d_klim)) || (rc = up(h.upper, h.d_upper)) || (rc = up(h.lower, h.d_lower))) return rc;
        d.kp_joint = h.d_kp_joint, d.weight = h.d_weight, d.klim = h.d_klim, d.upper = h.d_upper, d.lower = h.d_lower;
        d.inv_dt = h.d.sampling_time > 0 ? 1.0 / h.d.sampling_time : 0.0;
        d.setpoint = h.d_setpoint;
    }
    if (!dt.empty()) CU(cudaMemcpyAsync(p->d_tasks, dt.data(), dt.size() * sizeof(DevTask), cudaMemcpyHostToDevice, p->stream));
    CU(cudaStreamSynchronize(p->stream));  // dt is a stack-lifetime host buffer
    p->tasks_dirty = false;
    return BLFIK_OK;
}

extern "C" int blfik_problem_create(const blfik_model* model, const blfik_task_desc* tasks, int32_t ntasks, int64_t N,
                                    int32_t device, blfik_problem** out)
{
    if (!model || !out || (ntasks > 0 && !tasks)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] null argument");
    if (N < 1) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] n_instances must be positive");
    if (ntasks < 1 || ntasks > 64) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] ntasks must be in [1, 64]");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(BLFIK_ERR_CUDA, "[blfik_problem_create] no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] device %d out of range (%d devices)", device, ndev);
    CU(cudaSetDevice(device));

    blfik_problem* p = new (std::nothrow) blfik_problem;
    if (!p) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] out of host memory");
    p->model = model, p->device = device, p->N = N, p->n = model->n, p->nv = model->n + 6;
    const int n = p->n;
    blfik_solver_settings ds;
    blfik_default_settings(&ds);
    p->st = {ds.rho, ds.sigma, ds.alpha, ds.eps, ds.admm_iters, ds.max_rounds, ds.polish_iters, ds.eq_iters};

    auto bail = [&](int rc) { blfik_problem_destroy(p); return rc; };
    // ---- addTask rules, QPInverseKinematics.cpp:167-249
    int me = 0;
    for (int t = 0; t < ntasks; t++) {
        HostTask h;
        h.d = tasks[t];
        h.rows = rows_of(h.d, n);
        if (h.rows < 0) return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: unknown type %d", t, h.d.type));
        if (h.d.priority != 0 && h.d.priority != 1)
            return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: priority must be 0 or 1 (got %d)", t, h.d.priority));
        if (is_inequality(h.d.type) && h.d.priority != 0)
            return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: inequality tasks can only be hard constraints (priority 0)", t));
        if (h.d.priority == 1 && !h.d.weight)
            return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: a priority-1 task needs a weight", t));
        if (h.d.type == BLFIK_TASK_DISTANCE && h.d.ref_frame >= model->nf)
            return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: reference frame index %d out of range", t, h.d.ref_frame));
        if (has_frame(h.d.type) && (h.d.frame < 0 || h.d.frame >= model->nf))
            return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: frame index %d out of range", t, h.d.frame));
        if (h.rows == 0) return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: the mask selects no row", t));
        h.sp_len = sp_len_of(h.d.type, n);
        if (h.d.type == BLFIK_TASK_SE3 && (h.d.use_position_exogenous_feedback || h.d.use_orientation_exogenous_feedback)) h.fb_len = 7;
        else if (h.d.type == BLFIK_TASK_COM && h.d.use_position_exogenous_feedback) h.fb_len = 3;
        else h.d.use_position_exogenous_feedback = h.d.use_orientation_exogenous_feedback = 0;   // the other tasks have no such option
        h.sp_len += h.fb_len;
        if (h.d.type == BLFIK_TASK_JOINT_TRACKING) {
            if (!h.d.kp_joint) return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: JointTrackingTask needs kp", t));
            h.kp_joint.assign(h.d.kp_joint, h.d.kp_joint + n);
        }
        if (h.d.type == BLFIK_TASK_JOINT_LIMITS) {
            if (!h.d.klim || !h.d.upper_limits || !h.d.lower_limits || !(h.d.sampling_time > 0))
                return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: JointLimitsTask needs klim, limits and a positive sampling_time", t));
            h.klim.assign(h.d.klim, h.d.klim + n);
            for (int j = 0; j < n; j++)
                if (!(h.klim[j] >= 0 && h.klim[j] <= 1))
                    return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: klim must be in [0, 1]", t));
        }
        if (is_inequality(h.d.type)) {
            if (!h.d.upper_limits || !h.d.lower_limits) return bail(fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_problem_create] task %d: limits missing", t));
            h.upper.assign(h.d.upper_limits, h.d.upper_limits + n), h.lower.assign(h.d.lower_limits, h.d.lower_limits + n);
        }
        if (h.d.priority == 1) h.weight.assign(h.d.weight, h.d.weight + h.rows);
        if (h.d.priority == 0 && is_dense(h.d.type)) {
            h.row_offset = me;
            me += h.rows;
        }
        if (is_dense(h.d.type)) {
            h.brow = p->K;
            p->K += h.rows;
        }
        h.d.kp_joint = h.d.weight = h.d.klim = h.d.upper_limits = h.d.lower_limits = nullptr;  // owned copies only
        h.setpoint_set = h.sp_len == 0;
        p->tasks.push_back(std::move(h));
    }
    layout_packed(p);
    p->packed_len_max = p->packed_len;
    if (me > 64) return bail(fail(BLFIK_ERR_UNSUPPORTED, "[blfik_problem_create] more than 64 dense equality rows (%d)", me));
    p->me = me;
    p->L = make_layout(n, me);

    // ---- launch geometry: as many warps (robots) per SM as shared memory allows
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return bail(fail(BLFIK_ERR_CUDA, "cudaGetDeviceProperties failed"));
    p->num_sms = prop.multiProcessorCount;
    const size_t per_warp = (size_t)p->L.total * sizeof(double);
    const size_t smem_max = prop.sharedMemPerBlockOptin;
    if (per_warp > smem_max) return bail(fail(BLFIK_ERR_UNSUPPORTED, "[blfik_problem_create] one robot needs %zu B of shared memory, the device offers %zu", per_warp, smem_max));
    const size_t sm_total = prop.sharedMemPerMultiprocessor;
    int warps_per_sm = (int)((sm_total - 2048) / (per_warp + 256));
    if (warps_per_sm < 1) warps_per_sm = 1;
    int wpb = warps_per_sm >= 8 ? 4 : (warps_per_sm >= 4 ? 2 : 1);
    while ((size_t)wpb * per_warp > smem_max && wpb > 1) wpb >>= 1;
    int blocks_per_sm = warps_per_sm / wpb;
    if (blocks_per_sm < 1) blocks_per_sm = 1;
    p->warps_per_block = wpb;
    p->smem_bytes = (size_t)wpb * per_warp;
    long long want = (N + wpb - 1) / wpb;
    long long cap = (long long)p->num_sms * blocks_per_sm;
    p->grid = (int)(want < cap ? want : cap);
    // The limit is a property of the kernel, not of this problem: several problems of different sizes share the kernels, so
    // every one of them raises it to the device's opt-in maximum (a smaller, per-problem value would lower it for the others).
    const int smem_optin = (int)prop.sharedMemPerBlockOptin - 1024;
    if (p->smem_bytes > (size_t)smem_optin ||
        cudaFuncSetAttribute(blfik_tick_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin) != cudaSuccess ||
        cudaFuncSetAttribute(blfik_kinematics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin) != cudaSuccess)
        return bail(fail(BLFIK_ERR_CUDA, "[blfik_problem_create] cannot reserve %zu B of shared memory: %s", p->smem_bytes, cudaGetErrorString(cudaGetLastError())));

    // ---- register-resident kernel: pick the smallest instantiation that holds (nv, K)
    {
        const int nv = p->nv, K = p->K;
        bool angmom = false;
        for (const HostTask& h : p->tasks) angmom |= h.d.type == BLFIK_TASK_ANGULAR_MOMENTUM;
        if (nv <= 32 && K <= 32) {
            if (angmom) {
                // the centroidal momentum rows live in their own instantiations (largest row count only: the task is rare)
                if (nv == 29) pick_fast<29, 32, true>(p);
                else pick_fast<32, 32, true>(p);
            } else if (nv == 29) {
                if (K <= 16) pick_fast<29, 16>(p);
                else if (K <= 18) pick_fast<29, 18>(p);
                else if (K <= 30) pick_fast<29, 30>(p);
                else pick_fast<29, 32>(p);
            } else {
                if (K <= 16) pick_fast<32, 16>(p);
                else if (K <= 20) pick_fast<32, 20>(p);
                else pick_fast<32, 32>(p);
            }
            std::vector<double> pack = pack_model(*model, p->F, p->njump);
            p->F.sh_model = 0;
            p->F.sh_tasks = p->F.sh_model_len;
            p->F.sh_total = p->F.sh_tasks + (int)(((size_t)ntasks * sizeof(DevTask) + 15) / 16 * 2);
            for (const HostTask& h : p->tasks) p->has_com |= h.d.type == BLFIK_TASK_COM || h.d.type == BLFIK_TASK_ANGULAR_MOMENTUM;   // both need the CoM
            const size_t fshared = (size_t)p->F.sh_total * sizeof(double);
            const size_t fper = (size_t)p->F.total * sizeof(double);
            int fw = (int)((sm_total - 3072 - fshared) / (fper + 64));
            if (fw > p->fast_wpb) fw = p->fast_wpb;   // register file: the instantiation's launch bound (FastDims::WARPS)
            if (fw < p->fast_wpb) p->fast_wpb = fw < 1 ? 1 : fw;   // one block per SM holding all its resident warps
            int fb = fw / p->fast_wpb;
            if (fb < 1) {
                p->fast_kernel = nullptr;    // does not fit: generic kernel
            } else {
                p->fast_smem = (size_t)p->fast_wpb * fper + fshared;
                if (ntasks * 12 > p->F.total - p->F.Us || ntasks > 32) p->fast_kernel = nullptr;   // control buffer must fit the U region
                void* dp = nullptr;
                if (cudaMalloc(&dp, pack.size() * sizeof(double)) != cudaSuccess) return bail(fail(BLFIK_ERR_CUDA, "[blfik_problem_create] cudaMalloc failed"));
                p->allocs.push_back(dp);
                if (cudaMemcpy(dp, pack.data(), pack.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess)
                    return bail(fail(BLFIK_ERR_CUDA, "[blfik_problem_create] cudaMemcpy failed"));
                p->d_model_pack = static_cast<double*>(dp);
                const long long fwant = (N + p->fast_wpb - 1) / p->fast_wpb;
                const long long fcap = (long long)p->num_sms * fb;
                p->fast_grid = (int)(fwant < fcap ? fwant : fcap);
                if (p->fast_smem > (size_t)smem_optin ||
                    cudaFuncSetAttribute((const void*)p->fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin) != cudaSuccess)
                    return bail(fail(BLFIK_ERR_CUDA, "[blfik_problem_create] cannot reserve %zu B of shared memory for the register-resident kernel: %s",
                                     p->fast_smem, cudaGetErrorString(cudaGetLastError())));
                // the two-launch path shares the packed model; its kernels size their own shared memory
                p->pack_len = p->F.sh_model_len;
                {
                    // depth-first walk of the tree for the assemble kernel: visit a link, its subtree, then close it
                    const int nl = n + 1;
                    std::vector<std::vector<int>> kids(nl);
                    for (int k = 1; k < nl; k++) kids[model->parent[k]].push_back(k);
                    std::vector<int> ev;
                    std::vector<std::pair<int, size_t>> stack;   // (link, next child)
                    stack.push_back({0, 0});
                    while (!stack.empty()) {
                        auto& top = stack.back();
                        if (top.second < kids[top.first].size()) {
                            const int c = kids[top.first][top.second++];
                            ev.push_back((0 << 16) | ((kids[c].size() >= 2 ? 1 : 0) << 17) | (model->depth[c] << 8) | c);   // bit 17: the walk comes back to this link
                            stack.push_back({c, 0});
                        } else {
                            if (top.first != 0) ev.push_back((1 << 16) | (model->depth[top.first] << 8) | top.first);
                            stack.pop_back();
                        }
                    }
                    p->asm_nevents = (int)ev.size(), p->asm_D = model->maxdepth + 1;
                    const int tile_doubles = 4 * 32 * (p->rec_kp | 1);   // one [32][KP | 1] staging tile per warp
                    int ndense = 0;
                    bool unsupported = false;
                    for (const HostTask& h : p->tasks) {
                        ndense += is_dense(h.d.type);
                        unsupported |= h.d.type == BLFIK_TASK_DISTANCE || h.d.type == BLFIK_TASK_ANGULAR_MOMENTUM;
                    }
                    const size_t fixed = (size_t)p->pack_len + ((size_t)ntasks * sizeof(DevTask) + 7) / 8 + (ev.size() + 1) / 2 + 1 + 4 * (size_t)ntasks;
                    p->asm_threads = 128;
                    p->asm_smem = (fixed + (size_t)tile_doubles) * sizeof(double);
                    int rc2 = BLFIK_OK;
                    if (!p->fast_kernel || unsupported || ndense > kAsmMaxDense || p->asm_D > kAsmMaxDepth || p->asm_smem > (size_t)smem_optin || p->solve_smem > (size_t)smem_optin ||
                        (rc2 = dev_upload(p, ev, &p->d_asm_events)) != BLFIK_OK ||
                        cudaFuncSetAttribute((const void*)p->asm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin) != cudaSuccess ||
                        cudaFuncSetAttribute((const void*)p->solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->solve_smem) != cudaSuccess) {
                        cudaGetLastError();
                        p->solve_kernel = nullptr;
                    }
                }
            }
            if (!p->fast_kernel) p->solve_kernel = nullptr;
        }
    }

    // ---- factored-form kernel for what the register-resident one cannot hold: at most 64 dense rows
    if (!p->fast_kernel && p->K >= 1 && p->K <= 64) {
        bool box = false;
        for (const HostTask& h : p->tasks)
            box |= is_inequality(h.d.type) || (h.d.type == BLFIK_TASK_JOINT_TRACKING && h.d.priority == 0);
        static const int box_env = [] { const char* e = getenv("BLFIK_EQ_BOX"); return e ? atoi(e) : 1; }();
        if (!box || box_env) {
            bool needs_A = false;
            for (const HostTask& h : p->tasks) needs_A |= h.d.type == BLFIK_TASK_DISTANCE && h.d.ref_frame >= 0;
            p->Leq = make_layout_eq(n, me, p->K, needs_A, box);
            const size_t eper = (size_t)p->Leq.total * sizeof(double);
            int ew = (int)((sm_total - 2048) / (eper + 64));
            if (ew > 16) ew = 16;                      // register file: two blocks of 256 threads at <= 128 registers
            int ewpb = 1;                              // block size (<= 8 warps) that leaves the fewest warps of the SM unused
            for (int c = 1; c <= 8; c++)
                if ((ew / c) * c >= (ew / ewpb) * ewpb) ewpb = c;
            const int ekp = eq_rider_rows(p->K);
            p->eq_func = ekp == 16 ? blfik_tick_eq_kernel<16> : (ekp == 24 ? blfik_tick_eq_kernel<24> : blfik_tick_eq_kernel<0>);
            if (ew >= 1 && (size_t)ewpb * eper <= (size_t)smem_optin &&
                cudaFuncSetAttribute((const void*)p->eq_func, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin) == cudaSuccess) {
                p->eq_kernel = true;
                p->eq_wpb = ewpb, p->eq_smem = (size_t)ewpb * eper;
                const long long ewant = (N + ewpb - 1) / ewpb;
                const long long ecap = (long long)p->num_sms * (ew / ewpb);
                p->eq_grid = (int)(ewant < ecap ? ewant : ecap);
            }
        }
    }

    if (cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(fail(BLFIK_ERR_CUDA, "cudaStreamCreate failed"));
    p->own_stream = true;

    // ---- device copies
    int rc;
    DevModel& dm = p->dm;
    dm.n = n, dm.nl = n + 1, dm.nf = model->nf, dm.maxdepth = model->maxdepth, dm.inv_total_mass = 1.0 / model->total_mass;
    int *d_parent, *d_dof, *d_depth, *d_lod, *d_fl;
    unsigned long long* d_anc;
    double *d_Xr, *d_Xp, *d_axis, *d_mass, *d_com, *d_msub, *d_Fr, *d_Fp, *d_inertia;
    if ((rc = dev_upload(p, model->parent, &d_parent)) || (rc = dev_upload(p, model->dof, &d_dof)) || (rc = dev_upload(p, model->depth, &d_depth)) ||
        (rc = dev_upload(p, model->link_of_dof, &d_lod)) || (rc = dev_upload(p, model->frame_link, &d_fl)) || (rc = dev_upload(p, model->anc, &d_anc)) ||
        (rc = dev_upload(p, model->Xr, &d_Xr)) || (rc = dev_upload(p, model->Xp, &d_Xp)) || (rc = dev_upload(p, model->axis, &d_axis)) ||
        (rc = dev_upload(p, model->mass, &d_mass)) || (rc = dev_upload(p, model->com, &d_com)) || (rc = dev_upload(p, model->msub, &d_msub)) ||
        (rc = dev_upload(p, model->Fr, &d_Fr)) || (rc = dev_upload(p, model->Fp, &d_Fp)) || (rc = dev_upload(p, model->inertia, &d_inertia)))
        return bail(rc);
    dm.parent = d_parent, dm.dof = d_dof, dm.depth = d_depth, dm.link_of_dof = d_lod, dm.frame_link = d_fl, dm.anc = d_anc;
    dm.Xr = d_Xr, dm.Xp = d_Xp, dm.axis = d_axis, dm.mass = d_mass, dm.com = d_com, dm.msub = d_msub, dm.Fr = d_Fr, dm.Fp = d_Fp, dm.inertia = d_inertia;

    const size_t Ns = (size_t)N;
    if ((rc = dev_alloc(p, Ns * 3, &p->base_pos)) || (rc = dev_alloc(p, Ns * 4, &p->base_quat)) || (rc = dev_alloc(p, Ns * n, &p->s)) ||
        (rc = dev_alloc(p, Ns * p->nv, &p->nu)) || (rc = dev_alloc(p, Ns * p->nv, &p->wx)) || (rc = dev_alloc(p, Ns * n, &p->wz)) ||
        (rc = dev_alloc(p, Ns * n, &p->wy)) || (rc = dev_alloc(p, Ns, &p->status)) || (rc = dev_alloc(p, Ns, &p->iters)) ||
        (rc = dev_alloc(p, 1, &p->d_count)) || (rc = dev_alloc(p, p->tasks.size(), &p->d_tasks)) || (rc = dev_alloc(p, kCounterSlots, &p->work_counter)))
        return bail(rc);
    for (HostTask& h : p->tasks) {
        if (h.sp_len && (rc = dev_alloc(p, Ns * h.sp_len, &h.d_setpoint))) return bail(rc);
        if (h.fb_len == 7) {   // the controllers' initial state: zero position, identity rotation (quaternion x, y, z, w)
            const std::vector<double> ones(Ns, 1.0);
            if (cudaMemcpy2D(h.d_setpoint + h.sp_len - 1, h.sp_len * sizeof(double), ones.data(), sizeof(double), sizeof(double), Ns, cudaMemcpyHostToDevice) != cudaSuccess)
                return bail(fail(BLFIK_ERR_CUDA, "[blfik_problem_create] cudaMemcpy2D failed: %s", cudaGetErrorString(cudaGetLastError())));
        }
        if ((rc = dev_upload(p, h.kp_joint, &h.d_kp_joint)) || (rc = dev_upload(p, h.weight, &h.d_weight)) || (rc = dev_upload(p, h.klim, &h.d_klim)) ||
            (rc = dev_upload(p, h.upper, &h.d_upper)) || (rc = dev_upload(p, h.lower, &h.d_lower)))
            return bail(rc);
    }
    p->tasks_dirty = true;
    *out = p;
    return BLFIK_OK;
}

extern "C" void blfik_problem_destroy(blfik_problem* p)
{
    if (!p) return;
    cudaSetDevice(p->device);
    if (p->stream) cudaStreamSynchronize(p->stream);
    for (void* a : p->allocs) cudaFree(a);
    for (HostTask& h : p->tasks) cudaFree(h.d_sp_times), cudaFree(h.d_sp_fac), cudaFree(h.d_sp_coef);
    if (p->h2d_stream) {
        cudaStreamSynchronize(p->h2d_stream), cudaStreamSynchronize(p->d2h_stream);
        for (int c = 0; c < kMaxChunks; c++) cudaEventDestroy(p->ev_in[c]), cudaEventDestroy(p->ev_done[c]);
        cudaEventDestroy(p->ev_start);
        cudaStreamDestroy(p->h2d_stream), cudaStreamDestroy(p->d2h_stream);
        for (int k = 0; k < 4; k++)
            if (p->chunk_stream[k]) cudaStreamSynchronize(p->chunk_stream[k]), cudaStreamDestroy(p->chunk_stream[k]);
        for (int k = 0; k < 4; k++)
            if (p->asm_stream[k]) cudaStreamSynchronize(p->asm_stream[k]), cudaStreamDestroy(p->asm_stream[k]);
        for (int c = 0; c < kMaxChunks; c++)
            if (p->ev_asm[c]) cudaEventDestroy(p->ev_asm[c]);
    }
    if (p->h_stream_ctl) cudaFreeHost(p->h_stream_ctl);
    for (int k = 0; k < 3; k++)
        if (p->ev_t[k]) cudaEventDestroy(p->ev_t[k]);
    if (p->own_stream && p->stream) cudaStreamDestroy(p->stream);
    delete p;
}

extern "C" int blfik_set_solver_settings(blfik_problem* p, const blfik_solver_settings* s)
{
    if (!p || !s) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_solver_settings] null argument");
    if (!(s->rho > 0) || !(s->sigma > 0) || !(s->alpha > 0 && s->alpha < 2) || !(s->eps > 0) || s->admm_iters < 1 || s->max_rounds < 1 || s->polish_iters < 1 || s->eq_iters < 1)
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_solver_settings] rho, sigma, eps must be positive, alpha in (0, 2), iteration counts >= 1");
    p->st = {s->rho, s->sigma, s->alpha, s->eps, s->admm_iters, s->max_rounds, s->polish_iters, s->eq_iters};
    return BLFIK_OK;
}

static cudaMemcpyKind kind_in(int mem) { return mem == BLFIK_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice; }
static cudaMemcpyKind kind_out(int mem) { return mem == BLFIK_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost; }

#define CHECK_P(p, fn)                                                                         \
    do {                                                                                       \
        if (!(p)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[" fn "] null problem handle");     \
        CU(cudaSetDevice((p)->device));                                                        \
    } while (0)

extern "C" int blfik_set_state(blfik_problem* p, const double* bp, const double* bq, const double* s, int32_t mem)
{
    CHECK_P(p, "blfik_set_state");
    if (!bp || !bq || !s) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_state] null buffer");
    const size_t N = (size_t)p->N;
    CU(cudaMemcpyAsync(p->base_pos, bp, N * 3 * sizeof(double), kind_in(mem), p->stream));
    CU(cudaMemcpyAsync(p->base_quat, bq, N * 4 * sizeof(double), kind_in(mem), p->stream));
    CU(cudaMemcpyAsync(p->s, s, N * p->n * sizeof(double), kind_in(mem), p->stream));
    p->state_set = true;
    // a state given from outside starts a new history: what the last ticks learnt about slow and unsolved robots is dropped
    p->slow_list_for = -1;
    if (p->d_slow) CU(cudaMemsetAsync(p->d_slow + 4 * N + 2, 0, N * sizeof(int), p->stream));
    return BLFIK_OK;
}

extern "C" int blfik_get_state(blfik_problem* p, double* bp, double* bq, double* s, int32_t mem)
{
    CHECK_P(p, "blfik_get_state");
    if (!p->state_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_get_state] the state has never been set");
    const size_t N = (size_t)p->N;
    if (bp) CU(cudaMemcpyAsync(bp, p->base_pos, N * 3 * sizeof(double), kind_out(mem), p->stream));
    if (bq) CU(cudaMemcpyAsync(bq, p->base_quat, N * 4 * sizeof(double), kind_out(mem), p->stream));
    if (s) CU(cudaMemcpyAsync(s, p->s, N * p->n * sizeof(double), kind_out(mem), p->stream));
    if (mem == BLFIK_MEM_HOST) CU(cudaStreamSynchronize(p->stream));
    return BLFIK_OK;
}

#define CHECK_TASK(p, t, fn)                                                                                     \
    do {                                                                                                         \
        if ((t) < 0 || (t) >= (int)(p)->tasks.size())                                                            \
            return fail(BLFIK_ERR_INVALID_ARGUMENT, "[" fn "] task index %d out of range (%d tasks)", (int)(t), (int)(p)->tasks.size()); \
    } while (0)

extern "C" int32_t blfik_setpoint_len(const blfik_problem* p, int32_t t)
{
    if (!p || t < 0 || t >= (int)p->tasks.size()) return -1;
    return p->tasks[t].sp_len - p->tasks[t].fb_len;
}

extern "C" int blfik_set_setpoint(blfik_problem* p, int32_t t, const double* data, int32_t mem)
{
    CHECK_P(p, "blfik_set_setpoint");
    CHECK_TASK(p, t, "blfik_set_setpoint");
    HostTask& h = p->tasks[t];
    if (h.sp_len == 0) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint] task %d takes no set-point", t);
    if (!data) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint] null buffer");
    const int base = h.sp_len - h.fb_len;
    if (h.shared) {   // back to one set-point per robot
        h.shared = false;
        layout_packed(p);
        p->tasks_dirty = true;
    }
    if (h.fb_len == 0)
        CU(cudaMemcpyAsync(h.d_setpoint, data, (size_t)p->N * h.sp_len * sizeof(double), kind_in(mem), p->stream));
    else   // the record is wider than the set-point: strided rows
        CU(cudaMemcpy2DAsync(h.d_setpoint, h.sp_len * sizeof(double), data, base * sizeof(double), base * sizeof(double), (size_t)p->N, kind_in(mem), p->stream));
    h.setpoint_set = true;
    return BLFIK_OK;
}

extern "C" int blfik_set_setpoint_shared(blfik_problem* p, int32_t t, const double* data, int32_t mem)
{
    CHECK_P(p, "blfik_set_setpoint_shared");
    CHECK_TASK(p, t, "blfik_set_setpoint_shared");
    HostTask& h = p->tasks[t];
    if (h.sp_len == 0) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_shared] task %d takes no set-point", t);
    if (h.fb_len != 0) return fail(BLFIK_ERR_UNSUPPORTED, "[blfik_set_setpoint_shared] task %d carries an exogenous feedback per robot", t);
    if (!data) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_shared] null buffer");
    CU(cudaMemcpyAsync(h.d_setpoint, data, (size_t)h.sp_len * sizeof(double), kind_in(mem), p->stream));
    if (mem == BLFIK_MEM_HOST) CU(cudaStreamSynchronize(p->stream));   // the caller's buffer may be pageable and short-lived
    if (!h.shared) {
        h.shared = true;
        layout_packed(p);
        p->tasks_dirty = true;
    }
    h.setpoint_set = true;
    return BLFIK_OK;
}

extern "C" int blfik_set_feedback(blfik_problem* p, int32_t t, const double* data, int32_t mem)
{
    CHECK_P(p, "blfik_set_feedback");
    CHECK_TASK(p, t, "blfik_set_feedback");
    HostTask& h = p->tasks[t];
    if (h.fb_len == 0) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_feedback] task %d does not use an exogenous feedback", t);
    if (!data) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_feedback] null buffer");
    CU(cudaMemcpy2DAsync(h.d_setpoint + (h.sp_len - h.fb_len), h.sp_len * sizeof(double), data, h.fb_len * sizeof(double), h.fb_len * sizeof(double),
                         (size_t)p->N, kind_in(mem), p->stream));
    return BLFIK_OK;
}

extern "C" int32_t blfik_feedback_len(const blfik_problem* p, int32_t t)
{
    if (!p || t < 0 || t >= (int)p->tasks.size()) return -1;
    return p->tasks[t].fb_len;
}

extern "C" int32_t blfik_record_len(const blfik_problem* p, int32_t t)
{
    if (!p || t < 0 || t >= (int)p->tasks.size()) return -1;
    return p->tasks[t].sp_len;
}

extern "C" int blfik_set_setpoint_minimum_jerk(blfik_problem* p, int32_t t, const double* target, double duration, int32_t mem)
{
    CHECK_P(p, "blfik_set_setpoint_minimum_jerk");
    CHECK_TASK(p, t, "blfik_set_setpoint_minimum_jerk");
    HostTask& h = p->tasks[t];
    if (h.d.type != BLFIK_TASK_SE3 && h.d.type != BLFIK_TASK_SO3 && h.d.type != BLFIK_TASK_R3 && h.d.type != BLFIK_TASK_COM)
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_minimum_jerk] task %d is not an SE3 / SO3 / R3 / CoM task", t);
    if (!(duration > 0)) {   // SO3Planner::setRotations rejects a non-positive duration (SO3Planner.tpp:118-122); here it switches the generator off
        h.mj_T = 0.0;
        p->tasks_dirty = true;
        return BLFIK_OK;
    }
    if (!target) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_minimum_jerk] null target");
    if (!h.setpoint_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_set_setpoint_minimum_jerk] the initial set point of task %d has not been set", t);
    const int base = h.sp_len - h.fb_len;
    int rc;
    if (!h.d_mj_target && (rc = dev_alloc(p, (size_t)p->N * base, &h.d_mj_target))) return rc;
    CU(cudaMemcpyAsync(h.d_mj_target, target, (size_t)p->N * base * sizeof(double), kind_in(mem), p->stream));
    h.mj_t0 = p->time, h.mj_T = duration;
    h.traj_omega = 0.0;   // one generator per task
    p->tasks_dirty = true;
    return BLFIK_OK;
}

// 2 x 2 helpers for the block elimination of the spline system
static void m22_mul(const double* a, const double* b, double* o)
{
    o[0] = a[0] * b[0] + a[1] * b[2], o[1] = a[0] * b[1] + a[1] * b[3];
    o[2] = a[2] * b[0] + a[3] * b[2], o[3] = a[2] * b[1] + a[3] * b[3];
}
static bool m22_inv(const double* a, double* o)
{
    const double det = a[0] * a[3] - a[1] * a[2];
    const double scale = std::fabs(a[0] * a[3]) + std::fabs(a[1] * a[2]);
    if (!(std::fabs(det) > 1e-13 * scale)) return false;
    o[0] = a[3] / det, o[1] = -a[1] / det, o[2] = -a[2] / det, o[3] = a[0] / det;
    return true;
}

static int set_setpoint_spline(blfik_problem* p, int32_t t, int32_t n_knots, const double* times, const double* positions,
                               const double* boundary, int32_t mem, bool cubic)
{
    CHECK_P(p, "blfik_set_setpoint_spline");
    CHECK_TASK(p, t, "blfik_set_setpoint_spline");
    HostTask& h = p->tasks[t];
    if (h.d.type != BLFIK_TASK_SE3 && h.d.type != BLFIK_TASK_R3 && h.d.type != BLFIK_TASK_COM && h.d.type != BLFIK_TASK_JOINT_TRACKING)
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_spline] task %d is not an SE3 / R3 / CoM / JointTracking task", t);
    const int dim = h.d.type == BLFIK_TASK_JOINT_TRACKING ? p->model->n : 3;   // Spline<VectorXd> in joint space, Spline<Vector3d> otherwise
    if (n_knots == 0) {   // switch the generator off
        h.sp_nk = 0;
        p->tasks_dirty = true;
        return BLFIK_OK;
    }
    if (n_knots < 2)   // Spline.h:338-342
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_spline] the number of knots has to be at least 2");
    if (n_knots > BLFIK_MAX_SPLINE_KNOTS)
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_spline] at most %d knots", BLFIK_MAX_SPLINE_KNOTS);
    if (!times || !positions) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_spline] null times or positions");
    if (!h.setpoint_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_set_setpoint_spline] the set point of task %d has not been set", t);
    for (int i = 0; i + 1 < n_knots; i++) {
        if (times[i + 1] == times[i])   // Spline.h:444-448
            return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_spline] two consecutive knots have the same time coordinate");
        if (!(times[i + 1] > times[i])) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_spline] the knot times must increase");
    }
    // block elimination of the interior-knot system (times only): per interior knot L, M = L D'^-1 of the previous knot, D'^-1, U
    const int m = n_knots - 2;
    std::vector<double> fac((size_t)16 * (m > 0 ? m : 1), 0.0);
    double Dprev_inv[4] = {0, 0, 0, 0}, Uprev[4] = {0, 0, 0, 0};
    for (int i = 0; i < m; i++) {
        const double Tp = times[i + 1] - times[i], Tn = times[i + 2] - times[i + 1];
        double* f = &fac[(size_t)16 * i];
        const double L[4] = {-24.0 / (Tp * Tp), -3.0 / Tp, -168.0 / (Tp * Tp * Tp), -24.0 / (Tp * Tp)};
        double D[4] = {-36.0 / (Tp * Tp) + 36.0 / (Tn * Tn), 9.0 / Tp + 9.0 / Tn,
                       -192.0 / (Tp * Tp * Tp) - 192.0 / (Tn * Tn * Tn), 36.0 / (Tp * Tp) - 36.0 / (Tn * Tn)};
        const double U[4] = {24.0 / (Tn * Tn), -3.0 / Tn, -168.0 / (Tn * Tn * Tn), 24.0 / (Tn * Tn)};
        if (cubic) {   // scalar system: 2/Tp v_{i-1} + 4 (1/Tp + 1/Tn) v_i + 2/Tn v_{i+1} = 6 (dp/Tp^2 + dn/Tn^2)
            double d = 4.0 / Tp + 4.0 / Tn;
            f[0] = 2.0 / Tp, f[12] = 2.0 / Tn;
            if (i > 0) {
                f[4] = f[0] * Dprev_inv[0];
                d -= f[4] * Uprev[0];
            }
            f[8] = 1.0 / d;   // strictly diagonally dominant: d >= 2/Tp + 2/Tn > 0 after the elimination
            Dprev_inv[0] = f[8], Uprev[0] = f[12];
            continue;
        }
        for (int k = 0; k < 4; k++) f[k] = L[k], f[12 + k] = U[k];
        if (i > 0) {
            double MU[4];
            m22_mul(L, Dprev_inv, f + 4);
            m22_mul(f + 4, Uprev, MU);
            for (int k = 0; k < 4; k++) D[k] -= MU[k];
        }
        if (!m22_inv(D, f + 8)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_spline] the knot times give a singular spline system");
        for (int k = 0; k < 4; k++) Dprev_inv[k] = f[8 + k], Uprev[k] = U[k];
    }
    if (n_knots > h.sp_cap) {
        CU(cudaStreamSynchronize(p->stream));   // a tick in flight may still read the old buffers
        cudaFree(h.d_sp_times), cudaFree(h.d_sp_fac), cudaFree(h.d_sp_coef);
        h.d_sp_times = h.d_sp_fac = h.d_sp_coef = nullptr, h.sp_cap = 0, h.sp_nk = 0;
        p->tasks_dirty = true;
        CU(cudaMalloc((void**)&h.d_sp_times, (size_t)n_knots * sizeof(double)));
        CU(cudaMalloc((void**)&h.d_sp_fac, (size_t)16 * n_knots * sizeof(double)));
        CU(cudaMalloc((void**)&h.d_sp_coef, (size_t)p->N * (n_knots - 1) * 6 * dim * sizeof(double)));
        h.sp_cap = n_knots;
    }
    const size_t npos = (size_t)p->N * n_knots * dim, nbc = (size_t)p->N * (cubic ? 2 : 4) * dim;
    const double *d_x = positions, *d_bc = boundary;
    double* stage = nullptr;
    if (mem != BLFIK_MEM_DEVICE) {
        CU(cudaMalloc((void**)&stage, (npos + nbc) * sizeof(double)));
        cudaMemcpyAsync(stage, positions, npos * sizeof(double), cudaMemcpyHostToDevice, p->stream);
        if (boundary) cudaMemcpyAsync(stage + npos, boundary, nbc * sizeof(double), cudaMemcpyHostToDevice, p->stream);
        d_x = stage, d_bc = boundary ? stage + npos : nullptr;
    }
    cudaMemcpyAsync(h.d_sp_times, times, (size_t)n_knots * sizeof(double), cudaMemcpyHostToDevice, p->stream);
    cudaMemcpyAsync(h.d_sp_fac, fac.data(), fac.size() * sizeof(double), cudaMemcpyHostToDevice, p->stream);
    const long long threads = (long long)p->N * dim;
    if (cubic)
        blfik_cubic_spline_coefficients_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, p->stream>>>(p->N, n_knots, h.d_sp_times, h.d_sp_fac,
                                                                                                       d_x, d_bc, h.d_sp_coef, dim);
    else
        blfik_spline_coefficients_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, p->stream>>>(p->N, n_knots, h.d_sp_times, h.d_sp_fac, d_x,
                                                                                                 d_bc, h.d_sp_coef, dim);
    p->launches++;
    cudaError_t e = cudaGetLastError();
    if (stage) {   // the pageable sources (times, fac, positions) are consumed before this returns
        if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
        cudaFree(stage);
    } else if (e == cudaSuccess) {
        e = cudaStreamSynchronize(p->stream);   // times and fac are host vectors that die with this call
    }
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_set_setpoint_spline] %s", cudaGetErrorString(e));
    h.sp_nk = n_knots;
    h.traj_omega = 0.0;   // the sinusoidal generator stands alone; the minimum-jerk one keeps the orientation of an SE3 task
    p->tasks_dirty = true;
    return BLFIK_OK;
}

extern "C" int blfik_set_setpoint_spline(blfik_problem* p, int32_t t, int32_t n_knots, const double* times, const double* positions,
                                         const double* boundary, int32_t mem)
{
    return set_setpoint_spline(p, t, n_knots, times, positions, boundary, mem, false);
}

extern "C" int blfik_set_setpoint_cubic_spline(blfik_problem* p, int32_t t, int32_t n_knots, const double* times, const double* positions,
                                               const double* boundary, int32_t mem)
{
    return set_setpoint_spline(p, t, n_knots, times, positions, boundary, mem, true);
}

extern "C" int blfik_set_setpoint_trajectory(blfik_problem* p, int32_t t, const double* amplitude, double period, const double* phase, int32_t mem)
{
    CHECK_P(p, "blfik_set_setpoint_trajectory");
    CHECK_TASK(p, t, "blfik_set_setpoint_trajectory");
    HostTask& h = p->tasks[t];
    if (h.d.type != BLFIK_TASK_SE3 && h.d.type != BLFIK_TASK_SO3 && h.d.type != BLFIK_TASK_R3 && h.d.type != BLFIK_TASK_COM)
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_trajectory] task %d is not an SE3 / SO3 / R3 / CoM task", t);
    if (!(period > 0)) {   // switch the generator off
        h.traj_omega = 0.0;
        p->tasks_dirty = true;
        return BLFIK_OK;
    }
    if (!amplitude || !phase) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_setpoint_trajectory] null amplitude or phase");
    int rc;
    if (!h.d_traj_phase && (rc = dev_alloc(p, (size_t)p->N, &h.d_traj_phase))) return rc;
    CU(cudaMemcpyAsync(h.d_traj_phase, phase, (size_t)p->N * sizeof(double), kind_in(mem), p->stream));
    for (int c = 0; c < 6; c++) h.traj_amp[c] = amplitude[c];
    h.traj_omega = 6.283185307179586476925286766559 / period;
    h.mj_T = 0.0, h.sp_nk = 0;   // this generator stands alone
    p->tasks_dirty = true;
    return BLFIK_OK;
}

extern "C" int blfik_evaluate_setpoint(blfik_problem* p, int32_t t, double time, double* out, int32_t mem)
{
    CHECK_P(p, "blfik_evaluate_setpoint");
    CHECK_TASK(p, t, "blfik_evaluate_setpoint");
    HostTask& h = p->tasks[t];
    if (!out) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_evaluate_setpoint] null output");
    if (h.sp_len == 0) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_evaluate_setpoint] task %d takes no set point", t);
    if (!h.setpoint_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_evaluate_setpoint] the set point of task %d has not been set", t);
    int rc = refresh_tasks(p);
    if (rc) return rc;
    const size_t len = (size_t)(h.sp_len - h.fb_len) * p->N;
    double* dst = out;
    double* stage = nullptr;
    if (mem != BLFIK_MEM_DEVICE) {
        CU(cudaMalloc((void**)&stage, len * sizeof(double)));
        dst = stage;
    }
    blfik_evaluate_setpoint_kernel<<<(unsigned)((p->N + 127) / 128), 128, 0, p->stream>>>(p->d_tasks, t, p->N, time, dst);
    p->launches++;
    cudaError_t e = cudaGetLastError();
    if (stage) {
        if (e == cudaSuccess) e = cudaMemcpyAsync(out, stage, len * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
        cudaFree(stage);
    }
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_evaluate_setpoint] %s", cudaGetErrorString(e));
    return BLFIK_OK;
}

extern "C" int blfik_set_time(blfik_problem* p, double t)
{
    CHECK_P(p, "blfik_set_time");
    p->time = t;
    return BLFIK_OK;
}

extern "C" double blfik_get_time(const blfik_problem* p) { return p ? p->time : 0.0; }

extern "C" int blfik_set_weight(blfik_problem* p, int32_t t, const double* weight)
{
    CHECK_P(p, "blfik_set_weight");
    CHECK_TASK(p, t, "blfik_set_weight");
    HostTask& h = p->tasks[t];
    // QPInverseKinematics.cpp:251-305: only priority-1 tasks carry a weight
    if (h.d.priority != 1) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_weight] task %d is a hard constraint; only priority-1 tasks have a weight", t);
    if (!weight) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_weight] null weight");
    h.weight.assign(weight, weight + h.rows);
    p->tasks_dirty = true;
    return BLFIK_OK;
}

extern "C" int blfik_set_weight_schedule(blfik_problem* p, int32_t t, const double* target, double settling_time)
{
    CHECK_P(p, "blfik_set_weight_schedule");
    CHECK_TASK(p, t, "blfik_set_weight_schedule");
    HostTask& h = p->tasks[t];
    if (h.d.priority != 1) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_weight_schedule] task %d is a hard constraint; only priority-1 tasks have a weight", t);
    if (!(settling_time > 0.0) || !target) {   // off: the weight stays where the schedule has brought it
        h.sched_a = 0.0, h.sched_target.clear();
        return BLFIK_OK;
    }
    h.sched_target.assign(target, target + h.rows);
    h.sched_a = 3.0 / settling_time;   // FirstOrderSmoother.cpp:73-77
    return BLFIK_OK;
}

extern "C" int blfik_get_weight(blfik_problem* p, int32_t t, double* weight)
{
    CHECK_P(p, "blfik_get_weight");
    CHECK_TASK(p, t, "blfik_get_weight");
    const HostTask& h = p->tasks[t];
    if (h.d.priority != 1 || !weight) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_get_weight] task %d has no weight, or null buffer", t);
    std::copy(h.weight.begin(), h.weight.end(), weight);
    return BLFIK_OK;
}

// One ForwardEuler step of every scheduled weight (FirstOrderSmoother::advance: x += dt (a u - a x)); true if any weight moved.
static bool step_weight_schedules(blfik_problem* p, double dt)
{
    bool any = false;
    for (HostTask& h : p->tasks) {
        if (h.sched_a == 0.0) continue;
        for (size_t r = 0; r < h.weight.size(); r++) h.weight[r] += dt * (h.sched_a * h.sched_target[r] - h.sched_a * h.weight[r]);
        any = true;
    }
    return any;
}

extern "C" int blfik_set_gains(blfik_problem* p, int32_t t, const double* kl, const double* ka, const double* kj)
{
    CHECK_P(p, "blfik_set_gains");
    CHECK_TASK(p, t, "blfik_set_gains");
    HostTask& h = p->tasks[t];
    if (kl) memcpy(h.d.kp_linear, kl, 3 * sizeof(double));
    if (ka) memcpy(h.d.kp_angular, ka, 3 * sizeof(double));
    if (kj) {
        if (h.d.type != BLFIK_TASK_JOINT_TRACKING) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_gains] task %d has no joint gains", t);
        h.kp_joint.assign(kj, kj + p->n);
    }
    p->tasks_dirty = true;
    return BLFIK_OK;
}

extern "C" int blfik_set_limits(blfik_problem* p, int32_t t, const double* upper, const double* lower, const double* klim, double sampling_time)
{
    CHECK_P(p, "blfik_set_limits");
    CHECK_TASK(p, t, "blfik_set_limits");
    HostTask& h = p->tasks[t];
    if (!is_inequality(h.d.type)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_limits] task %d is not a JointLimitsTask / JointVelocityLimitsTask", t);
    const int n = p->n;
    if (upper) h.upper.assign(upper, upper + n);
    if (lower) h.lower.assign(lower, lower + n);
    if (h.d.type == BLFIK_TASK_JOINT_LIMITS) {
        if (klim) {
            for (int j = 0; j < n; j++)
                if (!(klim[j] >= 0 && klim[j] <= 1)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_limits] klim must be in [0, 1]");   // JointLimitsTask.cpp:150-166
            h.klim.assign(klim, klim + n);
        }
        if (sampling_time != 0.0) {
            if (!(sampling_time > 0)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_limits] sampling_time must be positive");
            h.d.sampling_time = sampling_time;
        }
    }
    p->tasks_dirty = true;
    return BLFIK_OK;
}

extern "C" int blfik_set_controller_mode(blfik_problem* p, int32_t t, int32_t enabled)
{
    CHECK_P(p, "blfik_set_controller_mode");
    CHECK_TASK(p, t, "blfik_set_controller_mode");
    p->tasks[t].d.controller_enabled = enabled != 0;
    p->tasks_dirty = true;
    return BLFIK_OK;
}

extern "C" int blfik_debug_enable(blfik_problem* p, int32_t enabled)
{
    CHECK_P(p, "blfik_debug_enable");
    if (enabled && !p->dbg.H) {
        const size_t N = (size_t)p->N, nv = p->nv, me = p->me, n = p->n;
        int rc;
        if ((rc = dev_alloc(p, N * nv * nv, &p->dbg.H)) || (rc = dev_alloc(p, N * nv, &p->dbg.g)) || (rc = dev_alloc(p, N * me * nv, &p->dbg.Ceq)) ||
            (rc = dev_alloc(p, N * me, &p->dbg.beq)) || (rc = dev_alloc(p, N * n, &p->dbg.lo)) || (rc = dev_alloc(p, N * n, &p->dbg.hi)))
            return rc;
    }
    p->debug = enabled != 0;
    return BLFIK_OK;
}

extern "C" int blfik_debug_get(blfik_problem* p, int32_t what, double* out, int32_t mem)
{
    CHECK_P(p, "blfik_debug_get");
    if (!p->dbg.H || !p->ticked) return fail(BLFIK_ERR_NOT_READY, "[blfik_debug_get] call blfik_debug_enable(1) and blfik_tick first");
    if (!out) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_debug_get] null buffer");
    const size_t N = (size_t)p->N, nv = p->nv, me = p->me, n = p->n;
    const double* src = nullptr;
    size_t cnt = 0;
    switch (what) {
    case BLFIK_DEBUG_H: src = p->dbg.H, cnt = N * nv * nv; break;
    case BLFIK_DEBUG_G: src = p->dbg.g, cnt = N * nv; break;
    case BLFIK_DEBUG_CEQ: src = p->dbg.Ceq, cnt = N * me * nv; break;
    case BLFIK_DEBUG_BEQ: src = p->dbg.beq, cnt = N * me; break;
    case BLFIK_DEBUG_LO: src = p->dbg.lo, cnt = N * n; break;
    case BLFIK_DEBUG_HI: src = p->dbg.hi, cnt = N * n; break;
    default: return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_debug_get] unknown selector %d", what);
    }
    if (cnt) CU(cudaMemcpyAsync(out, src, cnt * sizeof(double), kind_out(mem), p->stream));
    if (mem == BLFIK_MEM_HOST) CU(cudaStreamSynchronize(p->stream));
    return BLFIK_OK;
}

// one launch of the tick over robots [first, first + count) on `stream`
static int launch_tick_range(blfik_problem* p, long long first, long long count, int32_t nticks, double dt, uint32_t flags, cudaStream_t stream,
                             const double* packed = nullptr, double* nu_host = nullptr, int* status_host = nullptr, int counter_slot = 0,
                             bool counter_zeroed = false, cudaStream_t asm_stream = nullptr, cudaEvent_t asm_done = nullptr)
{
    TickParams P;
    P.m = p->dm, P.tasks = p->d_tasks, P.ntasks = (int)p->tasks.size(), P.first = first, P.N = first + count;   // robots [first, N)
    P.base_pos = p->base_pos, P.base_quat = p->base_quat, P.s = p->s;
    P.nu = p->nu, P.wx = p->wx, P.wz = p->wz, P.wy = p->wy;
    P.status = p->status, P.iters = p->iters, P.st = p->st, P.nticks = nticks, P.dt = dt, P.flags = flags;
    P.dbg = (p->debug && first == 0 && count == p->N) ? p->dbg : DebugOut{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    P.L = p->L;
    P.packed = packed, P.compact = (packed && (flags & BLFIK_TICK_NO_FEEDFORWARD)) ? 1 : 0;
    P.packed_len = P.compact ? p->packed_len_c : p->packed_len;
    P.nu_host = nu_host, P.status_host = status_host;
    P.t0 = p->time;
    if (p->solve_kernel && p->path == 0 && !P.dbg.H && !p->streaming_launch) {
        // two launches per tick (blfik_split.cuh): K1 + K2 one thread per robot, K3 + K4 one warp per robot
        if (!p->d_qp) {
            int rc = dev_alloc(p, (size_t)p->N * p->rec_len, &p->d_qp, false);
            if (rc) return rc;
        }
        AsmParams A;
        SolveParams S;
        A.T = P, A.qp = p->d_qp, A.rec_len = p->rec_len, A.K = p->K, A.off_c = p->rec_off_c, A.off_v = p->rec_off_v;
        A.model_pack = p->d_model_pack, A.pack_len = p->pack_len;
        A.sh_jump = p->F.sh_jump, A.sh_lod = p->F.sh_lod, A.sh_anc = p->F.sh_anc, A.sh_msub = p->F.sh_msub;
        A.events = p->d_asm_events, A.nevents = p->asm_nevents;
        A.work_counter = S.work_counter = p->work_counter + counter_slot;   // launches that may overlap use different slots
        S.T = P, S.qp = p->d_qp, S.prefetch_ahead = p->num_sms * 16, S.nv = p->nv, S.K = p->K;
        auto fill_weights = [&]() {
            for (int r = 0; r < 32; r++) S.evec[r] = 1.0;
            A.jw_task = -1;
            for (size_t t = 0; t < p->tasks.size(); t++) {
                const HostTask& h = p->tasks[t];
                if (is_dense(h.d.type))
                    for (int r = 0; r < h.rows; r++) S.evec[h.brow + r] = h.d.priority == 0 ? 0.0 : 1.0 / h.weight[r];
                else if (h.d.type == BLFIK_TASK_JOINT_TRACKING && h.d.priority == 1 && h.sched_a != 0.0) {
                    // a scheduled joint-space weight travels with the launch (the device copy of the task keeps the old one
                    // until the next refresh); one such task per problem
                    A.jw_task = (int)t;
                    for (int j = 0; j < 32; j++) A.jw[j] = j < (int)h.weight.size() ? h.weight[j] : 0.0;
                }
            }
        };
        fill_weights();
        S.max_as_iters = p->max_as_iters, S.max_polish_iters = p->max_polish_iters, S.max_total_iters = p->max_total_iters;
        { const char* e = getenv("BLFIK_PDAS_FIRST"); S.pdas_first = e ? atoi(e) : 1; }
        { const char* e = getenv("BLFIK_MAX_ITERS"); if (e) S.max_total_iters = atoi(e); }
        { const char* e = getenv("BLFIK_MAX_AS"); if (e) S.max_as_iters = atoi(e); }
        { const char* e = getenv("BLFIK_ADMM_STATUS"); S.admm_status = e ? atoi(e) : 0; }
        const long long swant = (count + p->solve_wpb - 1) / p->solve_wpb, scap = (long long)p->num_sms * p->solve_blocks;
        const int agrid = (int)((count + p->asm_threads - 1) / p->asm_threads), sgrid = (int)(swant < scap ? swant : scap);
        static const int lpt_env = [] { const char* e = getenv("BLFIK_LPT"); return e ? atoi(e) : 1; }();
        const bool lpt = lpt_env && first == 0 && count == p->N && !asm_stream && p->N < 0x7fffff00LL;   // whole-batch launches only
        if (lpt && !p->d_slow) {
            const int rc = dev_alloc(p, (size_t)5 * p->N + 2, &p->d_slow);   // zero-filled: no stamp matches
            if (rc) return rc;
        }
        static const int bins_env = [] { const char* e = getenv("BLFIK_BINS"); return e ? atoi(e) : 1; }();
        const bool bins = lpt && bins_env;
        if (bins && !p->d_bins) {
            const int rc = dev_alloc(p, (size_t)64 * p->N + 64, &p->d_bins, false);
            if (rc) return rc;
        }
        A.bin_cap = S.bin_cap = p->N;
        S.slow_threshold = 16;
        static const int short_env = [] { const char* e = getenv("BLFIK_SHORT_BUDGET"); return e ? atoi(e) : 48; }();
        S.unsolved_run = lpt ? p->d_slow + 4 * p->N + 2 : nullptr, S.short_budget = short_env;
        for (int tick = 0; tick < nticks; tick++) {
            A.tick = S.tick = tick;
            A.tnow = p->time + tick * dt;
            {
                const int T = ++p->solve_stamp, par = T & 1;
                int *lists = p->d_slow, *stamps = p->d_slow + 2 * p->N, *lens = p->d_slow + 4 * p->N;
                const bool in_valid = lpt && p->slow_list_for == T;
                S.stamp = T;
                S.slow_in = in_valid ? lists + (size_t)par * p->N : nullptr, S.slow_in_n = in_valid ? lens + par : nullptr;
                S.slow_stamp_in = in_valid ? stamps + (size_t)par * p->N : nullptr;
                S.slow_out = lpt ? lists + (size_t)(par ^ 1) * p->N : nullptr, S.slow_out_n = lpt ? lens + (par ^ 1) : nullptr;
                S.slow_stamp_out = lpt ? stamps + (size_t)(par ^ 1) * p->N : nullptr;
                A.slow_out_n = S.slow_out_n;
                p->slow_list_for = lpt ? T + 1 : -1;
                // the hardness bins order a tick that has no history to go by (a cold tick, the first tick of a rollout); once
                // the previous tick's slow robots are known they lead, and the rest goes by index
                const bool use_bins = bins && !in_valid;
                A.bin_cnt = use_bins ? p->d_bins : nullptr, A.bin_list = use_bins ? p->d_bins + 64 : nullptr;
                if (use_bins) CU(cudaMemsetAsync(p->d_bins, 0, 64 * sizeof(int), stream));
                // a chunk of the pipelined host tick: its own counters (by counter slot) and its own stretch of the lists
                if (bins_env && !lpt && asm_stream && asm_done && nticks == 1 && p->N < 0x7fffff00LL / 64) {
                    if (!p->d_bins_chunk) {
                        const int rc = dev_alloc(p, (size_t)64 * kCounterSlots + (size_t)64 * p->N, &p->d_bins_chunk, false);
                        if (rc) return rc;
                    }
                    A.bin_cnt = p->d_bins_chunk + 64 * counter_slot, A.bin_list = p->d_bins_chunk + 64 * kCounterSlots + 64 * first, A.bin_cap = count;
                    S.bin_cap = count;
                    CU(cudaMemsetAsync(A.bin_cnt, 0, 64 * sizeof(int), asm_stream));
                }
                S.bin_cnt = A.bin_cnt, S.bin_list = A.bin_list;
            }
            if (step_weight_schedules(p, dt)) {   // MultiStateWeightProvider::advance() of every period, then the solver reads the weight
                fill_weights();
                p->tasks_dirty = true;            // the device copies of the weights follow at the next refresh
            }
            const bool timed = p->timing && tick == nticks - 1 && count == p->N;
            if (timed) CU(cudaEventRecord(p->ev_t[0], stream));
            if (asm_stream && asm_done) {
                // Host-tick pipeline: the assemble kernel of this chunk runs on a high-priority stream, so that its few blocks
                // are scheduled ahead of the solve blocks of the previous chunks (its duration is the latency of one thread's
                // walk, whatever the chunk size: it must overlap the other chunks' work, not queue behind it)
                p->asm_kernel<<<agrid, p->asm_threads, p->asm_smem, asm_stream>>>(A);
                CU(cudaEventRecord(asm_done, asm_stream));
                CU(cudaStreamWaitEvent(stream, asm_done, 0));
            } else {
                p->asm_kernel<<<agrid, p->asm_threads, p->asm_smem, stream>>>(A);
            }
            if (timed) CU(cudaEventRecord(p->ev_t[1], stream));
            p->solve_kernel<<<sgrid, p->solve_wpb * 32, p->solve_smem, stream>>>(S);
            if (timed) CU(cudaEventRecord(p->ev_t[2], stream));
            p->launches += 2;
        }
        CU(cudaGetLastError());
        return BLFIK_OK;
    }
    {
        bool scheduled = false;
        for (const HostTask& h : p->tasks) scheduled |= h.sched_a != 0.0;
        if (scheduled) {
            // A single kernel cannot change a weight between its ticks (only the two-launch path above, whose ticks are
            // separate launches, can): one tick per call here, the stepped weight uploaded before it.
            if (nticks > 1)
                return fail(BLFIK_ERR_UNSUPPORTED, "[blfik_tick] a weight schedule inside a multi-tick launch needs the two-launch kernel path "
                                                   "(nv <= 32, <= 32 dense rows, kernel path 0); tick once per call instead");
            step_weight_schedules(p, dt);
            p->tasks_dirty = true;
            const int rc = refresh_tasks(p);
            if (rc) return rc;
        }
    }
    if (p->fast_kernel && (p->path == 0 || p->path == 2)) {
        FastParams FP;
        FP.T = P, FP.F = p->F, FP.nv = p->nv, FP.K = p->K;
        FP.max_as_iters = p->max_as_iters, FP.max_polish_iters = p->max_polish_iters, FP.max_total_iters = p->max_total_iters;
        FP.model_pack = p->d_model_pack, FP.njump = p->njump, FP.has_com = p->has_com;
        { const char* e = getenv("BLFIK_SYNC_PHASES"); FP.sync_phases = e ? atoi(e) : 2; }
        { const char* e = getenv("BLFIK_PDAS_FIRST"); FP.pdas_first = e ? atoi(e) : 1; }
        { const char* e = getenv("BLFIK_MAX_ITERS"); if (e) FP.max_total_iters = atoi(e); }
        { const char* e = getenv("BLFIK_MAX_AS"); if (e) FP.max_as_iters = atoi(e); }
        const long long want = (count + p->fast_wpb - 1) / p->fast_wpb;
        static const int dynamic_env = [] { const char* e = getenv("BLFIK_DYNAMIC"); return e ? atoi(e) : 1; }();
        FP.work_counter = nullptr;
        if (dynamic_env && p->work_counter) {
            FP.work_counter = p->work_counter + counter_slot;   // launches that may overlap use different slots
            if (!counter_zeroed) CU(cudaMemsetAsync(FP.work_counter, 0, sizeof(int), stream));
        }
        FP.arrived = nullptr, FP.abort_flag = nullptr, FP.abort_host = nullptr, FP.arrive_timeout = 0;
        if (p->streaming_launch && FP.work_counter) {
            FP.arrived = p->d_arrived, FP.abort_flag = p->d_arrived + 1;
            CU(cudaHostGetDevicePointer((void**)&FP.abort_host, p->h_stream_ctl + kMaxPieces, 0));
            static const long long timeout_env = [] { const char* e = getenv("BLFIK_ARRIVE_TIMEOUT"); return e ? atoll(e) : 500000000LL; }();
            FP.arrive_timeout = timeout_env;    // ~0.25 s of SM clocks (the whole transfer takes ~1 ms): copies and kernels are not overlapping
        }
        p->fast_kernel<<<(int)(want < p->fast_grid ? want : p->fast_grid), p->fast_wpb * 32, p->fast_smem, stream>>>(FP);
    } else if (p->eq_kernel && p->path != 1 && !P.dbg.H) {   // the debug read-back wants H: the generic kernel forms it
        P.L = p->Leq;
        const long long want = (count + p->eq_wpb - 1) / p->eq_wpb;
        p->eq_func<<<(int)(want < p->eq_grid ? want : p->eq_grid), p->eq_wpb * 32, p->eq_smem, stream>>>(P);
    } else {
        const long long want = (count + p->warps_per_block - 1) / p->warps_per_block;
        blfik_tick_kernel<<<(int)(want < p->grid ? want : p->grid), p->warps_per_block * 32, p->smem_bytes, stream>>>(P);
    }
    CU(cudaGetLastError());
    p->launches++;
    return BLFIK_OK;
}

static int check_tick_args(blfik_problem* p, int32_t nticks, double dt, uint32_t flags, const char* fn, bool need_inputs)
{
    // the reference's advance() fails when a set-point was never set (SE3Task.cpp:245-249) and kinDyn must hold a state
    if (need_inputs) {
        if (!p->state_set) return fail(BLFIK_ERR_NOT_READY, "[%s] the robot state has never been set", fn);
        for (size_t t = 0; t < p->tasks.size(); t++)
            if (!p->tasks[t].setpoint_set) return fail(BLFIK_ERR_NOT_READY, "[%s] the set point of task %d has not been set at least once", fn, (int)t);
    }
    if (nticks < 1) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[%s] nticks must be >= 1", fn);
    if ((flags & BLFIK_TICK_INTEGRATE) && !(dt > 0)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[%s] dt must be positive when integrating", fn);
    return BLFIK_OK;
}

extern "C" int blfik_tick(blfik_problem* p, int32_t nticks, double dt, uint32_t flags)
{
    CHECK_P(p, "blfik_tick");
    int rc = check_tick_args(p, nticks, dt, flags, "blfik_tick", true);
    if (rc) return rc;
    if ((rc = refresh_tasks(p))) return rc;
    if ((rc = launch_tick_range(p, 0, p->N, nticks, dt, flags, p->stream))) return rc;
    if (flags & BLFIK_TICK_INTEGRATE) p->time += nticks * dt;
    p->ticked = true;
    return BLFIK_OK;
}

extern "C" int blfik_tick_host(blfik_problem* p, const double* base_pos, const double* base_quat, const double* s,
                               const double* const* setpoints, double* nu, int32_t* status, double dt, uint32_t flags, int32_t nchunks)
{
    CHECK_P(p, "blfik_tick_host");
    int rc = check_tick_args(p, 1, dt, flags, "blfik_tick_host", false);
    if (rc) return rc;
    if (!base_pos || !base_quat || !s || !nu) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_tick_host] null state or output buffer");
    for (size_t t = 0; t < p->tasks.size(); t++)
        if (p->tasks[t].sp_len && !(setpoints && setpoints[t]) && !p->tasks[t].setpoint_set)
            return fail(BLFIK_ERR_NOT_READY, "[blfik_tick_host] the set point of task %d has not been set at least once", (int)t);
    if (nchunks < 1) nchunks = 1;
    if (nchunks > kMaxChunks) nchunks = kMaxChunks;
    if ((long long)nchunks > p->N) nchunks = (int)p->N;
    if ((rc = refresh_tasks(p))) return rc;
    if (!p->h2d_stream) {
        CU(cudaStreamCreateWithFlags(&p->h2d_stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&p->d2h_stream, cudaStreamNonBlocking));
        for (int c = 0; c < kMaxChunks; c++) {
            CU(cudaEventCreateWithFlags(&p->ev_in[c], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&p->ev_done[c], cudaEventDisableTiming));
        }
        CU(cudaEventCreateWithFlags(&p->ev_start, cudaEventDisableTiming));
    }
    // everything enqueued earlier on the problem's stream comes first
    CU(cudaEventRecord(p->ev_start, p->stream));
    CU(cudaStreamWaitEvent(p->h2d_stream, p->ev_start, 0));
    CU(cudaStreamWaitEvent(p->d2h_stream, p->ev_start, 0));
    const int n = p->n, nv = p->nv;
    for (int c = 0; c < nchunks; c++) {
        const long long a = p->N * c / nchunks, b = p->N * (c + 1) / nchunks, cnt = b - a;
        // chunk c: host -> device on the copy-in stream ...
        CU(cudaMemcpyAsync(p->base_pos + 3 * a, base_pos + 3 * a, (size_t)cnt * 3 * sizeof(double), cudaMemcpyHostToDevice, p->h2d_stream));
        CU(cudaMemcpyAsync(p->base_quat + 4 * a, base_quat + 4 * a, (size_t)cnt * 4 * sizeof(double), cudaMemcpyHostToDevice, p->h2d_stream));
        CU(cudaMemcpyAsync(p->s + (size_t)n * a, s + (size_t)n * a, (size_t)cnt * n * sizeof(double), cudaMemcpyHostToDevice, p->h2d_stream));
        for (size_t t = 0; t < p->tasks.size(); t++) {
            HostTask& h = p->tasks[t];
            if (h.sp_len && setpoints && setpoints[t])
                CU(cudaMemcpyAsync(h.d_setpoint + (size_t)h.sp_len * a, setpoints[t] + (size_t)h.sp_len * a, (size_t)cnt * h.sp_len * sizeof(double),
                                   cudaMemcpyHostToDevice, p->h2d_stream));
        }
        CU(cudaEventRecord(p->ev_in[c], p->h2d_stream));
        // ... the tick on the compute stream ...
        CU(cudaStreamWaitEvent(p->stream, p->ev_in[c], 0));
        if ((rc = launch_tick_range(p, a, cnt, 1, dt, flags, p->stream))) return rc;
        CU(cudaEventRecord(p->ev_done[c], p->stream));
        // ... device -> host on the copy-out stream
        CU(cudaStreamWaitEvent(p->d2h_stream, p->ev_done[c], 0));
        CU(cudaMemcpyAsync(nu + (size_t)nv * a, p->nu + (size_t)nv * a, (size_t)cnt * nv * sizeof(double), cudaMemcpyDeviceToHost, p->d2h_stream));
        if (status) CU(cudaMemcpyAsync(status + a, p->status + a, (size_t)cnt * sizeof(int32_t), cudaMemcpyDeviceToHost, p->d2h_stream));
    }
    for (size_t t = 0; t < p->tasks.size(); t++)
        if (p->tasks[t].sp_len && setpoints && setpoints[t]) p->tasks[t].setpoint_set = true;
    p->state_set = true, p->ticked = true;
    if (flags & BLFIK_TICK_INTEGRATE) p->time += dt;
    CU(cudaStreamSynchronize(p->d2h_stream));
    CU(cudaStreamSynchronize(p->stream));
    return BLFIK_OK;
}

extern "C" int32_t blfik_packed_record_len(const blfik_problem* p) { return p ? p->packed_len : -1; }

extern "C" int32_t blfik_packed_record_len_ex(const blfik_problem* p, uint32_t flags)
{
    if (!p) return -1;
    return (flags & BLFIK_TICK_NO_FEEDFORWARD) ? p->packed_len_c : p->packed_len;
}

extern "C" int32_t blfik_packed_offset_ex(const blfik_problem* p, int32_t task, uint32_t flags)
{
    if (!p) return -1;
    if (task < 0) return 0;
    if (task >= (int)p->tasks.size() || p->tasks[task].sp_len == 0 || p->tasks[task].shared) return -1;
    return (flags & BLFIK_TICK_NO_FEEDFORWARD) ? p->tasks[task].pk_off_c : p->tasks[task].pk_off;
}

extern "C" int32_t blfik_packed_part_len(const blfik_problem* p, int32_t task, uint32_t flags)
{
    if (!p || task < 0 || task >= (int)p->tasks.size()) return -1;
    const HostTask& h = p->tasks[task];
    if (h.sp_len == 0 || h.shared) return 0;
    return (flags & BLFIK_TICK_NO_FEEDFORWARD) ? h.base_c + h.fb_len : h.sp_len;
}

extern "C" int32_t blfik_packed_offset(const blfik_problem* p, int32_t task)
{
    if (!p) return -1;
    if (task < 0) return 0;   // the state: p(3) q(4) s(n)
    if (task >= (int)p->tasks.size() || p->tasks[task].sp_len == 0 || p->tasks[task].shared) return -1;
    return p->tasks[task].pk_off;
}

extern "C" int blfik_tick_packed(blfik_problem* p, const double* records, double* nu, int32_t* status, double dt, uint32_t flags, int32_t nchunks)
{
    CHECK_P(p, "blfik_tick_packed");
    int rc = check_tick_args(p, 1, dt, flags, "blfik_tick_packed", false);
    if (rc) return rc;
    if (!records || !nu) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_tick_packed] null input or output buffer");
    if ((flags & BLFIK_TICK_KEEP_STATE) && !p->state_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_tick_packed] BLFIK_TICK_KEEP_STATE: the device holds no state yet");
    if (nchunks <= 0) {
        // zero-copy: when every host buffer is page-locked and mapped into the device's address space (cudaHostAlloc /
        // cudaHostRegister / torch pin_memory), the kernel reads the records and writes nu / status over PCIe itself,
        // so the transfers overlap the computation robot by robot instead of chunk by chunk
        const void* hp[3] = {records, nu, status};
        void* dp[3] = {nullptr, nullptr, nullptr};
        bool mapped = true;
        for (int k = 0; k < 3 && mapped; k++) {
            if (!hp[k]) continue;
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, hp[k]) != cudaSuccess || at.type != cudaMemoryTypeHost || !at.devicePointer) {
                cudaGetLastError();
                mapped = false;
            } else {
                dp[k] = at.devicePointer;
            }
        }
        if (mapped) {
            if ((rc = refresh_tasks(p))) return rc;
            if ((rc = launch_tick_range(p, 0, p->N, 1, dt, flags, p->stream, static_cast<const double*>(dp[0]), static_cast<double*>(dp[1]),
                                        static_cast<int*>(dp[2]))))
                return rc;
            p->state_set = true, p->ticked = true;
            if (flags & BLFIK_TICK_INTEGRATE) p->time += dt;
            CU(cudaStreamSynchronize(p->stream));
            return BLFIK_OK;
        }
        nchunks = 2;   // pageable buffers: staged copies
    }
    if (nchunks > kMaxChunks) nchunks = kMaxChunks;
    if ((long long)nchunks > p->N) nchunks = (int)p->N;
    if ((rc = refresh_tasks(p))) return rc;
    if (!p->d_packed && (rc = dev_alloc(p, (size_t)p->N * p->packed_len_max, &p->d_packed, false))) return rc;
    if (!p->h2d_stream) {
        CU(cudaStreamCreateWithFlags(&p->h2d_stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&p->d2h_stream, cudaStreamNonBlocking));
        for (int c = 0; c < kMaxChunks; c++) {
            CU(cudaEventCreateWithFlags(&p->ev_in[c], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&p->ev_done[c], cudaEventDisableTiming));
        }
        CU(cudaEventCreateWithFlags(&p->ev_start, cudaEventDisableTiming));
    }
    if (p->work_counter) CU(cudaMemsetAsync(p->work_counter, 0, kCounterSlots * sizeof(int), p->stream));   // one reset for all chunk launches
    CU(cudaEventRecord(p->ev_start, p->stream));
    CU(cudaStreamWaitEvent(p->h2d_stream, p->ev_start, 0));
    CU(cudaStreamWaitEvent(p->d2h_stream, p->ev_start, 0));
    const int nv = p->nv, R = (flags & BLFIK_TICK_NO_FEEDFORWARD) ? p->packed_len_c : p->packed_len;
    // ---- streaming input: ONE persistent launch that starts at once and consumes the records as the copy engine delivers
    // them (chunk launches lose a tail each: 5 launches of 13k robots take 2.3 ms of kernel time against 1.36 ms for one
    // launch of 65k).  Needs the register-resident kernel with dynamic hand-out, page-locked records (a pure DMA on the
    // copy stream) and mapped page-locked outputs.
    static const int streaming_env = [] { const char* e = getenv("BLFIK_STREAMING"); return e ? atoi(e) : 1; }();
    if (streaming_env && !p->streaming_off && p->fast_kernel && (p->path == 2 || (p->path == 0 && !p->solve_kernel)) && p->work_counter && p->N < 0x7fffff00LL) {
        cudaPointerAttributes at;
        double* nu_dev = nullptr;
        int* st_dev = nullptr;
        bool ok = cudaPointerGetAttributes(&at, records) == cudaSuccess && at.type == cudaMemoryTypeHost;
        if (ok) ok = cudaPointerGetAttributes(&at, nu) == cudaSuccess && at.type == cudaMemoryTypeHost && (nu_dev = static_cast<double*>(at.devicePointer));
        if (ok && status) ok = cudaPointerGetAttributes(&at, status) == cudaSuccess && at.type == cudaMemoryTypeHost && (st_dev = static_cast<int*>(at.devicePointer));
        cudaGetLastError();
        static const int dyn = [] { const char* e = getenv("BLFIK_DYNAMIC"); return e ? atoi(e) : 1; }();
        if (ok && dyn) {
            if (!p->h_stream_ctl) {
                CU(cudaHostAlloc((void**)&p->h_stream_ctl, (kMaxPieces + 1) * sizeof(int), cudaHostAllocMapped));
                if ((rc = dev_alloc(p, 2, &p->d_arrived))) return rc;
            }
            // the arrival word is raised by a stream memory operation when the driver offers one (looked up through the runtime,
            // libcuda is not linked), else by a 4-byte copy
            typedef int (*write_value_fn)(cudaStream_t, unsigned long long, unsigned int, unsigned int);
            static const write_value_fn write_value = [] {
                const char* e = getenv("BLFIK_WRITE_VALUE");
                if (e && atoi(e) == 0) return (write_value_fn) nullptr;
                void* fn = nullptr;
                cudaDriverEntryPointQueryResult qr;
                if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &fn, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
                    cudaGetLastError();
                    fn = nullptr;
                }
                return (write_value_fn)fn;
            }();
            // pieces of about 2,048 robots (2 MB for C3: 40 us on the link), the first ones smaller so that work starts early
            static const int ramp_env = [] { const char* e = getenv("BLFIK_RAMP"); return e ? atoi(e) : 0; }();
            static const int pieces_env = [] { const char* e = getenv("BLFIK_PIECES"); return e ? atoi(e) : 0; }();
            int npieces = pieces_env > 0 ? pieces_env : 16;   // measured best on C3 (8: 40.7, 12: 41.9, 16: 42.1, 24: 40.8 M ticks/s)
            if (npieces > kMaxPieces) npieces = kMaxPieces;
            if ((long long)npieces > p->N) npieces = (int)p->N;
            p->h_stream_ctl[kMaxPieces] = 0;
            CU(cudaMemsetAsync(p->d_arrived, 0, 2 * sizeof(int), p->stream));
            CU(cudaEventRecord(p->ev_start, p->stream));
            CU(cudaStreamWaitEvent(p->h2d_stream, p->ev_start, 0));
            p->streaming_launch = true;
            rc = launch_tick_range(p, 0, p->N, 1, dt, flags, p->stream, p->d_packed, nu_dev, st_dev, 0);
            p->streaming_launch = false;
            if (rc) return rc;
            long long a = 0;
            for (int c = 0; c < npieces; c++) {
                // piece c ends at N * smoothstep((c + 1) / npieces): small pieces first (work starts early) and last (little is
                // left to compute when the final piece lands), large ones in between (fewer stream operations)
                const double f = (double)(c + 1) / npieces;
                long long b = c + 1 == npieces ? p->N : (long long)((double)p->N * (ramp_env ? f * sqrt(f) : f * f * (3.0 - 2.0 * f)));
                // A piece ends where a record starts on a 128-byte line (every 16 robots whatever the record length): the
                // kernel never reads a line that a later piece still has to fill, so no stale copy of it can sit in an L1.
                b = (b + 15) & ~15LL;
                if (b <= a) b = a + 16;
                if (b > p->N) b = p->N;
                CU(cudaMemcpyAsync(p->d_packed + (size_t)R * a, records + (size_t)R * a, (size_t)(b - a) * R * sizeof(double), cudaMemcpyHostToDevice, p->h2d_stream));
                p->h_stream_ctl[c] = (int)b;
                static const int stall_env = [] { const char* e = getenv("BLFIK_TEST_STALL"); return e ? atoi(e) : 0; }();
                if (stall_env && b >= p->N) {
                    // test hook: the last piece is never announced -> the kernel must give up and the call must fail cleanly
                } else if (write_value) {   // stream memory operation: a 32-bit store ordered after the copy, no DMA set-up
                    if (write_value(p->h2d_stream, (unsigned long long)(uintptr_t)p->d_arrived, (unsigned int)b, 0u) != 0)
                        return fail(BLFIK_ERR_CUDA, "[blfik_tick_packed] cuStreamWriteValue32 failed");
                } else {
                    CU(cudaMemcpyAsync(p->d_arrived, p->h_stream_ctl + c, sizeof(int), cudaMemcpyHostToDevice, p->h2d_stream));
                }
                a = b;
                if (a >= p->N) break;
            }
            p->state_set = true, p->ticked = true;
            if (flags & BLFIK_TICK_INTEGRATE) p->time += dt;
            CU(cudaStreamSynchronize(p->h2d_stream));
            CU(cudaStreamSynchronize(p->stream));
            if (!p->h_stream_ctl[kMaxPieces]) return BLFIK_OK;
            // The kernel gave up waiting: copies and kernels do not overlap here (a profiler or a debugger serialises them, or
            // the copy stream is stalled).  Nothing is lost -- a packed tick rebuilds the state from the records -- so this
            // call and the later ones of this problem use chunk launches, which only need stream order.
            static const int strict_env = [] { const char* e = getenv("BLFIK_TEST_STALL"); return e ? atoi(e) : 0; }();
            if (strict_env) return fail(BLFIK_ERR_CUDA, "[blfik_tick_packed] the kernel gave up waiting for the input records (copy stream stalled)");
            p->streaming_off = true;
            if (flags & BLFIK_TICK_INTEGRATE) p->time -= dt;
            // the aborted launch has drawn from counter slot 0 (every team that gave up still added its share): the chunk
            // launches below assume zeroed counters
            if (p->work_counter) CU(cudaMemsetAsync(p->work_counter, 0, kCounterSlots * sizeof(int), p->stream));
            CU(cudaEventRecord(p->ev_start, p->stream));
            CU(cudaStreamWaitEvent(p->h2d_stream, p->ev_start, 0));
            CU(cudaStreamWaitEvent(p->d2h_stream, p->ev_start, 0));
        }
    }
    // results: when nu (and status) are mapped page-locked buffers the kernels write them straight to the host (posted
    // PCIe writes, no latency to hide), otherwise they are copied back chunk by chunk
    double* nu_map = nullptr;
    int* status_map = nullptr;
    static const int host_write = [] { const char* e = getenv("BLFIK_HOST_WRITE"); return e ? atoi(e) : 1; }();
    if (host_write) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, nu) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) nu_map = static_cast<double*>(at.devicePointer);
        else cudaGetLastError();
        if (status && nu_map) {
            if (cudaPointerGetAttributes(&at, status) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) status_map = static_cast<int*>(at.devicePointer);
            else cudaGetLastError(), nu_map = nullptr;
        }
    }
    // solve kernels of consecutive chunks alternate over these streams; measured on C3 (two-launch path, 8 chunks): 38.5 M
    // ticks/s with 2 streams, 47.7 M with 4 (a chunk's solve must not queue behind an earlier chunk's assemble latency)
    static const int nstreams_env = [] { const char* e = getenv("BLFIK_CHUNK_STREAMS"); int v = e ? atoi(e) : 0; return v < 0 ? 0 : (v > 4 ? 4 : v); }();
    const int nstreams = nstreams_env ? nstreams_env : ((p->solve_kernel && p->path == 0) ? 4 : 2);
    for (int k = 0; k < nstreams; k++) {
        if (!p->chunk_stream[k]) CU(cudaStreamCreateWithFlags(&p->chunk_stream[k], cudaStreamNonBlocking));
        CU(cudaStreamWaitEvent(p->chunk_stream[k], p->ev_start, 0));
    }
    // BLFIK_TRACE=1: print the timeline of this call (copy arrival, kernel start / end per chunk) -- a debugging aid
    static const int trace = [] { const char* e = getenv("BLFIK_TRACE"); return e ? atoi(e) : 0; }();
    cudaEvent_t tr_start = nullptr, tr_in[kMaxChunks], tr_k0[kMaxChunks], tr_k1[kMaxChunks];
    if (trace) {
        CU(cudaEventCreate(&tr_start));
        for (int c = 0; c < nchunks; c++) { CU(cudaEventCreate(&tr_in[c])); CU(cudaEventCreate(&tr_k0[c])); CU(cudaEventCreate(&tr_k1[c])); }
        CU(cudaEventRecord(tr_start, p->h2d_stream));
    }
    const bool split = p->solve_kernel && p->path == 0 && !p->debug;
    if (split && !p->asm_stream[0]) {
        int lo_pri = 0, hi_pri = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo_pri, &hi_pri));
        for (int k = 0; k < 4; k++) CU(cudaStreamCreateWithPriority(&p->asm_stream[k], cudaStreamNonBlocking, hi_pri));
        for (int c = 0; c < kMaxChunks; c++) CU(cudaEventCreateWithFlags(&p->ev_asm[c], cudaEventDisableTiming));
    }
    // The first chunk is half the size of the others: nothing can run until it has arrived.  On the two-launch path the last
    // chunks shrink as well (what runs after the final copy is one assemble latency plus the last chunk's solve).
    auto bound = [&](int c) {
        if (c <= 0) return 0LL;
        if (c >= nchunks) return (long long)p->N;
        if (split && nchunks >= 4) {
            const double f = (double)c / nchunks;
            return (long long)((double)p->N * f * f * (3.0 - 2.0 * f)) & ~15LL;   // smoothstep, on multiples of 16 robots
        }
        return p->N * (2 * c - 1) / (2 * nchunks - 1);
    };
    for (int c = 0; c < nchunks; c++) {
        const long long a = bound(c), b = bound(c + 1), cnt = b - a;
        if (cnt <= 0) continue;
        cudaStream_t cs = p->chunk_stream[c % nstreams];
        CU(cudaMemcpyAsync(p->d_packed + (size_t)R * a, records + (size_t)R * a, (size_t)cnt * R * sizeof(double), cudaMemcpyHostToDevice, p->h2d_stream));
        CU(cudaEventRecord(p->ev_in[c], p->h2d_stream));
        CU(cudaStreamWaitEvent(cs, p->ev_in[c], 0));
        if (trace) CU(cudaEventRecord(tr_in[c], p->h2d_stream));
        if (trace) CU(cudaEventRecord(tr_k0[c], cs));
        if (split) CU(cudaStreamWaitEvent(p->asm_stream[c % 4], p->ev_in[c], 0));
        if ((rc = launch_tick_range(p, a, cnt, 1, dt, flags, cs, p->d_packed, nu_map, status_map, c % kCounterSlots, true,
                                    split ? p->asm_stream[c % 4] : nullptr, split ? p->ev_asm[c] : nullptr)))
            return rc;
        if (trace) CU(cudaEventRecord(tr_k1[c], cs));
        CU(cudaEventRecord(p->ev_done[c], cs));
        CU(cudaStreamWaitEvent(p->stream, p->ev_done[c], 0));   // later work on the problem's stream is ordered after every chunk
        if (!nu_map) {
            CU(cudaStreamWaitEvent(p->d2h_stream, p->ev_done[c], 0));
            CU(cudaMemcpyAsync(nu + (size_t)nv * a, p->nu + (size_t)nv * a, (size_t)cnt * nv * sizeof(double), cudaMemcpyDeviceToHost, p->d2h_stream));
            if (status) CU(cudaMemcpyAsync(status + a, p->status + a, (size_t)cnt * sizeof(int32_t), cudaMemcpyDeviceToHost, p->d2h_stream));
        }
    }
    p->state_set = true, p->ticked = true;
    if (flags & BLFIK_TICK_INTEGRATE) p->time += dt;
    CU(cudaStreamSynchronize(p->d2h_stream));
    CU(cudaStreamSynchronize(p->stream));
    if (trace) {
        for (int c = 0; c < nchunks; c++) {
            float a = 0, b = 0, d = 0;
            cudaEventElapsedTime(&a, tr_start, tr_in[c]), cudaEventElapsedTime(&b, tr_start, tr_k0[c]), cudaEventElapsedTime(&d, tr_start, tr_k1[c]);
            fprintf(stderr, "[blfik trace] chunk %d: %lld robots, copy done %.3f ms, kernel %.3f -> %.3f ms\n", c, (long long)(bound(c + 1) - bound(c)), a, b, d);
            cudaEventDestroy(tr_in[c]), cudaEventDestroy(tr_k0[c]), cudaEventDestroy(tr_k1[c]);
        }
        cudaEventDestroy(tr_start);
    }
    return BLFIK_OK;
}

extern "C" int blfik_set_kernel_path(blfik_problem* p, int32_t path)
{
    CHECK_P(p, "blfik_set_kernel_path");
    if (path < 0 || path > 2) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_set_kernel_path] path must be 0 (auto), 1 (generic kernel) or 2 (fused register-resident kernel)");
    p->path = path;
    return BLFIK_OK;
}

extern "C" int32_t blfik_get_kernel_path(const blfik_problem* p)
{
    if (!p) return -1;
    if (p->path == 1) return 1;
    if (p->solve_kernel && p->path == 0) return 4;
    return p->fast_kernel ? 2 : (p->eq_kernel ? 3 : 1);
}

extern "C" int blfik_get_output(blfik_problem* p, double* nu, int32_t mem)
{
    CHECK_P(p, "blfik_get_output");
    if (!p->ticked) return fail(BLFIK_ERR_NOT_READY, "[blfik_get_output] no tick has run yet");
    if (!nu) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_get_output] null buffer");
    CU(cudaMemcpyAsync(nu, p->nu, (size_t)p->N * p->nv * sizeof(double), kind_out(mem), p->stream));
    if (mem == BLFIK_MEM_HOST) CU(cudaStreamSynchronize(p->stream));
    return BLFIK_OK;
}

extern "C" int blfik_get_status(blfik_problem* p, int32_t* status, int32_t* iters, int32_t mem)
{
    CHECK_P(p, "blfik_get_status");
    if (!p->ticked) return fail(BLFIK_ERR_NOT_READY, "[blfik_get_status] no tick has run yet");
    if (status) CU(cudaMemcpyAsync(status, p->status, (size_t)p->N * sizeof(int), kind_out(mem), p->stream));
    if (iters) CU(cudaMemcpyAsync(iters, p->iters, (size_t)p->N * sizeof(int), kind_out(mem), p->stream));
    if (mem == BLFIK_MEM_HOST) CU(cudaStreamSynchronize(p->stream));
    return BLFIK_OK;
}

extern "C" int blfik_count_unsolved(blfik_problem* p, int64_t* count)
{
    CHECK_P(p, "blfik_count_unsolved");
    if (!p->ticked) return fail(BLFIK_ERR_NOT_READY, "[blfik_count_unsolved] no tick has run yet");
    if (!count) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_count_unsolved] null output");
    CU(cudaMemsetAsync(p->d_count, 0, sizeof(unsigned long long), p->stream));
    blfik_count_unsolved_kernel<<<p->num_sms, 256, 0, p->stream>>>(p->status, p->N, p->d_count);
    CU(cudaGetLastError());
    p->launches++;
    unsigned long long h = 0;
    CU(cudaMemcpyAsync(&h, p->d_count, sizeof h, cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    *count = (int64_t)h;
    return BLFIK_OK;
}

extern "C" int blfik_kinematics(blfik_problem* p, const int32_t* frames, int32_t nframes, double* frame_pos, double* frame_rot,
                                double* frame_jac, double* com, double* com_jac, int32_t mem)
{
    CHECK_P(p, "blfik_kinematics");
    if (!p->state_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_kinematics] the robot state has never been set");
    if (nframes < 0 || (nframes > 0 && !frames)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_kinematics] bad frame list");
    for (int f = 0; f < nframes; f++)
        if (frames[f] < 0 || frames[f] >= p->model->nf) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_kinematics] frame index %d out of range", frames[f]);
    const size_t N = (size_t)p->N, nv = p->nv, nf = (size_t)nframes;
    // scratch on the device: outputs are produced there, then copied to wherever the caller wants them
    struct Buf { double* user; double* dev; size_t cnt; } bufs[5] = {
        {frame_pos, nullptr, N * nf * 3}, {frame_rot, nullptr, N * nf * 9}, {frame_jac, nullptr, N * nf * 6 * nv}, {com, nullptr, N * 3}, {com_jac, nullptr, N * 3 * nv}};
    std::vector<void*> tmp;
    auto cleanup = [&]() { for (void* t : tmp) cudaFree(t); };
    int* d_frames = nullptr;
    if (nframes) {
        if (cudaMalloc((void**)&d_frames, nf * sizeof(int)) != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_kinematics] cudaMalloc failed");
        tmp.push_back(d_frames);
        cudaMemcpyAsync(d_frames, frames, nf * sizeof(int), cudaMemcpyHostToDevice, p->stream);
    }
    for (Buf& b : bufs) {
        if (!b.user || !b.cnt) { b.user = nullptr; continue; }
        if (mem == BLFIK_MEM_DEVICE) b.dev = b.user;
        else {
            if (cudaMalloc((void**)&b.dev, b.cnt * sizeof(double)) != cudaSuccess) { cleanup(); return fail(BLFIK_ERR_CUDA, "[blfik_kinematics] cudaMalloc failed"); }
            tmp.push_back(b.dev);
        }
    }
    KinParams K;
    K.m = p->dm, K.N = p->N, K.base_pos = p->base_pos, K.base_quat = p->base_quat, K.s = p->s, K.frames = d_frames, K.nframes = nframes;
    K.frame_pos = bufs[0].user ? bufs[0].dev : nullptr, K.frame_rot = bufs[1].user ? bufs[1].dev : nullptr;
    K.frame_jac = bufs[2].user ? bufs[2].dev : nullptr, K.com = bufs[3].user ? bufs[3].dev : nullptr, K.com_jac = bufs[4].user ? bufs[4].dev : nullptr;
    K.L = p->L;
    blfik_kinematics_kernel<<<p->grid, p->warps_per_block * 32, p->smem_bytes, p->stream>>>(K);
    cudaError_t e = cudaGetLastError();
    p->launches++;
    if (e == cudaSuccess && mem == BLFIK_MEM_HOST)
        for (Buf& b : bufs)
            if (b.user && e == cudaSuccess) e = cudaMemcpyAsync(b.user, b.dev, b.cnt * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
    cleanup();
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_kinematics] %s", cudaGetErrorString(e));
    return BLFIK_OK;
}

// ------------------------------------------------------------------- allocation-free kinematics
// Staging memory for blfik_kinematics_noalloc.  It is part of the library image: every GPU that loads
// the library carries these 4 MiB until the process exits.  In HOST code the name below is only a
// stand-in for the symbol, not a device address; the address a kernel can use comes from
// cudaGetSymbolAddress, looked up once per device in noalloc_scratch().
static __device__ __align__(256) unsigned char g_noalloc_scratch[BLFIK_NOALLOC_SCRATCH_BYTES];

static constexpr int kMaxDevices = 64;
static unsigned char* g_noalloc_ptr[kMaxDevices];  // device address per device ordinal, 0 = not looked up yet (host threads that
                                                   // race on the first look-up all store the same value)

// Device address of the staging array on `device`, which must be the current device.
static int noalloc_scratch(int device, unsigned char** out)
{
    if (device < 0 || device >= kMaxDevices)
        return fail(BLFIK_ERR_UNSUPPORTED, "[blfik_kinematics_noalloc] device ordinal %d is beyond the %d this entry point tracks", device, kMaxDevices);
    if (!g_noalloc_ptr[device]) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return fail(BLFIK_ERR_CUDA, "[blfik_kinematics_noalloc] cannot resolve the staging array on device %d: %s", device, cudaGetErrorString(e));
        g_noalloc_ptr[device] = static_cast<unsigned char*>(ptr);
    }
    *out = g_noalloc_ptr[device];
    return BLFIK_OK;
}

extern "C" int blfik_kinematics_noalloc(blfik_problem* p, const int32_t* frames, int32_t nframes, double* frame_pos, double* frame_rot,
                                        double* frame_jac, double* com, double* com_jac, int32_t mem)
{
    CHECK_P(p, "blfik_kinematics_noalloc");  // makes p->device current, which the symbol look-up needs
    if (!p->state_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_kinematics_noalloc] the robot state has never been set");
    if (nframes < 0 || (nframes > 0 && !frames)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_kinematics_noalloc] bad frame list");
    for (int f = 0; f < nframes; f++)
        if (frames[f] < 0 || frames[f] >= p->model->nf) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_kinematics_noalloc] frame index %d out of range", frames[f]);
    const size_t N = (size_t)p->N, nv = p->nv, nf = (size_t)nframes;
    struct Buf { double* user; double* dev; size_t bytes; } bufs[5] = {
        {frame_pos, nullptr, N * nf * 3 * sizeof(double)}, {frame_rot, nullptr, N * nf * 9 * sizeof(double)},
        {frame_jac, nullptr, N * nf * 6 * nv * sizeof(double)}, {com, nullptr, N * 3 * sizeof(double)}, {com_jac, nullptr, N * 3 * nv * sizeof(double)}};
    auto round256 = [](size_t b) { return (b + 255) & ~(size_t)255; };
    // what has to live in the staging array: the frame list always, results only when they go to the host
    size_t need = round256(nf * sizeof(int));
    for (Buf& b : bufs) {
        if (!b.user || !b.bytes) { b.user = nullptr; continue; }
        if (mem != BLFIK_MEM_DEVICE) need += round256(b.bytes);
    }
    if (need > (size_t)BLFIK_NOALLOC_SCRATCH_BYTES)
        return fail(BLFIK_ERR_UNSUPPORTED,
                    "[blfik_kinematics_noalloc] this request needs %zu bytes of staging, the built-in array holds %zu "
                    "(BLFIK_NOALLOC_SCRATCH_BYTES); ask for fewer outputs or robots, or use blfik_kinematics",
                    need, (size_t)BLFIK_NOALLOC_SCRATCH_BYTES);
    unsigned char* base = nullptr;
    int rc = noalloc_scratch(p->device, &base);
    if (rc) return rc;
    size_t off = 0;
    int* d_frames = reinterpret_cast<int*>(base);
    off += round256(nf * sizeof(int));
    // From here on work may be in flight on the staging array: errors are collected, not returned on the spot, so that the
    // call always reaches the stream synchronisation below before it hands the array to the next caller.
    cudaError_t e = cudaSuccess;
    if (nframes) e = cudaMemcpyAsync(d_frames, frames, nf * sizeof(int), cudaMemcpyHostToDevice, p->stream);
    for (Buf& b : bufs) {
        if (!b.user) continue;
        if (mem == BLFIK_MEM_DEVICE) b.dev = b.user;
        else {
            b.dev = reinterpret_cast<double*>(base + off);
            off += round256(b.bytes);
        }
    }
    if (e == cudaSuccess) {
        KinParams K;
        K.m = p->dm, K.N = p->N, K.base_pos = p->base_pos, K.base_quat = p->base_quat, K.s = p->s, K.frames = d_frames, K.nframes = nframes;
        K.frame_pos = bufs[0].user ? bufs[0].dev : nullptr, K.frame_rot = bufs[1].user ? bufs[1].dev : nullptr;
        K.frame_jac = bufs[2].user ? bufs[2].dev : nullptr, K.com = bufs[3].user ? bufs[3].dev : nullptr, K.com_jac = bufs[4].user ? bufs[4].dev : nullptr;
        K.L = p->L;
        blfik_kinematics_kernel<<<p->grid, p->warps_per_block * 32, p->smem_bytes, p->stream>>>(K);
        e = cudaGetLastError();
        p->launches++;
    }
    if (e == cudaSuccess && mem != BLFIK_MEM_DEVICE)
        for (Buf& b : bufs)
            if (b.user && e == cudaSuccess) e = cudaMemcpyAsync(b.user, b.dev, b.bytes, cudaMemcpyDeviceToHost, p->stream);
    // the staging array is free for the next caller only once everything has left it -- also after a failure
    const cudaError_t es = cudaStreamSynchronize(p->stream);
    if (e == cudaSuccess) e = es;
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_kinematics_noalloc] %s", cudaGetErrorString(e));
    return BLFIK_OK;
}

extern "C" int blfik_integrate(blfik_problem* p, const double* nu, int32_t mem, double dt, int32_t nsteps)
{
    CHECK_P(p, "blfik_integrate");
    if (!p->state_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_integrate] the state has never been set");
    // FixedStepIntegrator.h:73-85: the step must be positive; :94-140 with tf - t0 < dT underflows in the
    // reference -- rejected here instead of reproduced
    if (!(dt > 0)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate] the integration step must be positive");
    if (nsteps < 1) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate] nsteps must be >= 1 (final time before initial time + dT)");
    const double* src = p->nu;
    if (nu) {
        if (mem == BLFIK_MEM_DEVICE) src = nu;
        else {
            CU(cudaMemcpyAsync(p->nu, nu, (size_t)p->N * p->nv * sizeof(double), cudaMemcpyHostToDevice, p->stream));
            p->ticked = true;  // nu now holds a caller-provided control input
        }
    } else if (!p->ticked) {
        return fail(BLFIK_ERR_NOT_READY, "[blfik_integrate] no control input: pass nu or run a tick first");
    }
    const long long blocks = (p->N * 32 + 255) / 256;
    const long long cap = (long long)p->num_sms * 8;
    blfik_integrate_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, p->stream>>>(p->n, p->N, p->base_pos, p->base_quat, p->s, src, dt, nsteps);
    CU(cudaGetLastError());
    p->launches++;
    return BLFIK_OK;
}

extern "C" int blfik_measure_fp64_peak(int32_t device, double millis, double* tflops)
{
    if (!tflops) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_measure_fp64_peak] null output");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    double* sink = nullptr;
    CU(cudaMalloc((void**)&sink, sizeof(double)));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int blocks = prop.multiProcessorCount * 4, threads = 512;
    int iters = 2000;
    double best = 0.0, spent = 0.0;
    blfik_dfma_probe_kernel<<<blocks, threads>>>(sink, 200, 0.999999, 1e-9);  // warm-up
    CU(cudaDeviceSynchronize());
    while (spent < millis) {
        CU(cudaEventRecord(e0));
        blfik_dfma_probe_kernel<<<blocks, threads>>>(sink, iters, 0.999999, 1e-9);
        CU(cudaEventRecord(e1));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        spent += ms;
        const double flops = 2.0 * 64.0 * (double)iters * blocks * threads;
        const double tf = flops / (ms * 1e-3) * 1e-12;
        if (tf > best) best = tf;
        if (ms < 5.0) iters *= 2;
    }
    cudaEventDestroy(e0), cudaEventDestroy(e1), cudaFree(sink);
    *tflops = best;
    return BLFIK_OK;
}

extern "C" int blfik_integrate_state(int32_t n, int64_t N, double* bp, double* bq, double* s, const double* nu, double dt, int32_t nsteps,
                                     int32_t mem, int32_t device)
{
    if (n < 1 || N < 1 || !bp || !bq || !s || !nu) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_state] bad sizes or null buffer");
    if (!(dt > 0)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_state] the integration step must be positive");
    if (nsteps < 1) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_state] nsteps must be >= 1");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(BLFIK_ERR_CUDA, "[blfik_integrate_state] no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_state] device %d out of range", device);
    CU(cudaSetDevice(device));
    const size_t Ns = (size_t)N, nv = (size_t)n + 6;
    double *dp = bp, *dq = bq, *ds = s;
    const double* dnu = nu;
    double* stage = nullptr;
    if (mem != BLFIK_MEM_DEVICE) {
        const size_t total = Ns * (3 + 4 + (size_t)n + nv);
        CU(cudaMalloc((void**)&stage, total * sizeof(double)));
        dp = stage, dq = dp + Ns * 3, ds = dq + Ns * 4;
        double* dn = ds + Ns * n;
        dnu = dn;
        cudaMemcpy(dp, bp, Ns * 3 * sizeof(double), cudaMemcpyHostToDevice);
        cudaMemcpy(dq, bq, Ns * 4 * sizeof(double), cudaMemcpyHostToDevice);
        cudaMemcpy(ds, s, Ns * n * sizeof(double), cudaMemcpyHostToDevice);
        cudaMemcpy(dn, nu, Ns * nv * sizeof(double), cudaMemcpyHostToDevice);
    }
    const long long blocks = ((long long)N * 32 + 255) / 256;
    blfik_integrate_kernel<<<(int)(blocks < 148 * 8 ? blocks : 148 * 8), 256>>>(n, N, dp, dq, ds, dnu, dt, nsteps);
    e = cudaGetLastError();
    if (e == cudaSuccess && mem != BLFIK_MEM_DEVICE) {
        cudaMemcpy(bp, dp, Ns * 3 * sizeof(double), cudaMemcpyDeviceToHost);
        cudaMemcpy(bq, dq, Ns * 4 * sizeof(double), cudaMemcpyDeviceToHost);
        e = cudaMemcpy(s, ds, Ns * n * sizeof(double), cudaMemcpyDeviceToHost);
    } else if (e == cudaSuccess) {
        e = cudaDeviceSynchronize();
    }
    if (stage) cudaFree(stage);
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_integrate_state] %s", cudaGetErrorString(e));
    return BLFIK_OK;
}

extern "C" int blfik_integrate_acceleration_state(int32_t n, int64_t N, double* bp, double* bq, double* s, double* dp, double* omega, double* ds,
                                                  const double* acc, double dt, int32_t nsteps, int32_t device)
{
    if (n < 1 || N < 1 || !bp || !bq || !s || !dp || !omega || !ds || !acc)
        return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_acceleration_state] bad sizes or null buffer");
    if (!(dt > 0)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_acceleration_state] the integration step must be positive");
    if (nsteps < 1) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_acceleration_state] nsteps must be >= 1");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return fail(BLFIK_ERR_CUDA, "[blfik_integrate_acceleration_state] no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_integrate_acceleration_state] device %d out of range", device);
    CU(cudaSetDevice(device));
    // host arrays in, host arrays out: staged through one temporary device buffer
    const size_t Ns = (size_t)N, nv = (size_t)n + 6;
    const size_t cnt[7] = {Ns * 3, Ns * 4, Ns * n, Ns * 3, Ns * 3, Ns * n, Ns * nv};
    double* host[7] = {bp, bq, s, dp, omega, ds, const_cast<double*>(acc)};
    size_t total = 0;
    for (size_t c : cnt) total += c;
    double* stage = nullptr;
    CU(cudaMalloc((void**)&stage, total * sizeof(double)));
    double* dev[7];
    size_t off = 0;
    for (int k = 0; k < 7; k++) {
        dev[k] = stage + off, off += cnt[k];
        cudaMemcpy(dev[k], host[k], cnt[k] * sizeof(double), cudaMemcpyHostToDevice);
    }
    const long long blocks = ((long long)N * 32 + 255) / 256;
    blfik_integrate_acceleration_kernel<<<(int)(blocks < 148 * 8 ? blocks : 148 * 8), 256>>>(n, N, dev[0], dev[1], dev[2], dev[3], dev[4], dev[5], dev[6], dt, nsteps);
    e = cudaGetLastError();
    for (int k = 0; k < 6 && e == cudaSuccess; k++) e = cudaMemcpy(host[k], dev[k], cnt[k] * sizeof(double), cudaMemcpyDeviceToHost);
    cudaFree(stage);
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_integrate_acceleration_state] %s", cudaGetErrorString(e));
    return BLFIK_OK;
}

extern "C" int blfik_dynamics(blfik_problem* p, double* mass_matrix, double* gravity_forces, double gravity_acceleration, int32_t mem)
{
    CHECK_P(p, "blfik_dynamics");
    if (!p->state_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_dynamics] the robot state has never been set");
    if (!mass_matrix && !gravity_forces) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_dynamics] nothing asked for");
    const size_t N = (size_t)p->N, nv = p->nv;
    double *dM = mass_matrix, *dG = gravity_forces;
    std::vector<void*> tmp;
    if (mem != BLFIK_MEM_DEVICE) {
        if (mass_matrix) { if (cudaMalloc((void**)&dM, N * nv * nv * sizeof(double)) != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_dynamics] cudaMalloc failed"); tmp.push_back(dM); }
        if (gravity_forces) { if (cudaMalloc((void**)&dG, N * nv * sizeof(double)) != cudaSuccess) { for (void* t : tmp) cudaFree(t); return fail(BLFIK_ERR_CUDA, "[blfik_dynamics] cudaMalloc failed"); } tmp.push_back(dG); }
    }
    KinParams K;
    K.m = p->dm, K.N = p->N, K.base_pos = p->base_pos, K.base_quat = p->base_quat, K.s = p->s, K.frames = nullptr, K.nframes = 0;
    K.frame_pos = K.frame_rot = K.frame_jac = K.com = K.com_jac = nullptr;
    K.L = p->L;
    cudaError_t e = cudaFuncSetAttribute(blfik_dynamics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem_bytes);
    if (e == cudaSuccess) {
        blfik_dynamics_kernel<<<p->grid, p->warps_per_block * 32, p->smem_bytes, p->stream>>>(K, dM, dG, gravity_acceleration);
        e = cudaGetLastError();
        p->launches++;
    }
    if (e == cudaSuccess && mem != BLFIK_MEM_DEVICE) {
        if (mass_matrix) e = cudaMemcpyAsync(mass_matrix, dM, N * nv * nv * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
        if (gravity_forces && e == cudaSuccess) e = cudaMemcpyAsync(gravity_forces, dG, N * nv * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
    for (void* t : tmp) cudaFree(t);
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_dynamics] %s", cudaGetErrorString(e));
    return BLFIK_OK;
}

extern "C" int blfik_bias_forces(blfik_problem* p, const double* nu, double gravity_acceleration, double* bias_forces, const int32_t* frames,
                                 int32_t nframes, double* frame_bias_acc, double* frame_vel, double* com_bias_acc, int32_t mem)
{
    CHECK_P(p, "blfik_bias_forces");
    if (!p->state_set) return fail(BLFIK_ERR_NOT_READY, "[blfik_bias_forces] the robot state has never been set");
    if (!nu) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_bias_forces] null generalised velocity");
    if (nframes < 0 || (nframes > 0 && !frames)) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_bias_forces] invalid frame list");
    if (!frame_bias_acc && !frame_vel) nframes = 0;
    if (!bias_forces && !com_bias_acc && nframes == 0) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_bias_forces] nothing asked for");
    for (int f = 0; f < nframes; f++)
        if (frames[f] < 0 || frames[f] >= p->model->nf) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_bias_forces] frame index %d out of range", frames[f]);
    const size_t N = (size_t)p->N, nv = p->nv;
    const bool host = mem != BLFIK_MEM_DEVICE;
    const size_t n_nu = N * nv, n_h = bias_forces ? N * nv : 0, n_f = frame_bias_acc ? (size_t)nframes * 6 * N : 0, n_v = frame_vel ? (size_t)nframes * 6 * N : 0,
                 n_c = com_bias_acc ? 3 * N : 0;
    double* stage = nullptr;
    int* d_frames = nullptr;
    auto cleanup = [&] { cudaFree(stage), cudaFree(d_frames); };
    const double* d_nu = nu;
    double *d_h = bias_forces, *d_f = frame_bias_acc, *d_v = frame_vel, *d_c = com_bias_acc;
    if (host) {
        if (cudaMalloc((void**)&stage, (n_nu + n_h + n_f + n_v + n_c) * sizeof(double)) != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_bias_forces] cudaMalloc failed");
        cudaMemcpyAsync(stage, nu, n_nu * sizeof(double), cudaMemcpyHostToDevice, p->stream);
        d_nu = stage, d_h = n_h ? stage + n_nu : nullptr, d_f = n_f ? stage + n_nu + n_h : nullptr, d_v = n_v ? stage + n_nu + n_h + n_f : nullptr;
        d_c = n_c ? stage + n_nu + n_h + n_f + n_v : nullptr;
    }
    if (nframes) {
        if (cudaMalloc((void**)&d_frames, nframes * sizeof(int)) != cudaSuccess) { cleanup(); return fail(BLFIK_ERR_CUDA, "[blfik_bias_forces] cudaMalloc failed"); }
        cudaMemcpyAsync(d_frames, frames, nframes * sizeof(int), cudaMemcpyHostToDevice, p->stream);
    }
    KinParams K;
    K.m = p->dm, K.N = p->N, K.base_pos = p->base_pos, K.base_quat = p->base_quat, K.s = p->s, K.frames = d_frames, K.nframes = nframes;
    K.frame_pos = K.frame_rot = K.frame_jac = K.com = K.com_jac = nullptr;
    K.L = p->L;
    cudaError_t e = cudaFuncSetAttribute(blfik_bias_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem_bytes);
    if (e == cudaSuccess) {
        blfik_bias_kernel<<<p->grid, p->warps_per_block * 32, p->smem_bytes, p->stream>>>(K, d_nu, gravity_acceleration, d_h, d_f, d_v, d_c);
        e = cudaGetLastError();
        p->launches++;
    }
    if (e == cudaSuccess && host) {
        if (bias_forces) e = cudaMemcpyAsync(bias_forces, d_h, n_h * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
        if (n_f && e == cudaSuccess) e = cudaMemcpyAsync(frame_bias_acc, d_f, n_f * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
        if (n_v && e == cudaSuccess) e = cudaMemcpyAsync(frame_vel, d_v, n_v * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
        if (com_bias_acc && e == cudaSuccess) e = cudaMemcpyAsync(com_bias_acc, d_c, n_c * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);   // the staging buffers and the frame list die with this call
    cleanup();
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_bias_forces] %s", cudaGetErrorString(e));
    return BLFIK_OK;
}

extern "C" int32_t blfik_num_dofs(const blfik_problem* p) { return p ? p->n : -1; }
extern "C" int32_t blfik_num_equality_rows(const blfik_problem* p) { return p ? p->me : -1; }
extern "C" int64_t blfik_num_instances(const blfik_problem* p) { return p ? p->N : -1; }
extern "C" int64_t blfik_launch_count(const blfik_problem* p) { return p ? p->launches : -1; }

extern "C" int blfik_host_alloc(size_t bytes, void** out)
{
    if (!out || bytes == 0) return fail(BLFIK_ERR_INVALID_ARGUMENT, "[blfik_host_alloc] null argument or zero size");
    const cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocPortable | cudaHostAllocMapped);
    if (e != cudaSuccess) return fail(BLFIK_ERR_CUDA, "[blfik_host_alloc] cudaHostAlloc(%zu): %s", bytes, cudaGetErrorString(e));
    return BLFIK_OK;
}

extern "C" void blfik_host_free(void* ptr)
{
    if (ptr) cudaFreeHost(ptr);
}

extern "C" int blfik_kernel_timing(blfik_problem* p, int32_t enabled)
{
    CHECK_P(p, "blfik_kernel_timing");
    if (enabled && !p->ev_t[0])
        for (int k = 0; k < 3; k++) CU(cudaEventCreate(&p->ev_t[k]));
    p->timing = enabled != 0;
    return BLFIK_OK;
}

extern "C" int blfik_get_kernel_times(blfik_problem* p, double* ms_assemble, double* ms_solve)
{
    CHECK_P(p, "blfik_get_kernel_times");
    if (!p->timing || !p->solve_kernel || p->path != 0) return fail(BLFIK_ERR_NOT_READY, "[blfik_get_kernel_times] timing is off or the two-launch path is not in use");
    CU(cudaEventSynchronize(p->ev_t[2]));
    float a = 0.f, b = 0.f;
    CU(cudaEventElapsedTime(&a, p->ev_t[0], p->ev_t[1]));
    CU(cudaEventElapsedTime(&b, p->ev_t[1], p->ev_t[2]));
    if (ms_assemble) *ms_assemble = a;
    if (ms_solve) *ms_solve = b;
    return BLFIK_OK;
}

extern "C" int blfik_sync(blfik_problem* p)
{
    CHECK_P(p, "blfik_sync");
    CU(cudaStreamSynchronize(p->stream));
    return BLFIK_OK;
}

extern "C" void* blfik_get_stream(blfik_problem* p) { return p ? (void*)p->stream : nullptr; }

extern "C" int blfik_set_stream(blfik_problem* p, void* s)
{
    CHECK_P(p, "blfik_set_stream");
    CU(cudaStreamSynchronize(p->stream));
    if (p->own_stream && p->stream) cudaStreamDestroy(p->stream);
    p->stream = (cudaStream_t)s;
    p->own_stream = false;
    return BLFIK_OK;
}
