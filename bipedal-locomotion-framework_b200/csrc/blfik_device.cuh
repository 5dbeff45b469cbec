// blfik_device.cuh -- device side of the batched IK tick (sm_100a, FP64).
//
// Execution model: ONE WARP PER ROBOT.  Lane i owns velocity variable i (row i of every
// nv x nv matrix); all per-robot matrices live in that warp's slice of shared memory, the
// kinematic tree (batch-shared constants) is read through L1.  Control flow is warp-uniform
// by construction (loop bounds depend only on the model/problem, never on the robot), so a
// warp never diverges except on the rare "polish rejected" retry path.
//
// Stages:  K1 fk_jac -> K2 task_assemble -> K3 qp (ADMM + polish) -> K4 se3_euler
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace blfik {

constexpr double kInf = 1e30;       // OsqpEigen::INFTY
constexpr double kInfHalf = 0.5e30;
constexpr unsigned kFull = 0xffffffffu;

enum TaskType : int { T_SE3 = 0, T_COM = 1, T_SO3 = 2, T_JOINT_TRACKING = 3, T_JOINT_LIMITS = 4, T_JOINT_VEL_LIMITS = 5, T_R3 = 6, T_DISTANCE = 7, T_GRAVITY = 8, T_ANGMOM = 9 };
__host__ __device__ inline bool is_dense_type(int t) { return t == T_SE3 || t == T_COM || t == T_SO3 || t == T_R3 || t == T_DISTANCE || t == T_GRAVITY || t == T_ANGMOM; }

struct DevModel {
    int n, nl, nf, maxdepth;
    const int* parent;       // [nl]
    const int* dof;          // [nl]
    const int* depth;        // [nl]
    const int* link_of_dof;  // [n]
    const unsigned long long* anc;  // [nl] bit j set <=> DoF j is on the path base -> link
    const double* Xr;        // [nl][9]
    const double* Xp;        // [nl][3]
    const double* axis;      // [nl][3]
    const double* mass;      // [nl]
    const double* com;       // [nl][3]
    const double* msub;      // [n] mass of the subtree distal to DoF j
    double inv_total_mass;
    const int* frame_link;   // [nf]
    const double* Fr;        // [nf][9]
    const double* Fp;        // [nf][3]
    const double* inertia;   // [nl][6] Ixx Ixy Ixz Iyy Iyz Izz about the link CoM, link frame (nullptr: point masses)
};

struct DevTask {
    int type, priority, frame, rows, enabled, sp_len, row_offset;
    int brow;  // first row in the dense task-row matrix of the register-resident kernel
    int mask[3];
    double kp_lin[3], kp_ang[3];
    const double* kp_joint;
    const double* weight;
    const double* klim;
    const double* upper;
    const double* lower;
    double inv_dt;
    const double* setpoint;  // device [N][sp_len]
    double Fr[9], Fp[3];     // the task frame relative to its link (copied from the model)
    // optional on-device set-point trajectory: base set-point + amp * sin(omega t + phase_i) (omega = 0: off)
    const double* traj_phase;   // device [N]
    double traj_amp[6];         // linear (3) and angular (3, rotation vector) amplitudes
    double traj_omega;
    int flink;               // link the task frame is attached to
    int flink2;              // DistanceTask: link of the reference frame, -1 = the world origin
    double Fp2[3];           // DistanceTask: origin of the reference frame in its link
    double thr;              // DistanceTask: distance_numeric_threshold
    int pk_off;              // offset of this task's set-point inside a packed input record
    int pk_off_c;            // the same in the compact layout (no feed-forward terms: pose part, then the feedback)
    int base_c;              // doubles of the pose part in the compact layout (SE3 7, CoM / R3 3, SO3 4, joints n, ...)
    // optional minimum-jerk point-to-point generator (Planners::SO3Planner with zero boundary velocity / acceleration, and
    // the quintic between two knots for positions): from the stored set-point to mj_target over [mj_t0, mj_t0 + mj_T]
    const double* mj_target;    // device [N][set-point length], nullptr: off
    double mj_t0, mj_T;
    int exo_pos, exo_ori;    // exogenous feedback (SE3Task.cpp:261-268, CoMTask.cpp:192-195): controller state from the record
    int fb_off;              // where the feedback p(3) [q(4)] sits inside the record (= the set-point length)
    int sp_stride;           // doubles between consecutive robots in `setpoint`: sp_len, or 0 for a batch-shared set-point
    int gen;                 // attached set-point generators: 1 sinusoidal, 2 minimum jerk, 4 quintic spline (0: none)
    // optional multi-knot quintic spline for the linear part (Math::QuinticSpline, Spline.h / QuinticSpline.h): knot times
    // shared by the robots, one polynomial per interval and robot, coefficients [N][sp_nk - 1][6][3] in interval-local time
    int sp_nk;
    const double* sp_times;  // device [sp_nk]
    const double* sp_coef;   // device, written by blfik_spline_coefficients_kernel
    int sp_dim;              // coordinates of the spline: 3 (SE3 / R3 / CoM), n for a JointTrackingTask (Spline<VectorXd>)
};

struct Settings {
    double rho, sigma, alpha, eps;
    int admm_iters, max_rounds, polish_iters, eq_iters;
};

// shared-memory layout of one warp, in doubles (computed on the host)
struct Layout {
    int nv, me, LD, LDE;
    int K;                                 // dense task rows, equality + cost (row layout of the equality-only kernel)
    int box;                               // that kernel: 1 if a task bounds the joint velocities (lo / hi live through K3)
    int linkR, linkP, axisW, mcw;          // kinematics
    int Hs, A, B, Cs, Ym, Wm, Ss, Jt;      // matrices
    int g, b, lo, hi, x, z, y, r, xb, wb, lam, gp, bp, bd, t;  // vectors
    int act;                               // int[nv] stored in a double slot array
    int total;                             // doubles per warp
};

struct DebugOut {
    double *H, *g, *Ceq, *beq, *lo, *hi;   // any may be null
};

struct TickParams {
    DevModel m;
    const DevTask* tasks;
    int ntasks;
    long long first, N;                    // this launch covers robots [first, N)
    double t0;                             // time of the first tick of this launch (set-point trajectories)
    double* nu_host;                       // optional second copy of nu / status written straight into mapped pinned host
    int* status_host;                      //   memory (zero-copy host tick); null otherwise
    const double* packed;                  // optional packed input records [.][packed_len]: p(3) q(4) s(n) set-points...
    int packed_len;
    int compact;                           // 1: packed records in the compact layout (BLFIK_TICK_NO_FEEDFORWARD)
    double *base_pos, *base_quat, *s;      // state (in/out)
    double* nu;                            // [N][nv] out
    double *wx, *wz, *wy;                  // warm start [N][nv], [N][n], [N][n]
    int *status, *iters;
    Settings st;
    int nticks;
    double dt;
    unsigned flags;
    DebugOut dbg;
    Layout L;
};

struct KinParams {
    DevModel m;
    long long N;
    const double *base_pos, *base_quat, *s;
    const int* frames;
    int nframes;
    double *frame_pos, *frame_rot, *frame_jac, *com, *com_jac;
    Layout L;
};

struct WM {  // one warp's shared memory, as pointers
    double *linkR, *linkP, *axisW, *mcw;
    double *Hs, *A, *B, *Cs, *Ym, *Wm, *Ss, *Jt;
    double *g, *b, *lo, *hi, *x, *z, *y, *r, *xb, *wb, *lam, *gp, *bp, *bd, *t;
    int* act;
};

__device__ __forceinline__ WM carve(double* base, const Layout& L)
{
    WM w;
    w.linkR = base + L.linkR, w.linkP = base + L.linkP, w.axisW = base + L.axisW, w.mcw = base + L.mcw;
    w.Hs = base + L.Hs, w.A = base + L.A, w.B = base + L.B, w.Cs = base + L.Cs;
    w.Ym = base + L.Ym, w.Wm = base + L.Wm, w.Ss = base + L.Ss, w.Jt = base + L.Jt;
    w.g = base + L.g, w.b = base + L.b, w.lo = base + L.lo, w.hi = base + L.hi;
    w.x = base + L.x, w.z = base + L.z, w.y = base + L.y, w.r = base + L.r;
    w.xb = base + L.xb, w.wb = base + L.wb, w.lam = base + L.lam;
    w.gp = base + L.gp, w.bp = base + L.bp, w.bd = base + L.bd, w.t = base + L.t;
    w.act = reinterpret_cast<int*>(base + L.act);
    return w;
}

// ------------------------------------------------------------------ small math
__device__ __forceinline__ void mat3_mul(const double* A, const double* B, double* C)
{
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
__device__ __forceinline__ void mat3T_vec(const double* A, const double* v, double* o)
{
#pragma unroll
    for (int i = 0; i < 3; i++) o[i] = A[i] * v[0] + A[3 + i] * v[1] + A[6 + i] * v[2];
}
__device__ __forceinline__ void mat3_vec(const double* A, const double* v, double* o)
{
#pragma unroll
    for (int i = 0; i < 3; i++) o[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}
__device__ __forceinline__ void cross3(const double* a, const double* b, double* o)
{
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ double warp_max(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

__device__ __forceinline__ void quat_to_rot(const double* q, double* R)
{
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    R[0] = 1 - 2 * (y * y + z * z), R[1] = 2 * (x * y - z * w), R[2] = 2 * (x * z + y * w);
    R[3] = 2 * (x * y + z * w), R[4] = 1 - 2 * (x * x + z * z), R[5] = 2 * (y * z - x * w);
    R[6] = 2 * (x * z - y * w), R[7] = 2 * (y * z + x * w), R[8] = 1 - 2 * (x * x + y * y);
}

// rotation matrix -> unit quaternion (x,y,z,w); Conversions::toManifRot semantics
// (ManifConversions.h:44-50: Eigen Quaternion(Matrix3) then normalize()).
__device__ __forceinline__ void rot_to_quat(const double* R, double* q)
{
    double t = R[0] + R[4] + R[8];
    if (t > 0) {
        t = sqrt(t + 1.0);
        q[3] = 0.5 * t;
        t = 0.5 / t;
        q[0] = (R[7] - R[5]) * t, q[1] = (R[2] - R[6]) * t, q[2] = (R[3] - R[1]) * t;
    } else if (R[0] >= R[4] && R[0] >= R[8]) {
        t = sqrt(R[0] - R[4] - R[8] + 1.0);
        q[0] = 0.5 * t;
        t = 0.5 / t;
        q[3] = (R[7] - R[5]) * t, q[1] = (R[3] + R[1]) * t, q[2] = (R[6] + R[2]) * t;
    } else if (R[4] > R[0] && R[4] >= R[8]) {
        t = sqrt(R[4] - R[8] - R[0] + 1.0);
        q[1] = 0.5 * t;
        t = 0.5 / t;
        q[3] = (R[2] - R[6]) * t, q[2] = (R[7] + R[5]) * t, q[0] = (R[1] + R[3]) * t;
    } else {
        t = sqrt(R[8] - R[0] - R[4] + 1.0);
        q[2] = 0.5 * t;
        t = 0.5 / t;
        q[3] = (R[3] - R[1]) * t, q[0] = (R[2] + R[6]) * t, q[1] = (R[5] + R[7]) * t;
    }
    const double inv = rsqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    q[0] *= inv, q[1] *= inv, q[2] *= inv, q[3] *= inv;
}

__device__ __forceinline__ void quat_mul(const double* a, const double* b, double* o)
{
    o[0] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
    o[1] = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
    o[2] = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
    o[3] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
}

// manif SO3::log on a unit quaternion
__device__ __forceinline__ void so3_log(const double* q, double* phi)
{
    const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
    const double w = q[3];
    double scale;
    if (n2 > 1e-20) {
        const double n = sqrt(n2);
        const double two_angle = 2.0 * (w < 0 ? atan2(-n, -w) : atan2(n, w));
        scale = two_angle / n;
    } else {
        scale = 2.0 / w;
    }
    phi[0] = scale * q[0], phi[1] = scale * q[1], phi[2] = scale * q[2];
}

// manif SO3Tangent::exp
__device__ __forceinline__ void so3_exp(const double* phi, double* q)
{
    const double th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
    double k;
    if (th2 > 1e-16) {
        const double th = sqrt(th2);
        double sn, cs;
        sincos(0.5 * th, &sn, &cs);
        k = sn / th;
        q[3] = cs;
    } else {
        k = 0.5 - th2 / 48.0;
        q[3] = 1.0 - th2 / 8.0;
    }
    q[0] = k * phi[0], q[1] = k * phi[1], q[2] = k * phi[2];
}

// LieGroupControllers P-law on SO(3), left trivialisation: w* = w_ff + Kp o Log(R_d R^-1)
__device__ __forceinline__ void so3_control(const double* Rcur, const double* qd, const double* wff,
                                            const double* kp, bool enabled, double* out)
{
    double qc[4], qe[4], e[3];
    rot_to_quat(Rcur, qc);
    qc[0] = -qc[0], qc[1] = -qc[1], qc[2] = -qc[2];
    quat_mul(qd, qc, qe);
    so3_log(qe, e);
#pragma unroll
    for (int c = 0; c < 3; c++) out[c] = enabled ? wff[c] + kp[c] * e[c] : wff[c];
}

// ------------------------------------------------------------------ K1: FK + Jacobians
// Level-synchronous tree walk, one link per lane.  Leaves in shared memory, for every link,
// the world rotation/origin, the world joint axis and mass * world CoM.
__device__ __forceinline__ void k1_fk(const DevModel& m, const WM& w, const double* bp, const double* bq,
                                      const double* s, int lane)
{
    for (int k = lane; k < m.nl; k += 32) {
        double R[9];
        if (k == 0) {
            double q[4] = {bq[0], bq[1], bq[2], bq[3]};
            quat_to_rot(q, R);
            w.linkP[0] = bp[0], w.linkP[1] = bp[1], w.linkP[2] = bp[2];
        } else {
            // local transform Xr * Rot(axis, s)
            const double* a = m.axis + 3 * k;
            double sn, cs;
            sincos(s[m.dof[k]], &sn, &cs);
            const double v = 1 - cs;
            double Rj[9];
            Rj[0] = cs + a[0] * a[0] * v, Rj[1] = a[0] * a[1] * v - a[2] * sn, Rj[2] = a[0] * a[2] * v + a[1] * sn;
            Rj[3] = a[1] * a[0] * v + a[2] * sn, Rj[4] = cs + a[1] * a[1] * v, Rj[5] = a[1] * a[2] * v - a[0] * sn;
            Rj[6] = a[2] * a[0] * v - a[1] * sn, Rj[7] = a[2] * a[1] * v + a[0] * sn, Rj[8] = cs + a[2] * a[2] * v;
            double X[9];
#pragma unroll
            for (int c = 0; c < 9; c++) X[c] = m.Xr[9 * k + c];
            mat3_mul(X, Rj, R);
        }
#pragma unroll
        for (int c = 0; c < 9; c++) w.linkR[9 * k + c] = R[c];
    }
    __syncwarp();
    for (int d = 1; d <= m.maxdepth; d++) {
        for (int k = lane; k < m.nl; k += 32) {
            if (m.depth[k] == d) {
                const int par = m.parent[k];
                double Rp[9], Rl[9], R[9], xp[3], off[3];
#pragma unroll
                for (int c = 0; c < 9; c++) Rp[c] = w.linkR[9 * par + c], Rl[c] = w.linkR[9 * k + c];
                mat3_mul(Rp, Rl, R);
                xp[0] = m.Xp[3 * k], xp[1] = m.Xp[3 * k + 1], xp[2] = m.Xp[3 * k + 2];
                mat3_vec(Rp, xp, off);
#pragma unroll
                for (int c = 0; c < 9; c++) w.linkR[9 * k + c] = R[c];
#pragma unroll
                for (int c = 0; c < 3; c++) w.linkP[3 * k + c] = w.linkP[3 * par + c] + off[c];
            }
        }
        __syncwarp();
    }
    for (int k = lane; k < m.nl; k += 32) {
        double R[9], a[3], c[3], aw[3], cw[3];
#pragma unroll
        for (int i = 0; i < 9; i++) R[i] = w.linkR[9 * k + i];
#pragma unroll
        for (int i = 0; i < 3; i++) a[i] = m.axis[3 * k + i], c[i] = m.com[3 * k + i];
        mat3_vec(R, a, aw);
        mat3_vec(R, c, cw);
        const double mk = m.mass[k];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            w.axisW[3 * k + i] = aw[i];
            w.mcw[3 * k + i] = mk * (w.linkP[3 * k + i] + cw[i]);
        }
    }
    __syncwarp();
}

// world pose of a frame (every lane computes the same values)
__device__ __forceinline__ void k1_frame_pose(const DevModel& m, const WM& w, int frame, double* R, double* p)
{
    const int L = m.frame_link[frame];
    double Rl[9], Fr[9], fp[3], off[3];
#pragma unroll
    for (int c = 0; c < 9; c++) Rl[c] = w.linkR[9 * L + c], Fr[c] = m.Fr[9 * frame + c];
#pragma unroll
    for (int c = 0; c < 3; c++) fp[c] = m.Fp[3 * frame + c];
    mat3_mul(Rl, Fr, R);
    mat3_vec(Rl, fp, off);
#pragma unroll
    for (int c = 0; c < 3; c++) p[c] = w.linkP[3 * L + c] + off[c];
}

// MIXED-representation 6 x nv Jacobian of a point pf rigidly attached to link L; column c by lane c.
// dst has row stride ld; rows 0-2 linear, 3-5 angular.
__device__ __forceinline__ void k1_frame_jac(const DevModel& m, const WM& w, int L, const double* pf, double* dst,
                                             int ld, int lane)
{
    const int nv = m.n + 6;
    const unsigned long long anc = m.anc[L];
    for (int c = lane; c < nv; c += 32) {
        double lin[3] = {0, 0, 0}, ang[3] = {0, 0, 0};
        if (c < 3) {
            lin[c] = 1.0;
        } else if (c < 6) {
            const double r[3] = {pf[0] - w.linkP[0], pf[1] - w.linkP[1], pf[2] - w.linkP[2]};
            ang[c - 3] = 1.0;
            cross3(ang, r, lin);
        } else if ((anc >> (c - 6)) & 1ull) {
            const int k = m.link_of_dof[c - 6];
            const double d[3] = {pf[0] - w.linkP[3 * k], pf[1] - w.linkP[3 * k + 1], pf[2] - w.linkP[3 * k + 2]};
            ang[0] = w.axisW[3 * k], ang[1] = w.axisW[3 * k + 1], ang[2] = w.axisW[3 * k + 2];
            cross3(ang, d, lin);
        }
#pragma unroll
        for (int i = 0; i < 3; i++) dst[i * ld + c] = lin[i], dst[(3 + i) * ld + c] = ang[i];
    }
}

// CoM position (all lanes) and 3 x nv CoM Jacobian through subtree sums:
// column j = (1/M) a_j x (sum_{k in sub(j)} m_k c_k - M_sub(j) o_j)
__device__ __forceinline__ void k1_com(const DevModel& m, const WM& w, double* com, double* dst, int ld, int lane)
{
    const int nv = m.n + 6;
    double acc[3] = {0, 0, 0};
    for (int k = lane; k < m.nl; k += 32) {
        acc[0] += w.mcw[3 * k], acc[1] += w.mcw[3 * k + 1], acc[2] += w.mcw[3 * k + 2];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) com[i] = warp_sum(acc[i]) * m.inv_total_mass;
    if (!dst) return;
    for (int c = lane; c < nv; c += 32) {
        double lin[3] = {0, 0, 0};
        if (c < 3) {
            lin[c] = 1.0;
        } else if (c < 6) {
            const double r[3] = {com[0] - w.linkP[0], com[1] - w.linkP[1], com[2] - w.linkP[2]};
            double e[3] = {0, 0, 0};
            e[c - 3] = 1.0;
            cross3(e, r, lin);
        } else {
            const int j = c - 6;
            const int kj = m.link_of_dof[j];
            double mc[3] = {0, 0, 0};
            for (int k = 1; k < m.nl; k++) {
                if ((m.anc[k] >> j) & 1ull) {
                    mc[0] += w.mcw[3 * k], mc[1] += w.mcw[3 * k + 1], mc[2] += w.mcw[3 * k + 2];
                }
            }
            const double ms = m.msub[j];
            const double d[3] = {(mc[0] - ms * w.linkP[3 * kj]) * m.inv_total_mass,
                                 (mc[1] - ms * w.linkP[3 * kj + 1]) * m.inv_total_mass,
                                 (mc[2] - ms * w.linkP[3 * kj + 2]) * m.inv_total_mass};
            const double a[3] = {w.axisW[3 * kj], w.axisW[3 * kj + 1], w.axisW[3 * kj + 2]};
            cross3(a, d, lin);
        }
#pragma unroll
        for (int i = 0; i < 3; i++) dst[i * ld + c] = lin[i];
    }
}

// Angular rows of the centroidal momentum matrix in MIXED representation (iDynTree getCentroidalTotalMomentumJacobian,
// rows 3..5; AngularMomentumTask.cpp:137-185): L = sum_k [ I_k^W w_k + m_k (c_k - c) x cdot_k ] about the CoM c.
// Column of a DoF with world axis a through o: every link k distal to it turns with w_k = a, cdot_k = a x (c_k - o);
// base angular columns: all links, a = e_i through p_B; base linear columns: sum m_k (c_k - c) x e_i = 0.
__device__ __forceinline__ void k1_angmom(const DevModel& m, const WM& w, const double* com, double* dst, int ld, int lane)
{
    const int nv = m.n + 6;
    for (int c = lane; c < nv; c += 32) {
        double col[3] = {0.0, 0.0, 0.0};
        if (c >= 3) {
            double a[3] = {0.0, 0.0, 0.0}, o[3];
            int j = -1;
            if (c < 6) {
                a[c - 3] = 1.0;
                o[0] = w.linkP[0], o[1] = w.linkP[1], o[2] = w.linkP[2];
            } else {
                j = c - 6;
                const int kj = m.link_of_dof[j];
#pragma unroll
                for (int i = 0; i < 3; i++) a[i] = w.axisW[3 * kj + i], o[i] = w.linkP[3 * kj + i];
            }
            for (int k = 0; k < m.nl; k++) {
                if (j >= 0 && !((m.anc[k] >> j) & 1ull)) continue;
                const double mk = m.mass[k];
                const double* R = w.linkR + 9 * k;
                if (m.inertia) {
                    const double* I6 = m.inertia + 6 * k;
                    double al[3], il[3], iw[3];
                    mat3T_vec(R, a, al);
                    il[0] = I6[0] * al[0] + I6[1] * al[1] + I6[2] * al[2];
                    il[1] = I6[1] * al[0] + I6[3] * al[1] + I6[4] * al[2];
                    il[2] = I6[2] * al[0] + I6[4] * al[1] + I6[5] * al[2];
                    mat3_vec(R, il, iw);
                    col[0] += iw[0], col[1] += iw[1], col[2] += iw[2];
                }
                if (mk > 0.0) {
                    const double inv = 1.0 / mk;
                    const double ck[3] = {w.mcw[3 * k] * inv, w.mcw[3 * k + 1] * inv, w.mcw[3 * k + 2] * inv};
                    const double r[3] = {ck[0] - o[0], ck[1] - o[1], ck[2] - o[2]};
                    const double d[3] = {ck[0] - com[0], ck[1] - com[1], ck[2] - com[2]};
                    double v[3], t[3];
                    cross3(a, r, v);
                    cross3(d, v, t);
                    col[0] += mk * t[0], col[1] += mk * t[1], col[2] += mk * t[2];
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 3; i++) dst[i * ld + c] = col[i];
    }
}

// Record of task `tk` for `robot`: from a packed input record (full or compact layout) or from the task's own buffer.
__device__ __forceinline__ const double* task_record(const TickParams& P, const DevTask& tk, long long robot)
{
    // a batch-shared set-point (blfik_set_setpoint_shared) is not part of the packed records: pk_off < 0
    if (P.packed && tk.pk_off >= 0) return P.packed + (size_t)robot * P.packed_len + (P.compact ? tk.pk_off_c : tk.pk_off);
    return tk.setpoint ? tk.setpoint + robot * tk.sp_stride : nullptr;
}
// the record of this task comes from a packed array in the compact layout
__device__ __forceinline__ bool task_compact(const TickParams& P, const DevTask& tk) { return P.packed && P.compact && tk.pk_off >= 0; }
// Compact records carry only the pose part of a set-point (the reference's setSetPoint(pose) with its default zero
// velocity); `out` receives the full dense record (<= 13 doubles) with the feed-forward entries zeroed.
__device__ __forceinline__ void expand_compact(const DevTask& tk, const double* rec, double* out)
{
    const int full = tk.fb_off;   // length of the full set-point
    for (int c = 0; c < full && c < 13; c++) out[c] = c < tk.base_c ? rec[c] : 0.0;
}

// Set-point of a task at time t when an on-device trajectory is attached (the set-point producers
// that feed setSetPoint): x_d(t) = x_0 + A sin(w t + phi_i), feed-forward += A w cos(w t + phi_i); orientations move by
// the world-frame rotation Exp(A_ang sin(.)) about a fixed axis, whose angular velocity is exactly A_ang w cos(.).
// `spl` receives the full record (13 / 7 / 6 doubles); returns the pointer the control law should read.
static __device__ __noinline__ const double* traj_eval_slow(const DevTask& tk, long long robot, double t, const double* sp, double* spl);
// The generators are rare and long: the common case (none attached) stays a test and a branch, the evaluation itself is an
// out-of-line call so that its registers do not weigh on the tick kernels.
__device__ __forceinline__ const double* traj_eval(const DevTask& tk, long long robot, double t, const double* sp, double* spl)
{
    if (tk.gen == 0 || !sp) return sp;
    return traj_eval_slow(tk, robot, t, sp, spl);
}
// Linear part of the set-point from the task's quintic spline: position = the spline, feed-forward = stored + its derivative.
// Outside the knots the end states hold, derivative included (Spline.h:607-625).
__device__ __forceinline__ void spline_segment(const DevTask& tk, double t, int& seg, double& tl)
{
    const int nk = tk.sp_nk;
    seg = 0, tl = 0.0;
    if (t >= tk.sp_times[nk - 1]) {
        seg = nk - 2, tl = tk.sp_times[nk - 1] - tk.sp_times[nk - 2];
    } else if (t >= tk.sp_times[0]) {
        int lo = 0, hi = nk - 1;          // times[lo] <= t < times[hi]
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (t >= tk.sp_times[mid]) lo = mid; else hi = mid;
        }
        seg = lo, tl = t - tk.sp_times[lo];
    }
}
// JointTrackingTask set-point of joint j at time t (JointTrackingTask.cpp:162-185: position and feed-forward velocity): the
// stored record, or -- with a joint-space spline attached (Math::QuinticSpline<VectorXd> / CubicSpline, whose output the
// reference's callers hand to setSetPoint every period) -- the position from the spline and the stored velocity plus the
// spline's derivative.  Compact packed records carry no feed-forward velocity.
__device__ __forceinline__ void joint_setpoint(const DevTask& tk, long long robot, double t, const double* sp, bool compact, int n, int j,
                                               double& pos, double& vel)
{
    pos = sp[j], vel = compact ? 0.0 : sp[n + j];
    if (tk.gen & 4) {
        int seg;
        double tl;
        spline_segment(tk, t, seg, tl);
        const double* c = tk.sp_coef + ((size_t)robot * (tk.sp_nk - 1) + seg) * 6 * n + j;
        const double c1 = c[n], c2 = c[2 * n], c3 = c[3 * n], c4 = c[4 * n], c5 = c[5 * n];
        pos = c[0] + tl * (c1 + tl * (c2 + tl * (c3 + tl * (c4 + tl * c5))));
        vel += c1 + tl * (2.0 * c2 + tl * (3.0 * c3 + tl * (4.0 * c4 + tl * 5.0 * c5)));
    }
}
__device__ __forceinline__ void spline_eval(const DevTask& tk, long long robot, double t, const double* sp, double* spl, int vo)
{
    int seg;
    double tl;
    spline_segment(tk, t, seg, tl);
    const double* c = tk.sp_coef + ((size_t)robot * (tk.sp_nk - 1) + seg) * 18;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const double c1 = c[3 + k], c2 = c[6 + k], c3 = c[9 + k], c4 = c[12 + k], c5 = c[15 + k];
        spl[k] = c[k] + tl * (c1 + tl * (c2 + tl * (c3 + tl * (c4 + tl * c5))));
        spl[vo + k] = sp[vo + k] + c1 + tl * (2.0 * c2 + tl * (3.0 * c3 + tl * (4.0 * c4 + tl * 5.0 * c5)));
    }
}
static __device__ __noinline__ const double* traj_eval_slow(const DevTask& tk, long long robot, double t, const double* sp, double* spl)
{
    if ((tk.gen & 4) && !(tk.gen & 2)) {
        for (int c = 0; c < tk.fb_off && c < 13; c++) spl[c] = sp[c];
        spline_eval(tk, robot, t, sp, spl, tk.type == T_SE3 ? 7 : 3);
        return spl;
    }
    if (tk.gen & 2) {
        // s(tau) = 10 tau^3 - 15 tau^4 + 6 tau^5 (SO3Planner.tpp:96-105 with zero boundary conditions), time clamped to the
        // segment as SO3Planner::advance does (:289-292): x(t) = x0 + s (x1 - x0), R(t) = Exp(s Log(R1 R0^-1)) R0 -- the
        // inertial-frame ("Right trivialization") form that matches the MIXED angular velocity, w(t) = sdot Log(R1 R0^-1)
        const int base = tk.fb_off;   // set-point length
        for (int c = 0; c < base && c < 13; c++) spl[c] = sp[c];
        const double* tg = tk.mj_target + (size_t)robot * base;
        double tau = (t - tk.mj_t0) / tk.mj_T;
        tau = fmin(fmax(tau, 0.0), 1.0);
        const double t2 = tau * tau, t3 = t2 * tau;
        const double sv = t3 * (10.0 - 15.0 * tau + 6.0 * t2);
        const double sd = (tau > 0.0 && tau < 1.0) ? 30.0 * t2 * (1.0 - 2.0 * tau + t2) / tk.mj_T : 0.0;
        const bool lin = tk.type == T_SE3 || tk.type == T_COM || tk.type == T_R3;
        const int qo = tk.type == T_SE3 ? 3 : 0, vo = tk.type == T_SE3 ? 7 : 3, wo = tk.type == T_SE3 ? 10 : 4;
        if (lin && (tk.gen & 4)) {   // the spline moves the origin, this generator the orientation (swing foot)
            spline_eval(tk, robot, t, sp, spl, vo);
        } else if (lin) {
#pragma unroll
            for (int c = 0; c < 3; c++) {
                const double d = tg[c] - sp[c];
                spl[c] = sp[c] + sv * d, spl[vo + c] = sp[vo + c] + sd * d;
            }
        }
        if (tk.type == T_SE3 || tk.type == T_SO3) {
            const double q0[4] = {sp[qo], sp[qo + 1], sp[qo + 2], sp[qo + 3]}, q0i[4] = {-q0[0], -q0[1], -q0[2], q0[3]};
            const double q1[4] = {tg[qo], tg[qo + 1], tg[qo + 2], tg[qo + 3]};
            double qe[4], dist[3], dq[4], qn[4];
            quat_mul(q1, q0i, qe);
            so3_log(qe, dist);
            const double ph[3] = {sv * dist[0], sv * dist[1], sv * dist[2]};
            so3_exp(ph, dq);
            quat_mul(dq, q0, qn);
#pragma unroll
            for (int c = 0; c < 4; c++) spl[qo + c] = qn[c];
#pragma unroll
            for (int c = 0; c < 3; c++) spl[wo + c] = sp[wo + c] + sd * dist[c];
        }
        return spl;
    }
    if (!(tk.gen & 1)) return sp;
    for (int c = 0; c < tk.sp_len && c < 13; c++) spl[c] = sp[c];
    double sn, cs;
    sincos(tk.traj_omega * t + tk.traj_phase[robot], &sn, &cs);
    const double wc = tk.traj_omega * cs;
    const bool lin = tk.type == T_SE3 || tk.type == T_COM || tk.type == T_R3;
    const int qo = tk.type == T_SE3 ? 3 : 0, vo = tk.type == T_SE3 ? 7 : 3, wo = tk.type == T_SE3 ? 10 : 4;
    if (lin) {
#pragma unroll
        for (int c = 0; c < 3; c++) spl[c] += tk.traj_amp[c] * sn, spl[vo + c] += tk.traj_amp[c] * wc;
    }
    if (tk.type == T_SE3 || tk.type == T_SO3) {
        const double phi[3] = {tk.traj_amp[3] * sn, tk.traj_amp[4] * sn, tk.traj_amp[5] * sn};
        double dq[4], q0[4] = {spl[qo], spl[qo + 1], spl[qo + 2], spl[qo + 3]}, qn[4];
        so3_exp(phi, dq);
        quat_mul(dq, q0, qn);
#pragma unroll
        for (int c = 0; c < 4; c++) spl[qo + c] = qn[c];
#pragma unroll
        for (int c = 0; c < 3; c++) spl[wo + c] += tk.traj_amp[3 + c] * wc;
    }
    return spl;
}

// ------------------------------------------------------------------ K2: task assembly
// Builds, in shared memory, what QPInverseKinematics::advance builds (QPInverseKinematics.cpp:442-491):
// H, g from the priority-1 tasks; the dense equality rows (Cs, b) from the priority-0 SE3/CoM/SO3/R3 tasks;
// and, because every inequality row of the reference is +-e_j' (JointLimitsTask.cpp:110-113,
// JointVelocityLimitsTask.cpp), the intersection of those rows as a per-joint box [lo, hi].
// A priority-0 JointTrackingTask (rows e_j' x = b_j) becomes lo = hi = b_j.
// ROWS = true (the equality-only kernel, blfik_eq.cuh) keeps the problem in factored form instead of forming H: every dense
// row, priority 0 or 1, goes to row tk.brow + r of w.Cs (cost rows scaled by the square root of their weight, w.bd marks
// them with 1), its right-hand side to w.b, and w.Hs is the vector of the diagonal cost terms (JointTrackingTask weights).
template <bool ROWS = false>
__device__ __forceinline__ void k2_assemble(const TickParams& P, const WM& w, long long robot, const double* s, int lane, double tnow)
{
    const DevModel& m = P.m;
    const Layout& L = P.L;
    const int nv = L.nv, LD = L.LD, n = m.n;
    for (int i = lane; i < nv; i += 32) {
        if (ROWS) w.Hs[i] = 0.0;
        else
            for (int j = 0; j < nv; j++) w.Hs[i * LD + j] = 0.0;
        w.g[i] = 0.0, w.lo[i] = -kInf, w.hi[i] = kInf;
    }
    __syncwarp();
    for (int t = 0; t < P.ntasks; t++) {
        const DevTask& tk = P.tasks[t];
        const double* sp = task_record(P, tk, robot);
        const bool compact = task_compact(P, tk);
        if (is_dense_type(tk.type)) {
            double bt[6], spl[13], spx[13];
            int src[6];
            int rows = 0;
            // the exogenous feedback sits behind the set-point: re-based so that sp0[tk.fb_off + c] addresses it in both layouts
            const double* sp0 = compact ? sp - (tk.fb_off - tk.base_c) : sp;
            if (compact) {
                expand_compact(tk, sp, spx);
                sp = spx;
            }
            sp = traj_eval(tk, robot, tnow, sp, spl);
            if (tk.type == T_ANGMOM) {   // AngularMomentumTask.cpp:137-185: A = masked rows of the angular CMM, b = desired momentum
                double com[3];
                k1_com(m, w, com, nullptr, LD, lane);
                k1_angmom(m, w, com, w.Jt, LD, lane);
                for (int c = 0; c < 3; c++)
                    if (tk.mask[c]) {
                        bt[rows] = sp[c];
                        src[rows++] = c;
                    }
            } else if (tk.type == T_COM) {
                double com[3];
                k1_com(m, w, com, w.Jt, LD, lane);
                if (tk.exo_pos)   // measured CoM instead of the model's (CoMTask.cpp:192-195)
                    for (int c = 0; c < 3; c++) com[c] = sp0[tk.fb_off + c];
                for (int c = 0; c < 3; c++)
                    if (tk.mask[c]) {  // CoMTask.cpp:194-203: b = v_ff + kp (p_d - p_com)
                        bt[rows] = sp[3 + c] + tk.kp_lin[c] * (sp[c] - com[c]);
                        src[rows++] = c;
                    }
            } else {
                double R[9], pf[3];
                k1_frame_pose(m, w, tk.frame, R, pf);
                k1_frame_jac(m, w, m.frame_link[tk.frame], pf, w.Jt, LD, lane);
                if (tk.type == T_DISTANCE) {   // DistanceTask.cpp:178-229: A = d' J_lin / max(thr, |d|), b = kp (d_des - |d|)
                    double pr[3] = {0.0, 0.0, 0.0};
                    if (tk.flink2 >= 0) {       // relative to a reference frame: the row is u'(J_lin,target - J_lin,reference)
                        double off[3];
                        mat3_vec(w.linkR + 9 * tk.flink2, tk.Fp2, off);
                        for (int c = 0; c < 3; c++) pr[c] = w.linkP[3 * tk.flink2 + c] + off[c];
                        k1_frame_jac(m, w, tk.flink2, pr, w.A, LD, lane);   // w.A is free until the QP setup
                    }
                    const double dv[3] = {pf[0] - pr[0], pf[1] - pr[1], pf[2] - pr[2]};
                    const double d = sqrt(dv[0] * dv[0] + dv[1] * dv[1] + dv[2] * dv[2]);
                    const double inv = 1.0 / fmax(d, tk.thr);
                    __syncwarp();
                    for (int c = lane; c < nv; c += 32) {
                        double v = 0.0;
                        for (int k = 0; k < 3; k++) v += dv[k] * (w.Jt[k * LD + c] - (tk.flink2 >= 0 ? w.A[k * LD + c] : 0.0));
                        w.Jt[5 * LD + c] = v * inv;   // row 5 (an unused angular row) carries the distance row
                    }
                    bt[0] = tk.kp_lin[0] * (sp[0] - d);
                    src[0] = 5, rows = 1;
                }
                if (tk.type == T_GRAVITY) {    // GravityTask.cpp:149-185
                    double zb[3] = {sp[0], sp[1], sp[2]}, fb[3] = {sp[3], sp[4], sp[5]}, z[3], ff[3];
                    const double zn = rsqrt(zb[0] * zb[0] + zb[1] * zb[1] + zb[2] * zb[2]);
                    zb[0] *= zn, zb[1] *= zn, zb[2] *= zn;
                    mat3_vec(R, zb, z);
                    mat3_vec(R, fb, ff);
                    bt[0] = ff[0] + tk.kp_lin[0] * z[1], bt[1] = ff[1] - tk.kp_lin[0] * z[0];
                    src[0] = 3, src[1] = 4, rows = 2;
                }
                if (tk.type == T_SE3 || tk.type == T_R3) {
                    const int voff = tk.type == T_SE3 ? 7 : 3;
                    double pc[3] = {pf[0], pf[1], pf[2]};
                    if (tk.type == T_SE3 && tk.exo_pos)   // SE3Task.cpp:266-269
                        for (int c = 0; c < 3; c++) pc[c] = sp0[tk.fb_off + c];
                    for (int c = 0; c < 3; c++)
                        if (tk.mask[c]) {  // SE3Task.cpp:276-316
                            bt[rows] = tk.enabled ? sp[voff + c] + tk.kp_lin[c] * (sp[c] - pc[c]) : sp[voff + c];
                            src[rows++] = c;
                        }
                }
                if (tk.type == T_SE3 || tk.type == T_SO3) {
                    const double* qd = tk.type == T_SE3 ? sp + 3 : sp;
                    const double* wff = tk.type == T_SE3 ? sp + 10 : sp + 4;
                    double qdl[4] = {qd[0], qd[1], qd[2], qd[3]}, wl[3] = {wff[0], wff[1], wff[2]}, out[3];
                    if (tk.type == T_SE3 && tk.exo_ori) {   // SE3Task.cpp:261-265: measured orientation
                        const double qf[4] = {sp0[tk.fb_off + 3], sp0[tk.fb_off + 4], sp0[tk.fb_off + 5], sp0[tk.fb_off + 6]};
                        quat_to_rot(qf, R);
                    }
                    so3_control(R, qdl, wl, tk.kp_ang, tk.type == T_SO3 ? true : (tk.enabled != 0), out);
                    for (int c = 0; c < 3; c++) bt[rows] = out[c], src[rows++] = 3 + c;
                }
            }
            __syncwarp();
            if (ROWS) {
                for (int r = 0; r < rows; r++) {
                    const double sc = tk.priority == 0 ? 1.0 : sqrt(tk.weight[r]);
                    for (int c = lane; c < nv; c += 32) w.Cs[(tk.brow + r) * LD + c] = sc * w.Jt[src[r] * LD + c];
                    if (lane == 0) w.b[tk.brow + r] = sc * bt[r], w.bd[tk.brow + r] = tk.priority == 0 ? 0.0 : 1.0;
                }
            } else if (tk.priority == 0) {
                for (int r = 0; r < rows; r++) {
                    for (int c = lane; c < nv; c += 32) w.Cs[(tk.row_offset + r) * LD + c] = w.Jt[src[r] * LD + c];
                    if (lane == 0) w.b[tk.row_offset + r] = bt[r];
                }
            } else {
                for (int i = lane; i < nv; i += 32) {
                    double gi = 0.0;
                    for (int r = 0; r < rows; r++) {
                        const double wa = tk.weight[r] * w.Jt[src[r] * LD + i];
                        gi += wa * bt[r];
                        for (int j = 0; j < nv; j++) w.Hs[i * LD + j] += wa * w.Jt[src[r] * LD + j];
                    }
                    w.g[i] -= gi;
                }
            }
            __syncwarp();
        } else {
            for (int j = lane; j < n; j += 32) {
                const int i = 6 + j;
                if (tk.type == T_JOINT_TRACKING) {  // JointTrackingTask.cpp:151
                    double pj, vj;
                    joint_setpoint(tk, robot, tnow, sp, compact, n, j, pj, vj);
                    const double bj = vj + tk.kp_joint[j] * (pj - s[j]);
                    if (tk.priority == 1) {
                        w.Hs[ROWS ? i : i * LD + i] += tk.weight[j];
                        w.g[i] -= tk.weight[j] * bj;
                    } else {
                        w.lo[i] = fmax(w.lo[i], bj), w.hi[i] = fmin(w.hi[i], bj);
                    }
                } else if (tk.type == T_JOINT_LIMITS) {  // JointLimitsTask.cpp:227-231
                    w.hi[i] = fmin(w.hi[i], tk.klim[j] * (tk.upper[j] - s[j]) * tk.inv_dt);
                    w.lo[i] = fmax(w.lo[i], tk.klim[j] * (tk.lower[j] - s[j]) * tk.inv_dt);
                } else {  // JointVelocityLimitsTask.cpp:140-141
                    w.hi[i] = fmin(w.hi[i], tk.upper[j]);
                    w.lo[i] = fmax(w.lo[i], tk.lower[j]);
                }
            }
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------ K3: QP
// In-place lower Cholesky of the n x n matrix M (row stride ld), rows by lane.  Every lane also
// recomputes the pivot redundantly so that one __syncwarp per column suffices.  Pivots below
// `floor_` are clamped (rank-deficient equality rows behave like OSQP's delta-regularised ones).
// Returns false if a pivot is not finite.
// drop = true: a row whose pivot falls below the floor depends on the previous ones; it is left out (zero column, zero
// diagonal) and the callers skip it.
__device__ __forceinline__ bool chol_inplace(double* M, int n, int ld, double floor_, int lane, bool drop = false)
{
    bool ok = true;
    for (int k = 0; k < n; k++) {
        double piv = M[k * ld + k];
        for (int j = 0; j < k; j++) piv -= M[k * ld + j] * M[k * ld + j];
        bool dropped = false;
        if (!(piv > floor_)) {
            if (!(piv == piv)) ok = false;
            piv = floor_;
            dropped = drop;
        }
        const double d = dropped ? 0.0 : sqrt(piv), inv = dropped ? 0.0 : 1.0 / d;
        double vals[2];
        int cnt = 0;
        for (int i = lane; i < n; i += 32, cnt++) {
            double v = 0.0;
            if (i > k) {
                v = M[i * ld + k];
                for (int j = 0; j < k; j++) v -= M[i * ld + j] * M[k * ld + j];
                v *= inv;
            }
            vals[cnt & 1] = v;
        }
        __syncwarp();
        cnt = 0;
        for (int i = lane; i < n; i += 32, cnt++) {
            if (i > k) M[i * ld + k] = vals[cnt & 1];
            else if (i == k) M[k * ld + k] = d;
            else M[i * ld + k] = 0.0;  // upper triangle explicitly zero
        }
        __syncwarp();
    }
    return ok;
}

// One factorisation ("setup") of the equality-projected proximal operator
//     x~ = argmin 1/2 x'Px - r'x  s.t.  C x = b,      P = H + sigma I + diag(rho)
// as x~ = N r + xb with N = P^-1 - W W',  W = P^-1 C' Ls^-T,  S = C P^-1 C' = Ls Ls',  xb = W Ls^-1 b.
// polish = true factors the active-set-reduced problem instead: rows/columns of the active
// joints are decoupled (H -> identity there, C columns zeroed) and there is no rho.
// Leaves N in w.A, W in w.Wm, Ls in w.Ss, Ls^-1 b in w.wb, xb in w.xb.
__device__ __forceinline__ bool qp_setup(const Layout& L, const WM& w, double sigma, const double* rho_vec,
                                         bool polish, const double* bvec, int lane)
{
    const int nv = L.nv, me = L.me, LD = L.LD, LDE = L.LDE;
    bool ok = true;
    // S1: P
    for (int i = lane; i < nv; i += 32) {
        const bool ai = polish && w.act[i];
        for (int j = 0; j < nv; j++) {
            double v = w.Hs[i * LD + j];
            if (polish && (ai || w.act[j])) v = (i == j) ? 1.0 : 0.0;
            if (i == j) v += sigma + (polish ? 0.0 : rho_vec[i]);
            w.A[i * LD + j] = v;
        }
    }
    __syncwarp();
    // S2: P = L L'
    ok = chol_inplace(w.A, nv, LD, 1e-300, lane) && ok;
    // S3: B = L^-1 (lower), column c by lane c, uniform loops
    for (int c = lane; c < nv; c += 32) {
        for (int i = 0; i < nv; i++) {
            double v = (i == c) ? 1.0 : 0.0;
            for (int j = 0; j < i; j++) v -= w.A[i * LD + j] * w.B[j * LD + c];
            w.B[i * LD + c] = (i >= c) ? v / w.A[i * LD + i] : 0.0;
        }
    }
    __syncwarp();
    if (me > 0) {
        // S4: Y = L^-1 C'   (nv x me)
        for (int i = lane; i < nv; i += 32) {
            for (int r = 0; r < me; r++) {
                double v = 0.0;
                for (int j = 0; j <= i; j++) {
                    const double cj = (polish && w.act[j]) ? 0.0 : w.Cs[r * LD + j];
                    v += w.B[i * LD + j] * cj;
                }
                w.Ym[i * LDE + r] = v;
            }
        }
        __syncwarp();
        // S5: S = Y'Y (me x me)
        double smax = 0.0;
        for (int p = lane; p < me * me; p += 32) {
            const int r = p / me, c = p - r * me;
            double v = 0.0;
            for (int i = 0; i < nv; i++) v += w.Ym[i * LDE + r] * w.Ym[i * LDE + c];
            w.Ss[r * LDE + c] = v;
            if (r == c) smax = fmax(smax, v);
        }
        smax = warp_max(smax);
        __syncwarp();
        // S6: S = Ls Ls'
        ok = chol_inplace(w.Ss, me, LDE, 1e-11 * smax + 1e-300, lane, true) && ok;   // dependent equality rows are dropped
        // S7: V = Y Ls^-T, row i by lane i (in place)
        for (int i = lane; i < nv; i += 32) {
            for (int r = 0; r < me; r++) {
                double v = w.Ym[i * LDE + r];
                for (int c = 0; c < r; c++) v -= w.Ss[r * LDE + c] * w.Ym[i * LDE + c];
                const double dg = w.Ss[r * LDE + r];
                w.Ym[i * LDE + r] = dg != 0.0 ? v / dg : 0.0;
            }
        }
        __syncwarp();
        // S8: W = L^-T V  (nv x me), row j by lane j
        for (int j = lane; j < nv; j += 32) {
            for (int r = 0; r < me; r++) {
                double v = 0.0;
                for (int i = j; i < nv; i++) v += w.B[i * LD + j] * w.Ym[i * LDE + r];
                w.Wm[j * LDE + r] = v;
            }
        }
        __syncwarp();
    }
    // S9: N = L^-T L^-1 - W W'  (into w.A, row a by lane a)
    for (int a = lane; a < nv; a += 32) {
        for (int c = 0; c < nv; c++) {
            double v = 0.0;
            const int i0 = a > c ? a : c;
            for (int i = i0; i < nv; i++) v += w.B[i * LD + a] * w.B[i * LD + c];
            for (int r = 0; r < me; r++) v -= w.Wm[a * LDE + r] * w.Wm[c * LDE + r];
            w.A[a * LD + c] = v;
        }
    }
    // wb = Ls^-1 b (forward substitution, one entry finalised per step)
    for (int r = lane; r < me; r += 32) w.wb[r] = bvec[r];
    __syncwarp();
    // a dropped row must also be consistent: what is left of its right-hand side has to vanish (else status 2)
    double bmax = 0.0;
    for (int r = lane; r < me; r += 32) bmax = fmax(bmax, fabs(bvec[r]));
    const double btol = 1e-8 * (1.0 + warp_max(bmax));
    bool inconsistent = false;
    for (int c = 0; c < me; c++) {
        if (lane == (c & 31)) {
            const double dg = w.Ss[c * LDE + c];
            if (dg == 0.0 && fabs(w.wb[c]) > btol) inconsistent = true;
            w.wb[c] = dg != 0.0 ? w.wb[c] / dg : 0.0;
        }
        __syncwarp();
        const double wc = w.wb[c];
        for (int r = lane; r < me; r += 32)
            if (r > c) w.wb[r] -= w.Ss[r * LDE + c] * wc;
        __syncwarp();
    }
    for (int i = lane; i < nv; i += 32) {
        double v = 0.0;
        for (int r = 0; r < me; r++) v += w.Wm[i * LDE + r] * w.wb[r];
        w.xb[i] = v;
    }
    __syncwarp();
    if (!polish && me > 0) {
        // Every iterate is N r + x_b with C N = 0, so C x_b - b is the equality residual of whatever the solve returns.  A
        // dependent row whose pivot happened to stay above the floor (rounding noise) slips through the test above: this
        // one catches contradicting rows whatever the factorisation made of them.
        for (int r = lane; r < me; r += 32) {
            double res = -bvec[r];
            for (int j = 0; j < nv; j++) res += w.Cs[r * LD + j] * w.xb[j];
            if (!(fabs(res) <= 1e-6 * (1.0 + fabs(bvec[r])))) inconsistent = true;
        }
    }
    if (__any_sync(kFull, inconsistent)) ok = false;
    return ok;
}

// y = N r + xb for this lane's rows (r, xb, N in shared memory)
__device__ __forceinline__ double apply_N(const Layout& L, const WM& w, int i)
{
    double v = w.xb[i];
    for (int j = 0; j < L.nv; j++) v += w.A[i * L.LD + j] * w.r[j];
    return v;
}

struct QpResult {
    int status, iters;
};

// The QP of one tick:  min 1/2 x'Hx + g'x  s.t.  Cs x = b,  lo <= x_J <= hi.
// ADMM (OSQP's iteration, with the equality rows projected exactly inside the x-update) predicts the
// active set; the polish solves the active-set-reduced equality QP and is accepted only with a KKT
// certificate (primal box feasibility + dual signs + stationarity), which makes the returned point the
// exact optimum; otherwise ADMM continues.  Without box rows the problem is a pure equality QP and a few
// proximal steps solve it.
__device__ __forceinline__ QpResult k3_solve(const TickParams& P, const WM& w, bool warm, int lane)
{
    const Layout& L = P.L;
    const Settings& st = P.st;
    const int nv = L.nv, me = L.me, LD = L.LD, LDE = L.LDE;
    QpResult res = {1, 0};

    // per-variable step: 0 (unbounded), rho (box), 1e3 rho (lo == hi, OSQP's equality-row scaling)
    bool any_box = false, infeasible = false;
    double gmax = 0.0;
    for (int i = lane; i < nv; i += 32) {
        const double lo = w.lo[i], hi = w.hi[i];
        double rv = 0.0;
        if (lo > -kInfHalf || hi < kInfHalf) rv = (lo == hi) ? 1e3 * st.rho : st.rho;
        if (lo > hi) infeasible = true;
        w.t[i] = rv;  // rho_vec lives in w.t during ADMM
        any_box |= rv > 0.0;
        gmax = fmax(gmax, fabs(w.g[i]));
        if (!warm) {
            w.x[i] = 0.0, w.z[i] = 0.0, w.y[i] = 0.0;
        }
        if (rv > 0.0) w.z[i] = fmin(fmax(w.z[i], lo), hi);
        else w.z[i] = 0.0, w.y[i] = 0.0;
    }
    any_box = __any_sync(kFull, any_box);
    infeasible = __any_sync(kFull, infeasible);
    gmax = warp_max(gmax);
    __syncwarp();
    if (infeasible) {
        res.status = 2;
        return res;
    }

    if (!any_box) {
        // pure equality QP: proximal-point iterations x <- argmin 1/2x'Hx+g'x+sigma/2|x-x_k|^2 s.t. Cx=b
        if (!qp_setup(L, w, st.sigma, w.t, false, w.b, lane)) {
            res.status = 2;
            return res;
        }
        double dx = 0.0;
        for (int it = 0; it < st.eq_iters; it++) {
            for (int i = lane; i < nv; i += 32) w.r[i] = st.sigma * w.x[i] - w.g[i];
            __syncwarp();
            dx = 0.0;
            double xn[2];
            int cnt = 0;
            for (int i = lane; i < nv; i += 32, cnt++) {
                xn[cnt & 1] = apply_N(L, w, i);
                dx = fmax(dx, fabs(xn[cnt & 1] - w.x[i]));
            }
            __syncwarp();
            cnt = 0;
            for (int i = lane; i < nv; i += 32, cnt++) w.x[i] = xn[cnt & 1];
            __syncwarp();
            res.iters++;
        }
        dx = warp_max(dx);
        res.status = (dx == dx) ? ((st.sigma * dx <= st.eps * (1.0 + gmax)) ? 0 : 1) : 2;
        return res;
    }

    for (int round = 0; round < st.max_rounds; round++) {
        // ---- ADMM phase
        for (int i = lane; i < nv; i += 32) {
            const double lo = w.lo[i], hi = w.hi[i];
            w.t[i] = (lo > -kInfHalf || hi < kInfHalf) ? ((lo == hi) ? 1e3 * st.rho : st.rho) : 0.0;
        }
        __syncwarp();
        if (!qp_setup(L, w, st.sigma, w.t, false, w.b, lane)) {
            res.status = 2;
            return res;
        }
        double rp = 0.0, rd = 0.0;
        for (int it = 0; it < st.admm_iters; it++) {
            for (int i = lane; i < nv; i += 32) {
                const double rv = w.t[i];
                w.r[i] = st.sigma * w.x[i] - w.g[i] + (rv > 0.0 ? rv * w.z[i] - w.y[i] : 0.0);
            }
            __syncwarp();
            rp = 0.0, rd = 0.0;
            double xt[2];
            int cnt = 0;
            for (int i = lane; i < nv; i += 32, cnt++) xt[cnt & 1] = apply_N(L, w, i);
            __syncwarp();
            cnt = 0;
            for (int i = lane; i < nv; i += 32, cnt++) {
                const double xti = xt[cnt & 1], rv = w.t[i];
                const double xo = w.x[i];
                const double xn = st.alpha * xti + (1.0 - st.alpha) * xo;
                rd = fmax(rd, st.sigma * fabs(xn - xo));
                w.x[i] = xn;
                if (rv > 0.0) {
                    const double zo = w.z[i], yo = w.y[i];
                    const double zr = st.alpha * xti + (1.0 - st.alpha) * zo;
                    const double zn = fmin(fmax(zr + yo / rv, w.lo[i]), w.hi[i]);
                    w.y[i] = yo + rv * (zr - zn);
                    w.z[i] = zn;
                    rd = fmax(rd, rv * fabs(zn - zo));
                    rp = fmax(rp, fabs(xn - zn));
                }
            }
            __syncwarp();
            res.iters++;
        }
        rp = warp_max(rp), rd = warp_max(rd);
        if (!(rp == rp) || !(rd == rd)) {
            res.status = 2;
            return res;
        }
        const bool admm_converged = rp <= st.eps && rd <= st.eps * (1.0 + gmax);

        // ---- polish: active set by OSQP's rule (polish.c), then the reduced equality QP
        for (int i = lane; i < nv; i += 32) {
            const double rv = w.t[i];
            int a = 0;
            double bd = 0.0;
            if (rv > 0.0) {
                const double lo = w.lo[i], hi = w.hi[i], z = w.z[i], y = w.y[i];
                if (lo == hi) a = 3, bd = lo;
                else if (z - lo < -y) a = 1, bd = lo;
                else if (hi - z < y) a = 2, bd = hi;
            }
            w.act[i] = a;
            w.bd[i] = bd;
        }
        __syncwarp();
        for (int i = lane; i < nv; i += 32) {
            double v;
            if (w.act[i]) {
                v = -w.bd[i];
            } else {
                v = w.g[i];
                for (int j = 6; j < nv; j++)
                    if (w.act[j]) v += w.Hs[i * LD + j] * w.bd[j];
            }
            w.gp[i] = v;
            w.lam[i] = w.act[i] ? w.bd[i] : w.x[i];  // polish iterate xp lives in w.lam[0..nv)
        }
        for (int r = lane; r < me; r += 32) {
            double v = w.b[r];
            for (int j = 6; j < nv; j++)
                if (w.act[j]) v -= w.Cs[r * LD + j] * w.bd[j];
            w.bp[r] = v;
        }
        __syncwarp();
        bool pol_ok = qp_setup(L, w, st.sigma, w.t, true, w.bp, lane);
        double* xp = w.lam;
        for (int it = 0; it < st.polish_iters; it++) {
            for (int i = lane; i < nv; i += 32) w.r[i] = st.sigma * xp[i] - w.gp[i];
            __syncwarp();
            double xn[2];
            int cnt = 0;
            for (int i = lane; i < nv; i += 32, cnt++) xn[cnt & 1] = apply_N(L, w, i);
            __syncwarp();
            cnt = 0;
            for (int i = lane; i < nv; i += 32, cnt++) xp[i] = xn[cnt & 1];
            __syncwarp();
        }
        // multipliers of the equality rows: lambda = Ls^-T (W' r - wb), r = sigma xp - g'
        for (int i = lane; i < nv; i += 32) w.r[i] = st.sigma * xp[i] - w.gp[i];
        __syncwarp();
        double* lam = w.bp;  // reuse (b' is consumed)
        for (int r = lane; r < me; r += 32) {
            double v = -w.wb[r];
            for (int i = 0; i < nv; i++) v += w.Wm[i * LDE + r] * w.r[i];
            lam[r] = v;
        }
        __syncwarp();
        for (int c = me - 1; c >= 0; c--) {
            if (lane == (c & 31)) lam[c] = lam[c] / w.Ss[c * LDE + c];
            __syncwarp();
            const double lc = lam[c];
            for (int r = lane; r < c; r += 32) lam[r] -= w.Ss[c * LDE + r] * lc;
            __syncwarp();
        }
        // certificate: y_full = -(H xp + g + C' lambda)
        bool cert = pol_ok;
        for (int i = lane; i < nv; i += 32) {
            double v = w.g[i];
            for (int j = 0; j < nv; j++) v += w.Hs[i * LD + j] * xp[j];
            for (int r = 0; r < me; r++) v += w.Cs[r * LD + i] * lam[r];
            const double yf = -v;
            const int a = w.act[i];
            const double xi = xp[i];
            const double tolx = 1e-9 * (1.0 + fabs(xi));
            const double toly = 1e-8 * (1.0 + gmax);
            if (!(xi == xi) || !(yf == yf)) cert = false;
            if (w.t[i] > 0.0 && (xi > w.hi[i] + tolx || xi < w.lo[i] - tolx)) cert = false;
            if (a == 0 && fabs(yf) > toly) cert = false;
            if (a == 1 && yf > toly) cert = false;
            if (a == 2 && yf < -toly) cert = false;
            w.r[i] = yf;  // keep for the warm start
        }
        cert = __all_sync(kFull, cert);
        __syncwarp();
        if (cert) {
            for (int i = lane; i < nv; i += 32) {
                const double xi = xp[i];
                w.x[i] = xi;
                if (w.t[i] > 0.0) {
                    w.z[i] = fmin(fmax(xi, w.lo[i]), w.hi[i]);
                    w.y[i] = w.act[i] ? w.r[i] : 0.0;
                }
            }
            __syncwarp();
            res.status = 0;
            return res;
        }
        if (admm_converged) {
            res.status = 0;
            return res;
        }
    }
    res.status = 1;
    return res;
}

// ------------------------------------------------------------------ K4: ForwardEuler on R3 x SO(3) x R^n
// packed input records carry the state too: unpack it into the state arrays before the tick reads them
__device__ __forceinline__ void unpack_state(const TickParams& P, long long robot, double* bp, double* bq, double* s, int n, int lane)
{
    const double* rec = P.packed + (size_t)robot * P.packed_len;
    for (int i = lane; i < 7 + n; i += 32) {
        const double v = rec[i];
        if (i < 3) bp[i] = v;
        else if (i < 7) bq[i - 3] = v;
        else s[i - 7] = v;
    }
    __syncwarp();
}

// p += v dt;  R <- exp(w dt) R (manif tangent + group = left-plus, ForwardEuler.h:82);  s += sdot dt
__device__ __forceinline__ void k4_integrate(int n, double* bp, double* bq, double* s, const double* nu, double dt,
                                             int lane)
{
    if (lane < 3) bp[lane] += nu[lane] * dt;
    double phi[3] = {nu[3] * dt, nu[4] * dt, nu[5] * dt}, dq[4], q[4] = {bq[0], bq[1], bq[2], bq[3]}, qn[4];
    so3_exp(phi, dq);
    quat_mul(dq, q, qn);
    const double inv = 1.0 / sqrt(qn[0] * qn[0] + qn[1] * qn[1] + qn[2] * qn[2] + qn[3] * qn[3]);
    __syncwarp();
    if (lane < 4) bq[lane] = qn[lane] * inv;
    for (int j = lane; j < n; j += 32) s[j] += nu[6 + j] * dt;
    __syncwarp();
}

}  // namespace blfik
