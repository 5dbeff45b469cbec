/*
 * blfik_oracle.c -- CPU oracle (TEST INFRASTRUCTURE ONLY, see blfik_oracle.h).
 *
 * Plain C99, FP64, no dependencies.  "PARITY UNPINNED": see the header.
 * All file:line citations are relative to /root/reference.
 */
#include "blfik_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <pthread.h>

/* ------------------------------------------------------------------------- */
/* small helpers                                                              */
/* ------------------------------------------------------------------------- */
static void mat3_mul(const double* A, const double* B, double* C)
{
    double t[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
    memcpy(C, t, sizeof t);
}
static void mat3_vec(const double* A, const double* v, double* o)
{
    double t0 = A[0] * v[0] + A[1] * v[1] + A[2] * v[2];
    double t1 = A[3] * v[0] + A[4] * v[1] + A[5] * v[2];
    double t2 = A[6] * v[0] + A[7] * v[1] + A[8] * v[2];
    o[0] = t0, o[1] = t1, o[2] = t2;
}
static void cross3(const double* a, const double* b, double* o)
{
    double t0 = a[1] * b[2] - a[2] * b[1];
    double t1 = a[2] * b[0] - a[0] * b[2];
    double t2 = a[0] * b[1] - a[1] * b[0];
    o[0] = t0, o[1] = t1, o[2] = t2;
}
/* ------------------------------------------------------------------------- */
/* SO(3) on unit quaternions, storage (x,y,z,w) as Eigen/manif   */
/* ------------------------------------------------------------------------- */
void orc_quat_to_rot(const double q[4], double R[9])
{
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    R[0] = 1 - 2 * (y * y + z * z), R[1] = 2 * (x * y - z * w), R[2] = 2 * (x * z + y * w);
    R[3] = 2 * (x * y + z * w), R[4] = 1 - 2 * (x * x + z * z), R[5] = 2 * (y * z - x * w);
    R[6] = 2 * (x * z - y * w), R[7] = 2 * (y * z + x * w), R[8] = 1 - 2 * (x * x + y * y);
}

/* rotation matrix -> quaternion -> normalise: Conversions::toManifRot,
 * src/Conversions/include/BipedalLocomotion/Conversions/ManifConversions.h:44-50
 * (Eigen's Quaternion(Matrix3) = trace / largest-diagonal branches).          */
void orc_rot_to_quat(const double R[9], double q[4])
{
    double t = R[0] + R[4] + R[8];
    if (t > 0) {
        t = sqrt(t + 1.0);
        q[3] = 0.5 * t;
        t = 0.5 / t;
        q[0] = (R[7] - R[5]) * t;
        q[1] = (R[2] - R[6]) * t;
        q[2] = (R[3] - R[1]) * t;
    } else {
        int i = 0;
        if (R[4] > R[0]) i = 1;
        if (R[8] > R[4 * i]) i = 2;
        int j = (i + 1) % 3, k = (j + 1) % 3;
        t = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
        q[i] = 0.5 * t;
        t = 0.5 / t;
        q[3] = (R[3 * k + j] - R[3 * j + k]) * t;
        q[j] = (R[3 * j + i] + R[3 * i + j]) * t;
        q[k] = (R[3 * k + i] + R[3 * i + k]) * t;
    }
    double nrm = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int c = 0; c < 4; c++) q[c] /= nrm;
}

void orc_quat_mul(const double a[4], const double b[4], double o[4])
{
    double x = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
    double y = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
    double z = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
    double w = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
    o[0] = x, o[1] = y, o[2] = z, o[3] = w;
}

/* manif SO3Tangent::exp */
void orc_so3_exp(const double phi[3], double q[4])
{
    double th2 = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
    double th = sqrt(th2), k;
    if (th2 > 1e-16) {
        k = sin(0.5 * th) / th;
        q[3] = cos(0.5 * th);
    } else {
        k = 0.5 - th2 / 48.0;
        q[3] = 1.0 - th2 / 8.0;
    }
    q[0] = k * phi[0], q[1] = k * phi[1], q[2] = k * phi[2];
}

/* manif SO3::log on a unit quaternion: angle in (-pi, pi] through atan2, with
 * the w<0 branch folded so the returned rotation vector has norm <= pi.       */
void orc_so3_log(const double q[4], double phi[3])
{
    double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
    double w = q[3], scale;
    if (n2 > 1e-20) {
        double n = sqrt(n2);
        double two_angle = 2.0 * (w < 0 ? atan2(-n, -w) : atan2(n, w));
        scale = two_angle / n;
    } else {
        scale = 2.0 / w;
    }
    phi[0] = scale * q[0], phi[1] = scale * q[1], phi[2] = scale * q[2];
}

/* ------------------------------------------------------------------------- */
/* K1: floating-base FK, frame Jacobian, CoM (iDynTree MIXED)     */
/* call sites: SE3Task.cpp:263-268,283-296  CoMTask.cpp:194,206  SO3Task.cpp:174,181 */
/* ------------------------------------------------------------------------- */
static void axis_rot(const double a[3], double th, double R[9])
{
    double s = sin(th), c = cos(th), v = 1 - c;
    R[0] = c + a[0] * a[0] * v, R[1] = a[0] * a[1] * v - a[2] * s, R[2] = a[0] * a[2] * v + a[1] * s;
    R[3] = a[1] * a[0] * v + a[2] * s, R[4] = c + a[1] * a[1] * v, R[5] = a[1] * a[2] * v - a[0] * s;
    R[6] = a[2] * a[0] * v - a[1] * s, R[7] = a[2] * a[1] * v + a[0] * s, R[8] = c + a[2] * a[2] * v;
}

void orc_fk(const orc_model* m, const double p[3], const double quat[4], const double* s,
            double* link_R, double* link_p)
{
    orc_quat_to_rot(quat, link_R);
    memcpy(link_p, p, 3 * sizeof(double));
    for (int k = 1; k <= m->n; k++) {
        const int par = m->parent[k];
        double Rj[9], Rpj[9], off[3];
        axis_rot(m->axis + 3 * k, s[m->dof[k]], Rj);
        mat3_mul(link_R + 9 * par, m->Xr + 9 * k, Rpj);
        mat3_mul(Rpj, Rj, link_R + 9 * k);
        mat3_vec(link_R + 9 * par, m->Xp + 3 * k, off);
        for (int c = 0; c < 3; c++) link_p[3 * k + c] = link_p[3 * par + c] + off[c];
    }
}

void orc_frame_pose(const orc_model* m, const double* link_R, const double* link_p, int frame,
                    double R[9], double p[3])
{
    const int L = m->frame_link[frame];
    double off[3];
    mat3_mul(link_R + 9 * L, m->Fr + 9 * frame, R);
    mat3_vec(link_R + 9 * L, m->Fp + 3 * frame, off);
    for (int c = 0; c < 3; c++) p[c] = link_p[3 * L + c] + off[c];
}

/* 6 x nv, rows [linear of the frame origin in world; angular in world],
 * columns nu = [base linear (world); base angular (world); sdot].             */
void orc_frame_jacobian(const orc_model* m, const double* link_R, const double* link_p, int frame,
                        double* J)
{
    const int nv = 6 + m->n;
    double R[9], pf[3];
    orc_frame_pose(m, link_R, link_p, frame, R, pf);
    memset(J, 0, sizeof(double) * 6 * nv);
    const double r[3] = {pf[0] - link_p[0], pf[1] - link_p[1], pf[2] - link_p[2]};
    /* base block [I, -S(r); 0, I] */
    for (int c = 0; c < 3; c++) J[c * nv + c] = 1.0, J[(3 + c) * nv + 3 + c] = 1.0;
    J[0 * nv + 4] = r[2], J[0 * nv + 5] = -r[1];
    J[1 * nv + 3] = -r[2], J[1 * nv + 5] = r[0];
    J[2 * nv + 3] = r[1], J[2 * nv + 4] = -r[0];
    for (int k = m->frame_link[frame]; k > 0; k = m->parent[k]) {
        const int j = 6 + m->dof[k];
        double aw[3], d[3], lin[3];
        mat3_vec(link_R + 9 * k, m->axis + 3 * k, aw);
        for (int c = 0; c < 3; c++) d[c] = pf[c] - link_p[3 * k + c];
        cross3(aw, d, lin);
        for (int c = 0; c < 3; c++) J[c * nv + j] = lin[c], J[(3 + c) * nv + j] = aw[c];
    }
}

/* p_com = sum m_k c_k / M;  J_com = sum (m_k/M) J_{c_k, linear}               */
void orc_com(const orc_model* m, const double* link_R, const double* link_p, double com[3],
             double* Jcom)
{
    const int nv = 6 + m->n, nl = m->n + 1;
    double M = 0, acc[3] = {0, 0, 0};
    double* cw = (double*)malloc(sizeof(double) * 3 * nl);
    for (int k = 0; k < nl; k++) {
        double off[3];
        mat3_vec(link_R + 9 * k, m->com + 3 * k, off);
        for (int c = 0; c < 3; c++) {
            cw[3 * k + c] = link_p[3 * k + c] + off[c];
            acc[c] += m->mass[k] * cw[3 * k + c];
        }
        M += m->mass[k];
    }
    for (int c = 0; c < 3; c++) com[c] = acc[c] / M;
    if (Jcom) {
        memset(Jcom, 0, sizeof(double) * 3 * nv);
        /* per-link accumulation (the way iDynTree does it; the GPU kernel uses
         * subtree sums instead -- an independent evaluation of the same matrix) */
        for (int k = 0; k < nl; k++) {
            const double wgt = m->mass[k] / M;
            const double r[3] = {cw[3 * k] - link_p[0], cw[3 * k + 1] - link_p[1],
                                 cw[3 * k + 2] - link_p[2]};
            for (int c = 0; c < 3; c++) Jcom[c * nv + c] += wgt;
            Jcom[0 * nv + 4] += wgt * r[2], Jcom[0 * nv + 5] -= wgt * r[1];
            Jcom[1 * nv + 3] -= wgt * r[2], Jcom[1 * nv + 5] += wgt * r[0];
            Jcom[2 * nv + 3] += wgt * r[1], Jcom[2 * nv + 4] -= wgt * r[0];
            for (int a = k; a > 0; a = m->parent[a]) {
                const int j = 6 + m->dof[a];
                double aw[3], d[3], lin[3];
                mat3_vec(link_R + 9 * a, m->axis + 3 * a, aw);
                for (int c = 0; c < 3; c++) d[c] = cw[3 * k + c] - link_p[3 * a + c];
                cross3(aw, d, lin);
                for (int c = 0; c < 3; c++) Jcom[c * nv + j] += wgt * lin[c];
            }
        }
    }
    free(cw);
}

/* Angular part of the centroidal momentum matrix (iDynTree::KinDynComputations::getCentroidalTotalMomentumJacobian rows
 * 3..5 in MIXED representation; used by AngularMomentumTask.cpp:160-185), accumulated link by link from each link's own
 * Jacobian at its centre of mass:  L = sum_k [ I_k^W J_w,k + m_k S(c_k - c) J_v,k ] nu.  (The GPU kernel evaluates the same
 * matrix column by column over the links distal to each DoF.)                                                             */
void orc_angular_momentum_matrix(const orc_model* m, const double* link_R, const double* link_p, double* A)
{
    const int nv = 6 + m->n, nl = m->n + 1;
    double com[3];
    orc_com(m, link_R, link_p, com, NULL);
    memset(A, 0, sizeof(double) * 3 * nv);
    double* Jk = (double*)malloc(sizeof(double) * 6 * nv);
    for (int k = 0; k < nl; k++) {
        /* Jacobian of link k at its CoM */
        double off[3], ck[3];
        mat3_vec(link_R + 9 * k, m->com + 3 * k, off);
        for (int c = 0; c < 3; c++) ck[c] = link_p[3 * k + c] + off[c];
        memset(Jk, 0, sizeof(double) * 6 * nv);
        const double r[3] = {ck[0] - link_p[0], ck[1] - link_p[1], ck[2] - link_p[2]};
        for (int c = 0; c < 3; c++) Jk[c * nv + c] = 1.0, Jk[(3 + c) * nv + 3 + c] = 1.0;
        Jk[0 * nv + 4] = r[2], Jk[0 * nv + 5] = -r[1];
        Jk[1 * nv + 3] = -r[2], Jk[1 * nv + 5] = r[0];
        Jk[2 * nv + 3] = r[1], Jk[2 * nv + 4] = -r[0];
        for (int a = k; a > 0; a = m->parent[a]) {
            const int j = 6 + m->dof[a];
            double aw[3], d[3], lin[3];
            mat3_vec(link_R + 9 * a, m->axis + 3 * a, aw);
            for (int c = 0; c < 3; c++) d[c] = ck[c] - link_p[3 * a + c];
            cross3(aw, d, lin);
            for (int c = 0; c < 3; c++) Jk[c * nv + j] = lin[c], Jk[(3 + c) * nv + j] = aw[c];
        }
        /* world inertia about the link CoM */
        double Iw[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        if (m->inertia) {
            const double* I6 = m->inertia + 6 * k;
            const double Il[9] = {I6[0], I6[1], I6[2], I6[1], I6[3], I6[4], I6[2], I6[4], I6[5]};
            double T[9], Rt[9];
            const double* R = link_R + 9 * k;
            for (int a = 0; a < 3; a++)
                for (int b = 0; b < 3; b++) Rt[3 * a + b] = R[3 * b + a];
            mat3_mul(R, Il, T);
            mat3_mul(T, Rt, Iw);
        }
        const double d[3] = {ck[0] - com[0], ck[1] - com[1], ck[2] - com[2]};
        const double mk = m->mass[k];
        for (int c = 0; c < nv; c++) {
            const double v[3] = {Jk[c], Jk[nv + c], Jk[2 * nv + c]}, w[3] = {Jk[3 * nv + c], Jk[4 * nv + c], Jk[5 * nv + c]};
            double t[3];
            cross3(d, v, t);
            for (int i = 0; i < 3; i++) A[i * nv + c] += Iw[3 * i] * w[0] + Iw[3 * i + 1] * w[1] + Iw[3 * i + 2] * w[2] + mk * t[i];
        }
    }
    free(Jk);
}

/* ------------------------------------------------------------------------- */
/* K2: tasks                                                                  */
/* ------------------------------------------------------------------------- */
int orc_setpoint_len(int type, int n)
{
    switch (type) {
    case ORC_TASK_SE3: return 13;            /* p_d(3) q_d(4) v_ff(3) w_ff(3)   */
    case ORC_TASK_COM: return 6;             /* p_d(3) v_ff(3)                  */
    case ORC_TASK_SO3: return 7;             /* q_d(4) w_ff(3)                  */
    case ORC_TASK_R3: return 6;              /* p_d(3) v_ff(3)                  */
    case ORC_TASK_JOINT_TRACKING: return 2 * n; /* s_d(n) sdot_d(n)             */
    case ORC_TASK_DISTANCE: return 1;        /* desired distance                 */
    case ORC_TASK_GRAVITY: return 6;         /* z_des in the body (3) ff (3)     */
    case ORC_TASK_ANGULAR_MOMENTUM: return 3; /* desired centroidal angular momentum */
    default: return 0;
    }
}

int orc_task_setpoint_len(const orc_task* t, int n)
{
    int len = orc_setpoint_len(t->type, n);
    if (t->type == ORC_TASK_SE3 && (t->exo_pos || t->exo_ori)) len += 7;
    if (t->type == ORC_TASK_COM && t->exo_pos) len += 3;
    return len;
}

static int mask_count(const int* mask) { return (mask[0] != 0) + (mask[1] != 0) + (mask[2] != 0); }

int orc_task_rows(const orc_task* t, int n)
{
    switch (t->type) {
    case ORC_TASK_SE3: return mask_count(t->mask) + 3;
    case ORC_TASK_COM: return mask_count(t->mask);
    case ORC_TASK_R3: return mask_count(t->mask);
    case ORC_TASK_SO3: return 3;
    case ORC_TASK_DISTANCE: return 1;
    case ORC_TASK_GRAVITY: return 2;
    case ORC_TASK_ANGULAR_MOMENTUM: return mask_count(t->mask);
    case ORC_TASK_JOINT_TRACKING: return n;
    case ORC_TASK_JOINT_LIMITS: return 2 * n;
    case ORC_TASK_JOINT_VELOCITY_LIMITS: return 2 * n;
    default: return -1;
    }
}

int orc_task_is_inequality(const orc_task* t)
{
    return t->type == ORC_TASK_JOINT_LIMITS || t->type == ORC_TASK_JOINT_VELOCITY_LIMITS;
}

int orc_constraint_rows(const orc_task* tasks, int ntasks, int n)
{
    int m = 0;
    for (int t = 0; t < ntasks; t++)
        if (tasks[t].priority == 0) m += orc_task_rows(&tasks[t], n);
    return m;
}

/* LieGroupControllers::ProportionalControllerSO3d, left trivialisation:
 * w* = w_ff + Kp o Log(R_d R^-1)   (SE3Task.cpp:261-273, SO3Task.cpp:174-179) */
static void so3_control(const double Rcur[9], const double qd[4], const double wff[3],
                        const double kp[3], int enabled, double out[3])
{
    double qc[4], qci[4], qe[4], e[3];
    orc_rot_to_quat(Rcur, qc);
    qci[0] = -qc[0], qci[1] = -qc[1], qci[2] = -qc[2], qci[3] = qc[3];
    orc_quat_mul(qd, qci, qe);
    orc_so3_log(qe, e);
    for (int c = 0; c < 3; c++) out[c] = enabled ? wff[c] + kp[c] * e[c] : wff[c];
}

int orc_task_update(const orc_model* m, const orc_task* t, const double* link_R,
                    const double* link_p, const double* s, const double* sp, double* A, double* b)
{
    const int n = m->n, nv = 6 + n;
    const int rows = orc_task_rows(t, n);
    if (rows < 0) return -1;
    memset(A, 0, sizeof(double) * rows * nv);
    switch (t->type) {
    case ORC_TASK_SE3: { /* SE3Task.cpp:231-321 */
        double J[6 * 64], R[9], pf[3], w[3];
        double* Jd = (nv <= 64) ? J : (double*)malloc(sizeof(double) * 6 * nv);
        orc_frame_pose(m, link_R, link_p, t->frame, R, pf);
        orc_frame_jacobian(m, link_R, link_p, t->frame, Jd);
        /* exogenous feedback (SE3Task.cpp:261-268, :338-366): the controllers' state is what the caller measured (record
         * tail: p_fb(3) q_fb(4)) instead of the pose kinDyn predicts; the Jacobian is always kinDyn's */
        if (t->exo_pos)
            for (int c = 0; c < 3; c++) pf[c] = sp[13 + c];
        if (t->exo_ori) orc_quat_to_rot(sp + 16, R);
        int idx = 0;
        for (int c = 0; c < 3; c++) {
            if (!t->mask[c]) continue;
            b[idx] = t->controller_enabled ? sp[7 + c] + t->kp_lin[c] * (sp[c] - pf[c]) : sp[7 + c];
            memcpy(A + idx * nv, Jd + c * nv, sizeof(double) * nv);
            idx++;
        }
        so3_control(R, sp + 3, sp + 10, t->kp_ang, t->controller_enabled, w);
        for (int c = 0; c < 3; c++) {
            b[idx + c] = w[c];
            memcpy(A + (idx + c) * nv, Jd + (3 + c) * nv, sizeof(double) * nv);
        }
        if (Jd != J) free(Jd);
        break;
    }
    case ORC_TASK_R3: { /* R3Task.cpp:178-243 */
        double J[6 * 64], R[9], pf[3];
        double* Jd = (nv <= 64) ? J : (double*)malloc(sizeof(double) * 6 * nv);
        orc_frame_pose(m, link_R, link_p, t->frame, R, pf);
        orc_frame_jacobian(m, link_R, link_p, t->frame, Jd);
        int idx = 0;
        for (int c = 0; c < 3; c++) {
            if (!t->mask[c]) continue;
            b[idx] = t->controller_enabled ? sp[3 + c] + t->kp_lin[c] * (sp[c] - pf[c]) : sp[3 + c];
            memcpy(A + idx * nv, Jd + c * nv, sizeof(double) * nv);
            idx++;
        }
        if (Jd != J) free(Jd);
        break;
    }
    case ORC_TASK_COM: { /* CoMTask.cpp:171-235 (getControl() always: no controller mode) */
        double Jc[3 * 64], com[3];
        double* Jd = (nv <= 64) ? Jc : (double*)malloc(sizeof(double) * 3 * nv);
        orc_com(m, link_R, link_p, com, Jd);
        if (t->exo_pos) /* CoMTask.cpp:192-195, :262-273: measured CoM in the record tail */
            for (int c = 0; c < 3; c++) com[c] = sp[6 + c];
        int idx = 0;
        for (int c = 0; c < 3; c++) {
            if (!t->mask[c]) continue;
            b[idx] = sp[3 + c] + t->kp_lin[c] * (sp[c] - com[c]);
            memcpy(A + idx * nv, Jd + c * nv, sizeof(double) * nv);
            idx++;
        }
        if (Jd != Jc) free(Jd);
        break;
    }
    case ORC_TASK_SO3: { /* SO3Task.cpp:153-192, A = J.bottomRows<3>() */
        double J[6 * 64], R[9], pf[3], w[3];
        double* Jd = (nv <= 64) ? J : (double*)malloc(sizeof(double) * 6 * nv);
        orc_frame_pose(m, link_R, link_p, t->frame, R, pf);
        orc_frame_jacobian(m, link_R, link_p, t->frame, Jd);
        so3_control(R, sp, sp + 4, t->kp_ang, 1, w);
        for (int c = 0; c < 3; c++) {
            b[c] = w[c];
            memcpy(A + c * nv, Jd + (3 + c) * nv, sizeof(double) * nv);
        }
        if (Jd != J) free(Jd);
        break;
    }
    case ORC_TASK_DISTANCE: { /* DistanceTask.cpp:178-229: A = p' J_lin / max(thr, |p|), b = kp (d_des - |p|) */
        double J[6 * 64], J2[6 * 64], R[9], pt[3], pr[3] = {0, 0, 0};
        if (nv > 64) return -1;
        orc_frame_pose(m, link_R, link_p, t->frame, R, pt);
        orc_frame_jacobian(m, link_R, link_p, t->frame, J);
        if (t->ref_frame >= 0) {
            /* relative position / relative Jacobian: d/dt |p_t - p_ref| = u'(v_t - v_ref), u = (p_t - p_ref)/d */
            orc_frame_pose(m, link_R, link_p, t->ref_frame, R, pr);
            orc_frame_jacobian(m, link_R, link_p, t->ref_frame, J2);
            for (int i = 0; i < 3 * nv; i++) J[i] -= J2[i];
        }
        const double dv[3] = {pt[0] - pr[0], pt[1] - pr[1], pt[2] - pr[2]};
        const double d = sqrt(dv[0] * dv[0] + dv[1] * dv[1] + dv[2] * dv[2]);
        const double den = d > t->threshold ? d : t->threshold;
        for (int c = 0; c < nv; c++) A[c] = (dv[0] * J[c] + dv[1] * J[nv + c] + dv[2] * J[2 * nv + c]) / den;
        b[0] = t->kp_lin[0] * (sp[0] - d);
        break;
    }
    case ORC_TASK_GRAVITY: { /* GravityTask.cpp:149-185: A = [1 0 0; 0 1 0] J_ang, b = (R ff)_{0,1} - kp [0 -1 0; 1 0 0] R z_des */
        double J[6 * 64], R[9], pf[3], z[3], ff[3], zb[3] = {sp[0], sp[1], sp[2]};
        if (nv > 64) return -1;
        const double zn = sqrt(zb[0] * zb[0] + zb[1] * zb[1] + zb[2] * zb[2]);
        for (int c = 0; c < 3; c++) zb[c] /= zn; /* setDesiredGravityDirectionInTargetFrame normalises (:187-194) */
        orc_frame_pose(m, link_R, link_p, t->frame, R, pf);
        orc_frame_jacobian(m, link_R, link_p, t->frame, J);
        mat3_vec(R, zb, z);
        mat3_vec(R, sp + 3, ff);
        memcpy(A, J + 3 * nv, sizeof(double) * nv);
        memcpy(A + nv, J + 4 * nv, sizeof(double) * nv);
        b[0] = ff[0] + t->kp_lin[0] * z[1];
        b[1] = ff[1] - t->kp_lin[0] * z[0];
        break;
    }
    case ORC_TASK_ANGULAR_MOMENTUM: { /* AngularMomentumTask.cpp:137-185: A = masked rows 3..5 of the CMM, b = desired momentum (:187-210) */
        double Am[3 * 64];
        if (nv > 64) return -1;
        orc_angular_momentum_matrix(m, link_R, link_p, Am);
        int idx = 0;
        for (int c = 0; c < 3; c++) {
            if (!t->mask[c]) continue;
            b[idx] = sp[c];
            memcpy(A + idx * nv, Am + c * nv, sizeof(double) * nv);
            idx++;
        }
        break;
    }
    case ORC_TASK_JOINT_TRACKING: /* JointTrackingTask.cpp:65-67 (A=[0 I]), :151 (b) */
        for (int j = 0; j < n; j++) {
            A[j * nv + 6 + j] = 1.0;
            b[j] = sp[n + j] + t->kp_joint[j] * (sp[j] - s[j]);
        }
        break;
    case ORC_TASK_JOINT_LIMITS: /* JointLimitsTask.cpp:110-113 (A=[0 I;0 -I]), :227-231 (b) */
        for (int j = 0; j < n; j++) {
            A[j * nv + 6 + j] = 1.0;
            A[(n + j) * nv + 6 + j] = -1.0;
            b[j] = t->klim[j] * (t->upper[j] - s[j]) / t->sampling_time;
            b[n + j] = t->klim[j] * (-t->lower[j] + s[j]) / t->sampling_time;
        }
        break;
    case ORC_TASK_JOINT_VELOCITY_LIMITS: /* JointVelocityLimitsTask.cpp:140-141 */
        for (int j = 0; j < n; j++) {
            A[j * nv + 6 + j] = 1.0;
            A[(n + j) * nv + 6 + j] = -1.0;
            b[j] = t->upper[j];
            b[n + j] = -t->lower[j];
        }
        break;
    default: return -1;
    }
    return rows;
}

/* QPInverseKinematics::advance, src/IK/src/QPInverseKinematics.cpp:442-491 */
int orc_assemble(const orc_model* m, const orc_task* tasks, int ntasks,
                 const double* const* setpoints, const double p[3], const double quat[4],
                 const double* s, double* H, double* g, double* C, double* l, double* u)
{
    const int n = m->n, nv = 6 + n, nl = n + 1;
    double* link_R = (double*)malloc(sizeof(double) * 12 * nl);
    double* link_p = link_R + 9 * nl;
    int maxrows = 6;
    if (2 * n > maxrows) maxrows = 2 * n;
    double* A = (double*)malloc(sizeof(double) * (maxrows * nv + maxrows));
    double* b = A + maxrows * nv;
    orc_fk(m, p, quat, s, link_R, link_p);
    memset(H, 0, sizeof(double) * nv * nv);
    memset(g, 0, sizeof(double) * nv);
    int row = 0;
    for (int t = 0; t < ntasks; t++) {
        const orc_task* tk = &tasks[t];
        int rows = orc_task_update(m, tk, link_R, link_p, s, setpoints ? setpoints[t] : NULL, A, b);
        if (rows < 0) {
            free(A), free(link_R);
            return -1;
        }
        if (tk->priority == 1) {
            /* tmp = A' diag(w); H += tmp A; g -= tmp b  (:450-453) */
            for (int r = 0; r < rows; r++) {
                const double w = tk->weight[r];
                const double* Ar = A + r * nv;
                for (int i = 0; i < nv; i++) {
                    const double wa = w * Ar[i];
                    if (wa == 0.0) continue;
                    for (int j = 0; j < nv; j++) H[i * nv + j] += wa * Ar[j];
                    g[i] -= wa * b[r];
                }
            }
        } else {
            /* C.middleRows = A; u = b; l = b or -INFTY  (:473-488) */
            memcpy(C + row * nv, A, sizeof(double) * rows * nv);
            for (int r = 0; r < rows; r++) {
                u[row + r] = b[r];
                l[row + r] = orc_task_is_inequality(tk) ? -ORC_INFTY : b[r];
            }
            row += rows;
        }
    }
    free(A), free(link_R);
    return row;
}

/* ------------------------------------------------------------------------- */
/* K3: the QP.  OSQP v0.6.3 algorithm restated (Stellato et al., "OSQP: an      */
/* operator splitting solver for quadratic programs", Alg. 1 + Sec. 4 polish,  */
/* Sec. 5.2 adaptive rho); no Ruiz equilibration (it changes the iterates, not */
/* the optimum).  The linear system is solved in the reduced form               */
/*   (H + sigma I + C' R C) x~ = sigma x - g + C'(R z - y),  z~ = C x~          */
/* which is algebraically identical to OSQP's quasi-definite KKT solve.         */
/* call sites: QPInverseKinematics.cpp:61-146 (init/update), :512-528 (solve)  */
/* ------------------------------------------------------------------------- */
void orc_default_settings(orc_qp_settings* s)
{
    s->rho = 0.1, s->sigma = 1e-6, s->alpha = 1.6;
    s->eps_abs = 1e-3, s->eps_rel = 1e-3;
    s->max_iter = 4000, s->check_termination = 25;
    s->adaptive_rho_interval = 50; /* OSQP's is time based; fixed here for determinism */
    s->polish = 0, s->delta = 1e-6, s->polish_refine_iter = 3;
    s->warm_start = 1;
}
void orc_parity_settings(orc_qp_settings* s)
{
    orc_default_settings(s);
    s->eps_abs = 1e-7, s->eps_rel = 1e-7;
    s->polish = 12; /* 1 = OSQP's single polish; > 1 = that many tries with active-set refinement in between */
    s->polish_refine_iter = 10; /* OSQP: 3; the regularised KKT solve needs more on ill-conditioned instances */
}

/* in-place lower Cholesky of an n x n row-major SPD matrix; 0 on success */
static int chol_factor(double* K, int n)
{
    for (int j = 0; j < n; j++) {
        double d = K[j * n + j];
        for (int k = 0; k < j; k++) d -= K[j * n + k] * K[j * n + k];
        if (!(d > 0)) return -1;
        d = sqrt(d);
        K[j * n + j] = d;
        for (int i = j + 1; i < n; i++) {
            double v = K[i * n + j];
            for (int k = 0; k < j; k++) v -= K[i * n + k] * K[j * n + k];
            K[i * n + j] = v / d;
        }
    }
    return 0;
}
static void chol_solve(const double* L, int n, double* x)
{
    for (int i = 0; i < n; i++) {
        double v = x[i];
        for (int k = 0; k < i; k++) v -= L[i * n + k] * x[k];
        x[i] = v / L[i * n + i];
    }
    for (int i = n - 1; i >= 0; i--) {
        double v = x[i];
        for (int k = i + 1; k < n; k++) v -= L[k * n + i] * x[k];
        x[i] = v / L[i * n + i];
    }
}
/* LU with partial pivoting, in place; piv[n]; 0 on success */
static int lu_factor(double* A, int n, int* piv)
{
    for (int k = 0; k < n; k++) {
        int p = k;
        double best = fabs(A[k * n + k]);
        for (int i = k + 1; i < n; i++)
            if (fabs(A[i * n + k]) > best) best = fabs(A[i * n + k]), p = i;
        if (best == 0.0) return -1;
        piv[k] = p;
        if (p != k)
            for (int j = 0; j < n; j++) {
                double t = A[k * n + j];
                A[k * n + j] = A[p * n + j], A[p * n + j] = t;
            }
        for (int i = k + 1; i < n; i++) {
            double f = A[i * n + k] / A[k * n + k];
            A[i * n + k] = f;
            if (f != 0.0)
                for (int j = k + 1; j < n; j++) A[i * n + j] -= f * A[k * n + j];
        }
    }
    return 0;
}
static void lu_solve(const double* A, int n, const int* piv, double* b)
{
    /* whole rows (multipliers included) were swapped while factoring, so L is stored in
     * its final row order: permute b first, then substitute */
    for (int k = 0; k < n; k++)
        if (piv[k] != k) {
            double t = b[k];
            b[k] = b[piv[k]], b[piv[k]] = t;
        }
    for (int k = 0; k < n; k++)
        for (int i = k + 1; i < n; i++) b[i] -= A[i * n + k] * b[k];
    for (int i = n - 1; i >= 0; i--) {
        double v = b[i];
        for (int j = i + 1; j < n; j++) v -= A[i * n + j] * b[j];
        b[i] = v / A[i * n + i];
    }
}

typedef struct {
    int nv, m;
    const double *H, *g, *C, *l, *u;
} qp_data;

static void qp_residuals(const qp_data* d, const double* x, const double* y, const double* z,
                         double* Cx, double* r_prim, double* r_dual, double* prim_scale,
                         double* dual_scale)
{
    const int nv = d->nv, m = d->m;
    double rp = 0, nCx = 0, nz = 0;
    for (int i = 0; i < m; i++) {
        double v = 0;
        for (int j = 0; j < nv; j++) v += d->C[i * nv + j] * x[j];
        Cx[i] = v;
        if (fabs(v - z[i]) > rp) rp = fabs(v - z[i]);
        if (fabs(v) > nCx) nCx = fabs(v);
        if (fabs(z[i]) > nz) nz = fabs(z[i]);
    }
    double rd = 0, nHx = 0, nCty = 0, ng = 0;
    for (int j = 0; j < nv; j++) {
        double hx = 0, cty = 0;
        for (int k = 0; k < nv; k++) hx += d->H[j * nv + k] * x[k];
        for (int i = 0; i < m; i++) cty += d->C[i * nv + j] * y[i];
        double r = hx + d->g[j] + cty;
        if (fabs(r) > rd || r != r) rd = fabs(r);
        if (fabs(hx) > nHx) nHx = fabs(hx);
        if (fabs(cty) > nCty) nCty = fabs(cty);
        if (fabs(d->g[j]) > ng) ng = fabs(d->g[j]);
    }
    *r_prim = rp, *r_dual = rd;
    *prim_scale = nCx > nz ? nCx : nz;
    *dual_scale = nHx > nCty ? (nHx > ng ? nHx : ng) : (nCty > ng ? nCty : ng);
}

static int build_and_factor(const qp_data* d, const double* rho_vec, double sigma, double* K)
{
    const int nv = d->nv, m = d->m;
    memcpy(K, d->H, sizeof(double) * nv * nv);
    for (int j = 0; j < nv; j++) K[j * nv + j] += sigma;
    for (int i = 0; i < m; i++) {
        const double* Ci = d->C + i * nv;
        for (int a = 0; a < nv; a++) {
            if (Ci[a] == 0.0) continue;
            const double ra = rho_vec[i] * Ci[a];
            for (int b = 0; b <= a; b++) K[a * nv + b] += ra * Ci[b];
        }
    }
    for (int a = 0; a < nv; a++)
        for (int b = a + 1; b < nv; b++) K[a * nv + b] = K[b * nv + a];
    return chol_factor(K, nv);
}

static void set_rho_vec(const qp_data* d, double rho, double* rho_vec)
{
    for (int i = 0; i < d->m; i++) {
        if (d->l[i] <= -ORC_INFTY && d->u[i] >= ORC_INFTY)
            rho_vec[i] = 1e-6; /* OSQP RHO_MIN on unconstrained rows */
        else if (d->l[i] == d->u[i])
            rho_vec[i] = 1e3 * rho; /* OSQP RHO_EQ_OVER_RHO_INEQ */
        else
            rho_vec[i] = rho;
    }
}

/* OSQP polish (polish.c): active set from (z, y), regularised KKT solve with
 * iterative refinement against the unregularised system.  Returns 0 when the
 * polished point improves both residuals (and overwrites x, y, z).            */
/* `aflag` (may be NULL): on entry, if aflag[0] != 2, the active set to use (-1 lower, +1 upper, 0 free per row; equality
 * rows are always active) instead of OSQP's rule; on a rejected candidate it receives the refined guess (violated rows
 * added, rows with a wrong-sign multiplier dropped) so that the caller can try again -- a primal-dual active-set step. */
static int qp_polish(const qp_data* d, const orc_qp_settings* st, double* x, double* y, double* z,
                     double r_prim_admm, double r_dual_admm, double* r_prim_out,
                     double* r_dual_out, int* aflag)
{
    const int nv = d->nv, m = d->m;
    int* act = (int*)malloc(sizeof(int) * (m + 1));
    double* rhs_b = (double*)malloc(sizeof(double) * (m + 1));
    int* sgn = (int*)malloc(sizeof(int) * (m + 1));
    int ma = 0;
    const int given = aflag && m > 0 && aflag[0] != 2;
    for (int i = 0; i < m; i++) {
        if (d->l[i] == d->u[i]) {
            act[ma] = i, rhs_b[ma] = d->l[i], sgn[ma] = 0, ma++;
        } else if (given ? aflag[i] < 0 : (z[i] - d->l[i] < -y[i])) {
            act[ma] = i, rhs_b[ma] = d->l[i], sgn[ma] = -1, ma++;
        } else if (given ? aflag[i] > 0 : (d->u[i] - z[i] < y[i])) {
            act[ma] = i, rhs_b[ma] = d->u[i], sgn[ma] = 1, ma++;
        }
    }
    const int N = nv + ma;
    double* KKT = (double*)calloc((size_t)N * N, sizeof(double));  /* regularised, factored */
    double* KKT0 = (double*)calloc((size_t)N * N, sizeof(double)); /* unregularised */
    double* rhs = (double*)malloc(sizeof(double) * N);
    double* sol = (double*)calloc(N, sizeof(double));
    double* res = (double*)malloc(sizeof(double) * N);
    int* piv = (int*)malloc(sizeof(int) * N);
    for (int a = 0; a < nv; a++)
        for (int b = 0; b < nv; b++) KKT0[a * N + b] = d->H[a * nv + b];
    for (int r = 0; r < ma; r++)
        for (int b = 0; b < nv; b++) {
            KKT0[(nv + r) * N + b] = d->C[act[r] * nv + b];
            KKT0[b * N + nv + r] = d->C[act[r] * nv + b];
        }
    memcpy(KKT, KKT0, sizeof(double) * N * N);
    for (int a = 0; a < nv; a++) KKT[a * N + a] += st->delta;
    for (int r = 0; r < ma; r++) KKT[(nv + r) * N + nv + r] -= st->delta;
    for (int a = 0; a < nv; a++) rhs[a] = -d->g[a];
    for (int r = 0; r < ma; r++) rhs[nv + r] = rhs_b[r];
    int ok = lu_factor(KKT, N, piv) == 0;
    if (ok) {
        memcpy(sol, rhs, sizeof(double) * N);
        lu_solve(KKT, N, piv, sol);
        for (int it = 0; it < st->polish_refine_iter; it++) {
            for (int a = 0; a < N; a++) {
                double v = rhs[a];
                for (int b = 0; b < N; b++) v -= KKT0[a * N + b] * sol[b];
                res[a] = v;
            }
            lu_solve(KKT, N, piv, res);
            for (int a = 0; a < N; a++) sol[a] += res[a];
        }
        double* xp = sol;
        double* yp = (double*)calloc(m + 1, sizeof(double));
        double* zp = (double*)malloc(sizeof(double) * (m + 1));
        double* Cx = (double*)malloc(sizeof(double) * (m + 1));
        for (int r = 0; r < ma; r++) yp[act[r]] = sol[nv + r];
        for (int i = 0; i < m; i++) {
            double v = 0;
            for (int j = 0; j < nv; j++) v += d->C[i * nv + j] * xp[j];
            zp[i] = v < d->l[i] ? d->l[i] : (v > d->u[i] ? d->u[i] : v);
        }
        double rp, rd, ps, ds;
        qp_residuals(d, xp, yp, zp, Cx, &rp, &rd, &ps, &ds);
        /* dual sign check on the inequality rows (a wrong active set shows up here) */
        int sign_ok = 1;
        for (int r = 0; r < ma; r++) {
            if (sgn[r] < 0 && sol[nv + r] > 1e-9) sign_ok = 0;
            if (sgn[r] > 0 && sol[nv + r] < -1e-9) sign_ok = 0;
        }
        ok = sign_ok && (rp == rp) && (rd == rd) && rp <= fmax(r_prim_admm, 1e-10)
             && rd <= fmax(r_dual_admm, 1e-10);
        if (!ok && aflag && rp == rp) { /* refined active set for the next try */
            for (int i = 0; i < m; i++) aflag[i] = 0;
            for (int r = 0; r < ma; r++)
                if (sgn[r] != 0 && !((sgn[r] < 0 && sol[nv + r] > 1e-9) || (sgn[r] > 0 && sol[nv + r] < -1e-9))) aflag[act[r]] = sgn[r];
            for (int i = 0; i < m; i++) {
                if (d->l[i] == d->u[i]) continue;
                double v = 0;
                for (int j = 0; j < nv; j++) v += d->C[i * nv + j] * xp[j];
                if (v < d->l[i] - 1e-9 * (1 + fabs(d->l[i]))) aflag[i] = -1;
                if (v > d->u[i] + 1e-9 * (1 + fabs(d->u[i]))) aflag[i] = 1;
            }
        }
        if (ok) {
            memcpy(x, xp, sizeof(double) * nv);
            memcpy(y, yp, sizeof(double) * m);
            memcpy(z, zp, sizeof(double) * m);
            *r_prim_out = rp, *r_dual_out = rd;
        }
        free(yp), free(zp), free(Cx);
    }
    free(act), free(rhs_b), free(sgn), free(KKT), free(KKT0), free(rhs), free(sol), free(res),
        free(piv);
    return ok ? 0 : -1;
}

int orc_qp_solve(int nv, int m, const double* H, const double* g, const double* C,
                 const double* l, const double* u, const orc_qp_settings* st, double* x,
                 double* y, double* z, orc_qp_info* info)
{
    qp_data d = {nv, m, H, g, C, l, u};
    const int mm = m > 0 ? m : 1;
    double* K = (double*)malloc(sizeof(double) * nv * nv);
    double* buf = (double*)malloc(sizeof(double) * (3 * nv + 5 * mm));
    double *xt = buf, *rhs = buf + nv, *xprev = buf + 2 * nv;
    double *rho_vec = buf + 3 * nv, *zt = rho_vec + mm, *Cx = zt + mm, *zprev = Cx + mm;
    double rho = st->rho;
    orc_qp_info inf = {0, 1, 0, 0, 0, 0};
    (void)xprev, (void)zprev;

    if (!st->warm_start) {
        memset(x, 0, sizeof(double) * nv);
        memset(y, 0, sizeof(double) * m);
        memset(z, 0, sizeof(double) * m);
    }
    set_rho_vec(&d, rho, rho_vec);
    if (build_and_factor(&d, rho_vec, st->sigma, K) != 0) {
        inf.status = 2;
        goto done;
    }
    double rp = 0, rd = 0, ps = 0, ds = 0;
    int it;
    for (it = 1; it <= st->max_iter; it++) {
        /* x~ */
        for (int j = 0; j < nv; j++) rhs[j] = st->sigma * x[j] - g[j];
        for (int i = 0; i < m; i++) {
            const double w = rho_vec[i] * z[i] - y[i];
            const double* Ci = C + i * nv;
            for (int j = 0; j < nv; j++) rhs[j] += Ci[j] * w;
        }
        memcpy(xt, rhs, sizeof(double) * nv);
        chol_solve(K, nv, xt);
        /* z~ = C x~ */
        for (int i = 0; i < m; i++) {
            double v = 0;
            const double* Ci = C + i * nv;
            for (int j = 0; j < nv; j++) v += Ci[j] * xt[j];
            zt[i] = v;
        }
        for (int j = 0; j < nv; j++) x[j] = st->alpha * xt[j] + (1 - st->alpha) * x[j];
        for (int i = 0; i < m; i++) {
            const double zr = st->alpha * zt[i] + (1 - st->alpha) * z[i];
            double zn = zr + y[i] / rho_vec[i];
            zn = zn < l[i] ? l[i] : (zn > u[i] ? u[i] : zn);
            y[i] += rho_vec[i] * (zr - zn);
            z[i] = zn;
        }
        const int check = st->check_termination > 0 && (it % st->check_termination == 0);
        const int adapt = st->adaptive_rho_interval > 0 && (it % st->adaptive_rho_interval == 0);
        if (check || adapt || it == st->max_iter) {
            qp_residuals(&d, x, y, z, Cx, &rp, &rd, &ps, &ds);
            if (rp != rp || rd != rd) {
                inf.status = 2;
                break;
            }
            if (rp <= st->eps_abs + st->eps_rel * ps && rd <= st->eps_abs + st->eps_rel * ds) {
                inf.status = 0;
                break;
            }
            if (adapt) { /* OSQP compute_rho_estimate / adapt_rho */
                const double pn = rp / (ps + 1e-10), dn = rd / (ds + 1e-10);
                double rn = rho * sqrt(pn / (dn + 1e-10));
                rn = rn < 1e-6 ? 1e-6 : (rn > 1e6 ? 1e6 : rn);
                if (rn > 5 * rho || rn < rho / 5) {
                    rho = rn;
                    set_rho_vec(&d, rho, rho_vec);
                    if (build_and_factor(&d, rho_vec, st->sigma, K) != 0) {
                        inf.status = 2;
                        break;
                    }
                    inf.rho_updates++;
                }
            }
        }
    }
    inf.iters = it > st->max_iter ? st->max_iter : it;
    if (inf.status != 2) {
        qp_residuals(&d, x, y, z, Cx, &rp, &rd, &ps, &ds);
        inf.r_prim = rp, inf.r_dual = rd;
        if (st->polish) {
            /* OSQP polishes once; the parity settings add primal-dual active-set steps (st->polish > 1 tries in total) so
             * that a rejected first guess does not leave the answer at ADMM accuracy -- on an ill-conditioned instance
             * eps 1e-7 residuals can mean 1e-4 in x, and the oracle has to stand for "the optimum to KKT <= 1e-9" */
            double rpp, rdp;
            int* aflag = st->polish > 1 && m > 0 ? (int*)malloc(sizeof(int) * m) : NULL;
            if (aflag) aflag[0] = 2;
            for (int attempt = 0; attempt < st->polish; attempt++) {
                if (qp_polish(&d, st, x, y, z, rp, rd, &rpp, &rdp, aflag) == 0) {
                    inf.polished = 1;
                    inf.r_prim = rpp, inf.r_dual = rdp;
                    break;
                }
                if (!aflag || aflag[0] == 2) break;
            }
            free(aflag);
        }
    }
done:
    free(K), free(buf);
    if (info) *info = inf;
    return inf.status;
}

double orc_qp_kkt(int nv, int m, const double* H, const double* g, const double* C,
                  const double* l, const double* u, const double* x, const double* y)
{
    double worst = 0;
    for (int j = 0; j < nv; j++) {
        double r = g[j];
        for (int k = 0; k < nv; k++) r += H[j * nv + k] * x[k];
        for (int i = 0; i < m; i++) r += C[i * nv + j] * y[i];
        if (fabs(r) > worst || r != r) worst = fabs(r);
    }
    for (int i = 0; i < m; i++) {
        double v = 0;
        for (int j = 0; j < nv; j++) v += C[i * nv + j] * x[j];
        double viol = 0;
        if (v < l[i]) viol = l[i] - v;
        if (v > u[i]) viol = v - u[i];
        if (viol > worst) worst = viol;
        /* complementarity: y>0 needs Cx=u, y<0 needs Cx=l */
        double comp = 0;
        if (y[i] > 0) comp = fmin(y[i], fabs(u[i] - v));
        if (y[i] < 0) comp = fmin(-y[i], fabs(v - l[i]));
        if (comp > worst) worst = comp;
    }
    return worst;
}

/* ------------------------------------------------------------------------- */
/* K4: ForwardEuler on R3 x SO(3) x Rn, ForwardEuler.h:72-84 (x <- dx*dt + x,  */
/* manif tangent + group = left-plus), FloatingBaseSystemVelocityKinematics.cpp:33-67 */
/* ------------------------------------------------------------------------- */
void orc_integrate(int n, double p[3], double quat[4], double* s, const double* nu, double dt)
{
    double phi[3] = {nu[3] * dt, nu[4] * dt, nu[5] * dt}, dq[4], qn[4];
    for (int c = 0; c < 3; c++) p[c] += nu[c] * dt;
    orc_so3_exp(phi, dq);
    orc_quat_mul(dq, quat, qn);
    double nrm = sqrt(qn[0] * qn[0] + qn[1] * qn[1] + qn[2] * qn[2] + qn[3] * qn[3]);
    for (int c = 0; c < 4; c++) quat[c] = qn[c] / nrm;
    for (int j = 0; j < n; j++) s[j] += nu[6 + j] * dt;
}

/* ------------------------------------------------------------------------- */
/* whole tick                                                                 */
/* ------------------------------------------------------------------------- */
int orc_tick_batch(const orc_model* m, const orc_task* tasks, int ntasks,
                   const double* const* setpoints, const orc_qp_settings* st, long i0, long i1,
                   double* base_pos, double* base_quat, double* s, double* nu, double* wx,
                   double* wy, double* wz, int* status, int* iters, double dt_integrate,
                   int nticks)
{
    const int n = m->n, nv = 6 + n;
    const int mc = orc_constraint_rows(tasks, ntasks, n);
    const int mm = mc > 0 ? mc : 1;
    double* H = (double*)malloc(sizeof(double) * (nv * nv + nv + (size_t)mm * nv + 4 * mm + nv));
    double *g = H + nv * nv, *C = g + nv, *l = C + (size_t)mm * nv, *u = l + mm;
    double *yloc = u + mm, *zloc = yloc + mm, *xloc = zloc + mm;
    const double** sp = (const double**)malloc(sizeof(double*) * (ntasks > 0 ? ntasks : 1));
    int bad = 0;
    for (long i = i0; i < i1; i++) {
        for (int tick = 0; tick < nticks; tick++) {
            for (int t = 0; t < ntasks; t++) {
                const int len = orc_task_setpoint_len(&tasks[t], n);
                sp[t] = (setpoints && setpoints[t]) ? setpoints[t] + (size_t)i * len : NULL;
            }
            int rows = orc_assemble(m, tasks, ntasks, sp, base_pos + 3 * i, base_quat + 4 * i,
                                    s + (size_t)n * i, H, g, C, l, u);
            orc_qp_settings local = *st;
            double *x = xloc, *y = yloc, *z = zloc;
            if (wx && wy && wz) {
                x = wx + (size_t)nv * i, y = wy + (size_t)mm * i, z = wz + (size_t)mm * i;
            } else {
                local.warm_start = 0;
            }
            orc_qp_info info;
            int stt = rows < 0 ? 2 : orc_qp_solve(nv, rows, H, g, C, l, u, &local, x, y, z, &info);
            if (rows < 0) info.iters = 0;
            memcpy(nu + (size_t)nv * i, x, sizeof(double) * nv);
            if (status) status[i] = stt;
            if (iters) iters[i] = info.iters;
            if (stt != 0 && tick == nticks - 1) bad++;
            if (dt_integrate > 0)
                orc_integrate(n, base_pos + 3 * i, base_quat + 4 * i, s + (size_t)n * i,
                              nu + (size_t)nv * i, dt_integrate);
        }
    }
    free(H), free(sp);
    return bad;
}

typedef struct {
    const orc_model* m;
    const orc_task* tasks;
    int ntasks;
    const double* const* setpoints;
    const orc_qp_settings* st;
    long i0, i1;
    double *base_pos, *base_quat, *s, *nu, *wx, *wy, *wz;
    int *status, *iters;
    double dt;
    int nticks, bad;
} mt_job;

static void* mt_worker(void* arg)
{
    mt_job* j = (mt_job*)arg;
    j->bad = orc_tick_batch(j->m, j->tasks, j->ntasks, j->setpoints, j->st, j->i0, j->i1,
                            j->base_pos, j->base_quat, j->s, j->nu, j->wx, j->wy, j->wz,
                            j->status, j->iters, j->dt, j->nticks);
    return NULL;
}

/* one contiguous slice of instances per POSIX thread ("one IK instance per core") */
int orc_tick_batch_mt(const orc_model* m, const orc_task* tasks, int ntasks,
                      const double* const* setpoints, const orc_qp_settings* st, long N,
                      double* base_pos, double* base_quat, double* s, double* nu, double* wx,
                      double* wy, double* wz, int* status, int* iters, double dt_integrate,
                      int nticks, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
    mt_job* jobs = (mt_job*)malloc(sizeof(mt_job) * nthreads);
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
    int bad = 0;
    for (int t = 0; t < nthreads; t++) {
        mt_job j = {m, tasks, ntasks, setpoints, st, N * t / nthreads, N * (t + 1) / nthreads,
                    base_pos, base_quat, s, nu, wx, wy, wz, status, iters, dt_integrate, nticks, 0};
        jobs[t] = j;
        if (pthread_create(&th[t], NULL, mt_worker, &jobs[t]) != 0) {
            mt_worker(&jobs[t]);
            th[t] = 0;
        }
    }
    for (int t = 0; t < nthreads; t++) {
        if (th[t]) pthread_join(th[t], NULL);
        bad += jobs[t].bad;
    }
    free(jobs), free(th);
    return bad;
}
