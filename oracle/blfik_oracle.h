/*
 * blfik_oracle.h -- CPU oracle for the batched IntegrationBasedIK tick.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (the CUDA library, the
 * host classes, the Python bindings) may include, link or call this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs do.
 *
 * It is a plain-C FP64 restatement of the reference's single-robot control tick
 * (all citations relative to /root/reference):
 *   - task update()s:        src/IK/src/SE3Task.cpp:231-321, CoMTask.cpp:171-235,
 *                            SO3Task.cpp:153-192, JointTrackingTask.cpp:127-155,
 *                            JointLimitsTask.cpp:211-259, JointVelocityLimitsTask.cpp:134-145
 *   - H/g/constraint stack:  src/IK/src/QPInverseKinematics.cpp:442-491
 *   - solution unpack:       src/IK/src/QPInverseKinematics.cpp:531-541
 *   - integration:           src/ContinuousDynamicalSystem/include/BipedalLocomotion/
 *                            ContinuousDynamicalSystem/ForwardEuler.h:72-124,
 *                            FixedStepIntegrator.h:94-140,
 *                            src/ContinuousDynamicalSystem/src/FloatingBaseSystemVelocityKinematics.cpp:33-67
 * and of the arithmetic the reference delegates to libraries that are NOT in the
 * reference tree (restated from their published definitions):
 *   - iDynTree v12.3.3  KinDynComputations (MIXED representation FK / frame
 *     Jacobian / CoM Jacobian),
 *   - LieGroupControllers v0.2.0 proportional controllers on R3 and SO(3),
 *   - manif 0.0.5 SO(3) exp/log on unit quaternions, "tangent + group" = left-plus,
 *   - OSQP v0.6.3 (+ osqp-eigen v0.7.0): ADMM on the QP  min 1/2 x'Hx + g'x,
 *     l <= Cx <= u, per-row rho (x1e3 on equality rows), relaxation alpha,
 *     sigma regularisation, adaptive rho, and the polishing step.
 *
 * PARITY UNPINNED: the reference tree holds no golden numeric vector for this
 * path and none of those libraries can be built in this container, so the oracle
 * is pinned only by (a) the reference tests' structural identities and
 * convergence scenarios and (b) mathematically independent checks (finite
 * differences, KKT residuals, closed forms, scipy) -- see tests/.
 */
#ifndef BLFIK_ORACLE_H
#define BLFIK_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* one-sided bound value, OsqpEigen::INFTY (QPInverseKinematics.cpp:481-482) */
#define ORC_INFTY 1e30

enum {
    ORC_TASK_SE3 = 0,
    ORC_TASK_COM = 1,
    ORC_TASK_SO3 = 2,
    ORC_TASK_JOINT_TRACKING = 3,
    ORC_TASK_JOINT_LIMITS = 4,
    ORC_TASK_JOINT_VELOCITY_LIMITS = 5,
    ORC_TASK_R3 = 6,
    ORC_TASK_DISTANCE = 7,
    ORC_TASK_GRAVITY = 8,
    ORC_TASK_ANGULAR_MOMENTUM = 9
};

/* Kinematic tree, link 0 = floating base, parent[k] < k.  Link k (k>=1) is the
 * child of a revolute joint whose DoF index is dof[k]; its pose is
 *   W_H_k = W_H_parent * (Xr[k], Xp[k]) * Rot(axis[k], s[dof[k]]).            */
typedef struct {
    int n;                 /* number of DoFs; links = n+1                       */
    int nf;                /* number of frames                                  */
    const int* parent;     /* [n+1]                                             */
    const int* dof;        /* [n+1], dof[0] = -1                                */
    const double* Xr;      /* [n+1][9] row-major rotation parent <- joint frame */
    const double* Xp;      /* [n+1][3] joint origin in the parent link frame    */
    const double* axis;    /* [n+1][3] unit axis in the joint (= child) frame   */
    const double* mass;    /* [n+1]                                             */
    const double* com;     /* [n+1][3] link CoM in the link frame               */
    const int* frame_link; /* [nf]                                              */
    const double* Fr;      /* [nf][9] link <- frame rotation                    */
    const double* Fp;      /* [nf][3]                                           */
    const double* inertia; /* [n+1][6] Ixx Ixy Ixz Iyy Iyz Izz about the link CoM, link frame; NULL = point masses */
} orc_model;

typedef struct {
    int type;
    int priority;           /* 0 = hard constraint, 1 = weighted cost           */
    int frame;              /* SE3 / SO3 / R3                                   */
    int mask[3];            /* SE3 / CoM / R3 linear mask                       */
    int controller_enabled; /* SE3Task::setTaskControllerMode                   */
    double kp_lin[3];
    double kp_ang[3];
    const double* kp_joint; /* [n]  JointTrackingTask                           */
    const double* weight;   /* [rows] priority-1 tasks                          */
    const double* klim;     /* [n]  JointLimitsTask                             */
    const double* upper;    /* [n]  JointLimitsTask / JointVelocityLimitsTask   */
    const double* lower;    /* [n]                                              */
    double sampling_time;   /* JointLimitsTask                                  */
    int ref_frame;          /* DistanceTask: reference frame, -1 = world origin  */
    double threshold;       /* DistanceTask: distance_numeric_threshold          */
    int exo_pos;            /* SE3Task use_position_exogenous_feedback / CoMTask use_exogenous_feedback (SE3Task.cpp:196-213, CoMTask.cpp:116) */
    int exo_ori;            /* SE3Task use_orientation_exogenous_feedback         */
} orc_task;

typedef struct {
    double rho, sigma, alpha;
    double eps_abs, eps_rel;
    int max_iter;
    int check_termination;   /* every k iterations                              */
    int adaptive_rho_interval; /* 0 = off                                       */
    int polish;              /* 0 off, 1 = OSQP-style polishing, k > 1 = up to k tries with active-set refinement in between */
    double delta;            /* polish regularisation                           */
    int polish_refine_iter;
    int warm_start;
} orc_qp_settings;

typedef struct {
    int iters;
    int status;              /* 0 solved, 1 max-iter (inaccurate), 2 failure    */
    int polished;            /* 1 if the polished point was accepted            */
    int rho_updates;
    double r_prim, r_dual;   /* unscaled inf-norm residuals of the returned x,y */
} orc_qp_info;

void orc_default_settings(orc_qp_settings* s);        /* OSQP v0.6.3 defaults  */
void orc_parity_settings(orc_qp_settings* s);         /* eps 1e-7 + polish     */

/* set-point record lengths per task type (doubles) */
int orc_setpoint_len(int type, int n);
/* record length of one task: the set-point, followed by the exogenous feedback (SE3: p(3) q(4); CoM: p(3)) when the task uses one */
int orc_task_setpoint_len(const orc_task* t, int n);
int orc_task_rows(const orc_task* t, int n);
int orc_task_is_inequality(const orc_task* t);

/* ---- kinematics (iDynTree MIXED semantics) ---- */
void orc_quat_to_rot(const double q[4], double R[9]);
void orc_rot_to_quat(const double R[9], double q[4]);
void orc_so3_exp(const double phi[3], double q[4]);
void orc_so3_log(const double q[4], double phi[3]);
void orc_quat_mul(const double a[4], const double b[4], double out[4]);

void orc_fk(const orc_model* m, const double p[3], const double quat[4], const double* s,
            double* link_R, double* link_p);
void orc_frame_pose(const orc_model* m, const double* link_R, const double* link_p, int frame,
                    double R[9], double p[3]);
void orc_frame_jacobian(const orc_model* m, const double* link_R, const double* link_p, int frame,
                        double* J /* 6 x (6+n) row-major */);
void orc_com(const orc_model* m, const double* link_R, const double* link_p, double com[3],
             double* Jcom /* 3 x (6+n) row-major, may be NULL */);

/* angular rows (3 x nv) of the centroidal momentum matrix, MIXED representation: L = sum_k I_k^W w_k + m_k (c_k - c) x cdot_k */
void orc_angular_momentum_matrix(const orc_model* m, const double* link_R, const double* link_p, double* A);

/* ---- tasks + assembly ---- */
/* A is rows x nv row-major, b is rows.  Returns rows, or -1 on error.          */
int orc_task_update(const orc_model* m, const orc_task* t, const double* link_R,
                    const double* link_p, const double* s, const double* setpoint, double* A,
                    double* b);

/* Follows QPInverseKinematics::advance: H = sum A'WA, g = -sum A'Wb over
 * priority-1 tasks; C/l/u stacked over priority-0 tasks in task order.
 * Returns the number of constraint rows m (C is m x nv row-major).             */
int orc_assemble(const orc_model* m, const orc_task* tasks, int ntasks,
                 const double* const* setpoints, const double p[3], const double quat[4],
                 const double* s, double* H, double* g, double* C, double* l, double* u);

/* ---- QP ---- */
/* x[nv], y[m], z[m] are warm-start in / solution out.                          */
int orc_qp_solve(int nv, int m, const double* H, const double* g, const double* C,
                 const double* l, const double* u, const orc_qp_settings* st, double* x,
                 double* y, double* z, orc_qp_info* info);
/* max KKT violation (stationarity, primal feasibility, dual sign/complementarity) */
double orc_qp_kkt(int nv, int m, const double* H, const double* g, const double* C,
                  const double* l, const double* u, const double* x, const double* y);

/* ---- integration: p += v dt; R <- exp(w dt) R; s += sdot dt ---- */
void orc_integrate(int n, double p[3], double quat[4], double* s, const double* nu, double dt);

/* ---- whole tick over instances [i0, i1) of robot-major batches ----
 * base_pos [N][3], base_quat [N][4] (x,y,z,w), s [N][n], setpoints[t] [N][len_t],
 * nu out [N][6+n], warm-start x/y/z [N][nv]/[N][m]/[N][m] (may be NULL -> cold),
 * status out [N] (may be NULL), iters out [N] (may be NULL).
 * If dt_integrate > 0 the state is advanced in place after each solve.
 * Returns 0, or the number of instances whose status != 0.                     */
int orc_tick_batch(const orc_model* m, const orc_task* tasks, int ntasks,
                   const double* const* setpoints, const orc_qp_settings* st, long i0, long i1,
                   double* base_pos, double* base_quat, double* s, double* nu, double* wx,
                   double* wy, double* wz, int* status, int* iters, double dt_integrate,
                   int nticks);
/* same, instances split over POSIX threads (bench cpu_baseline leg)           */
int orc_tick_batch_mt(const orc_model* m, const orc_task* tasks, int ntasks,
                      const double* const* setpoints, const orc_qp_settings* st, long N,
                      double* base_pos, double* base_quat, double* s, double* nu, double* wx,
                      double* wy, double* wz, int* status, int* iters, double dt_integrate,
                      int nticks, int nthreads);
int orc_constraint_rows(const orc_task* tasks, int ntasks, int n);

#ifdef __cplusplus
}
#endif
#endif
