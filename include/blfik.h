/*
 * blfik.h -- C ABI of libblfik_cuda.so: the batched IntegrationBasedIK control tick
 * of bipedal-locomotion-framework on NVIDIA B200 (sm_100a), FP64.
 *
 * One call of blfik_tick() runs, for N independent robots at once, what the
 * reference runs for one robot per control period (citations relative to the
 * reference tree):
 *
 *   kinDyn->setRobotState(...)            caller, e.g. src/IK/tests/QPInverseKinematicsTest.cpp:466-472
 *   task->update() for every task         src/IK/src/QPInverseKinematics.cpp:429-439
 *   H, g, constraint stack                src/IK/src/QPInverseKinematics.cpp:442-491
 *   OsqpEigen solve                       src/IK/src/QPInverseKinematics.cpp:494-528
 *   unpack [base twist; joint velocity]   src/IK/src/QPInverseKinematics.cpp:531-541
 *   ForwardEuler::oneStepIntegration      src/ContinuousDynamicalSystem/include/BipedalLocomotion/
 *                                         ContinuousDynamicalSystem/ForwardEuler.h:96-124
 *
 * The boundary is plain C: opaque handles, plain pointers and sizes.  All batch
 * arrays are robot-major ("array of robots"): element (i, c) of an array with C
 * components per robot lives at data[i*C + c].  Quaternions are stored (x, y, z, w)
 * as Eigen/manif do; twists are [linear; angular] in the MIXED representation
 * (base linear and angular velocity expressed in the inertial frame), as
 * QPInverseKinematics.cpp:531-541 and SE3Task.cpp:92-102 require.
 *
 * Error convention (the reference returns bool and logs, e.g.
 * QPInverseKinematics.cpp:416-426): every function returns BLFIK_OK (0) or a
 * negative blfik_error; blfik_last_error() gives the message.  There is no CPU
 * fallback: without a CUDA device every compute entry point fails.
 *
 * Threading: a blfik_problem is single-writer; work is enqueued on the problem's
 * CUDA stream and is asynchronous until blfik_sync() or a host-memory getter.
 */
#ifndef BLFIK_H
#define BLFIK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BLFIK_VERSION 2
#define BLFIK_INFTY 1e30 /* OsqpEigen::INFTY, QPInverseKinematics.cpp:481-482 */

typedef enum {
    BLFIK_OK = 0,
    BLFIK_ERR_INVALID_ARGUMENT = -1,
    BLFIK_ERR_CUDA = -2,
    BLFIK_ERR_UNSUPPORTED = -3,
    BLFIK_ERR_NOT_READY = -4,  /* e.g. a set-point was never set (SE3Task.cpp:245-249) */
    BLFIK_ERR_SOLVER = -5      /* at least one instance did not reach status 0 (QPInverseKinematics.cpp:518-523) */
} blfik_error;

/* task kinds; the names are the reference's factory strings (BLF_REGISTER_IK_TASK) */
typedef enum {
    BLFIK_TASK_SE3 = 0,                   /* src/IK/src/SE3Task.cpp                 */
    BLFIK_TASK_COM = 1,                   /* src/IK/src/CoMTask.cpp                 */
    BLFIK_TASK_SO3 = 2,                   /* src/IK/src/SO3Task.cpp                 */
    BLFIK_TASK_JOINT_TRACKING = 3,        /* src/IK/src/JointTrackingTask.cpp       */
    BLFIK_TASK_JOINT_LIMITS = 4,          /* src/IK/src/JointLimitsTask.cpp         */
    BLFIK_TASK_JOINT_VELOCITY_LIMITS = 5, /* src/IK/src/JointVelocityLimitsTask.cpp */
    BLFIK_TASK_R3 = 6,                    /* src/IK/src/R3Task.cpp                  */
    BLFIK_TASK_DISTANCE = 7,              /* src/IK/src/DistanceTask.cpp            */
    BLFIK_TASK_GRAVITY = 8,               /* src/IK/src/GravityTask.cpp             */
    BLFIK_TASK_ANGULAR_MOMENTUM = 9       /* src/IK/src/AngularMomentumTask.cpp     */
} blfik_task_type;

/* where a caller-provided batch buffer lives */
typedef enum { BLFIK_MEM_HOST = 0, BLFIK_MEM_DEVICE = 1 } blfik_mem;

/* per-instance solver status after a tick */
typedef enum {
    BLFIK_STATUS_SOLVED = 0,     /* KKT point certified (polish) or ADMM residuals below eps */
    BLFIK_STATUS_MAX_ITER = 1,   /* OSQP "solved inaccurate": iteration budget exhausted.  There is no certificate of
                                  * infeasibility on the device: an instance OSQP would prove primal infeasible (the reference's
                                  * advance() fails there) also ends here, not in FAILED -- a caller that wants the reference's
                                  * strictness treats this status as a failure                         */
    BLFIK_STATUS_FAILED = 2      /* infeasible bounds, factorisation breakdown or NaN         */
} blfik_status;

/* flags of blfik_tick */
#define BLFIK_TICK_INTEGRATE 1u   /* apply ForwardEuler after each solve (closed loop)           */
#define BLFIK_TICK_NO_FEEDFORWARD 4u /* blfik_tick_packed only: compact records, see there                 */
#define BLFIK_TICK_KEEP_STATE 8u  /* blfik_tick_packed only: the state part of the records is ignored, the tick starts from
                                     the state the device holds (a closed loop integrated on the device whose host copy
                                     the caller does not refresh every period)                                     */
#define BLFIK_TICK_WARM_START 2u  /* start ADMM from the previous tick's (x, z, y) (reference does,
                                     QPInverseKinematics.cpp:494-509)                             */
#define BLFIK_TICK_STRICT 16u     /* a tick that ends with MAX_ITER counts as FAILED: no velocity, the rollout of that robot ends
                                     (what the reference does with an instance OSQP proves infeasible or cannot solve inside its
                                     iteration limit, QPInverseKinematics.cpp:518-523; without the flag such a tick's best iterate
                                     is reported and integrated like OSQP's "solved inaccurate")                          */

/* Kinematic tree in the layout iDynTree's reduced models boil down to (stands in for
 * std::shared_ptr<iDynTree::KinDynComputations>, IKLinearTask.h:40).  Link 0 is the
 * floating base; parent[k] < k; link k>=1 is the child of the revolute joint with DoF
 * index dof[k]:  W_H_k = W_H_parent * (X_rot[k], X_pos[k]) * Rot(axis[k], s[dof[k]]). */
typedef struct {
    int32_t n_dofs;
    int32_t n_frames;
    const int32_t* parent;     /* [n_dofs+1]                                            */
    const int32_t* dof;        /* [n_dofs+1], dof[0] = -1                               */
    const double* X_rot;       /* [n_dofs+1][9] row-major                               */
    const double* X_pos;       /* [n_dofs+1][3]                                         */
    const double* axis;        /* [n_dofs+1][3] unit, in the joint (= child link) frame */
    const double* mass;        /* [n_dofs+1]                                            */
    const double* com;         /* [n_dofs+1][3] in the link frame                       */
    const int32_t* frame_link; /* [n_frames]                                            */
    const double* frame_rot;   /* [n_frames][9] link <- frame                           */
    const double* frame_pos;   /* [n_frames][3]                                         */
    const double* inertia;     /* [n_dofs+1][6] Ixx Ixy Ixz Iyy Iyz Izz about the link CoM, link frame; NULL = point masses.
                                  Only the AngularMomentumTask reads it (centroidal momentum matrix).             */
} blfik_model_desc;

/* One task; fields mirror the reference's parameter-handler keys. */
typedef struct {
    int32_t type;               /* blfik_task_type                                      */
    int32_t priority;           /* 0 hard constraint, 1 weighted cost (addTask, QPInverseKinematics.cpp:167-249) */
    int32_t frame;              /* frame index ("frame_name"), SE3/SO3/R3               */
    int32_t mask[3];            /* "mask" of the linear part, SE3/CoM/R3                */
    int32_t controller_enabled; /* setTaskControllerMode, SE3Task.cpp:251-258           */
    double kp_linear[3];        /* "kp_linear" (scalar gains are replicated)            */
    double kp_angular[3];       /* "kp_angular"                                         */
    const double* kp_joint;     /* "kp" [n]                 JointTrackingTask           */
    const double* weight;       /* "weight" [rows]          priority-1 tasks            */
    const double* klim;         /* "klim" [n]               JointLimitsTask             */
    const double* upper_limits; /* [n] JointLimitsTask (positions) / JointVelocityLimitsTask (velocities) */
    const double* lower_limits; /* [n]                                                  */
    double sampling_time;       /* "sampling_time" [s]      JointLimitsTask             */
    int32_t ref_frame;          /* "reference_frame_name" of DistanceTask, -1 = world origin ("")   */
    double distance_threshold;  /* "distance_numeric_threshold" of DistanceTask (0 -> 0.001)       */
    int32_t use_position_exogenous_feedback;    /* SE3Task key of the same name (SE3Task.cpp:206-214); CoMTask "use_exogenous_feedback" (CoMTask.cpp:116) */
    int32_t use_orientation_exogenous_feedback; /* SE3Task key of the same name (SE3Task.cpp:196-204)                                               */
} blfik_task_desc;

/* ADMM + polish parameters.  rho/sigma/alpha have OSQP's meaning (osqp v0.6.3); the
 * equality rows are projected exactly inside the x-update (the rho_eq -> inf limit
 * of OSQP's per-row rho), the box rows (joint limits) are the ADMM-split part.       */
typedef struct {
    double rho;            /* step on the box rows                    (default 1.0)    */
    double sigma;          /* proximal regularisation                 (default 1e-2)   */
    double alpha;          /* over-relaxation                         (default 1.6)    */
    double eps;            /* inf-norm residual tolerance of ADMM     (default 1e-7)   */
    int32_t admm_iters;    /* ADMM iterations before each polish try  (default 10)     */
    int32_t max_rounds;    /* ADMM+polish rounds before giving up     (default 8)      */
    int32_t polish_iters;  /* proximal refinement steps in the polish (default 4)      */
    int32_t eq_iters;      /* proximal steps when there are no box rows (default 5)    */
} blfik_solver_settings;

typedef struct blfik_model blfik_model;
typedef struct blfik_problem blfik_problem;

/* ---- life cycle ------------------------------------------------------------------ */
/* replaces kinDyn->loadRobotModel(...) + setFrameVelocityRepresentation(MIXED)        */
int blfik_model_create(const blfik_model_desc* desc, blfik_model** out);
void blfik_model_destroy(blfik_model* model);

/* replaces QPInverseKinematics::addTask x ntasks + finalize()
 * (QPInverseKinematics.cpp:167-249, 323-395) for a batch of N robots on one device.    */
int blfik_problem_create(const blfik_model* model, const blfik_task_desc* tasks, int32_t ntasks,
                         int64_t n_instances, int32_t device, blfik_problem** out);
void blfik_problem_destroy(blfik_problem* problem);

void blfik_default_settings(blfik_solver_settings* s);
int blfik_set_solver_settings(blfik_problem* problem, const blfik_solver_settings* s);

/* ---- per-tick inputs -------------------------------------------------------------- */
/* replaces kinDyn->setRobotState(world_H_base, s, ...) / dynamics->setState(...)
 * (FloatingBaseSystemVelocityKinematics.cpp:69-73): base_pos [N][3], base_quat [N][4], s [N][n] */
int blfik_set_state(blfik_problem* problem, const double* base_pos, const double* base_quat,
                    const double* s, int32_t mem);
/* replaces integrator->getSolution() (Integrator.h:117-122)                            */
int blfik_get_state(blfik_problem* problem, double* base_pos, double* base_quat, double* s,
                    int32_t mem);

/* replaces task->setSetPoint(...) (SE3Task.cpp:323-336, CoMTask.cpp:237-247, SO3Task.cpp:194-206,
 * JointTrackingTask.cpp:162-185, R3Task.cpp:245-256).  data is [N][len] with len:
 *   SE3: 13 = p_d(3) q_d(4) v_ff(3) w_ff(3)     CoM / R3: 6 = p_d(3) v_ff(3)
 *   SO3:  7 = q_d(4) w_ff(3)                    JointTracking: 2n = s_d(n) sdot_d(n)
 *   Distance: 1 = desired distance (DistanceTask::setDesiredDistance; gain kp_linear[0], frame = target_frame_name)
 *   Gravity:  6 = desired gravity direction in the target frame (3) + feed-forward (3)  (GravityTask::setSetPoint; kp_linear[0])
 *   AngularMomentum: 3 = desired centroidal angular momentum (AngularMomentumTask::setSetPoint; rows selected by mask) */
int blfik_set_setpoint(blfik_problem* problem, int32_t task, const double* data, int32_t mem);
int32_t blfik_setpoint_len(const blfik_problem* problem, int32_t task);

/* replaces task->setFeedback(...) (SE3Task.cpp:338-366, CoMTask.cpp:262-273) for a task created with
 * use_position_exogenous_feedback / use_orientation_exogenous_feedback: the state the task's P-controllers compare the
 * set-point with is what the caller measured instead of the pose kinDyn predicts (the Jacobian stays kinDyn's).
 * data is [N][7] = p(3) q(4) for an SE3Task (the part whose flag is off is ignored), [N][3] for a CoMTask.
 * Until the first call the feedback is the controllers' initial state: zero position, identity rotation.
 * In the batch record of such a task (blfik_tick_host, blfik_tick_packed: blfik_record_len doubles per robot) the
 * feedback follows the set-point.                                                                             */
int blfik_set_feedback(blfik_problem* problem, int32_t task, const double* data, int32_t mem);
int32_t blfik_feedback_len(const blfik_problem* problem, int32_t task);
/* blfik_setpoint_len + blfik_feedback_len: doubles per robot of the task's record */
int32_t blfik_record_len(const blfik_problem* problem, int32_t task);

/* On-device set-point producer for closed-loop rollouts (no single counterpart: it stands in for the planners /
 * splines whose output the reference's callers pass to setSetPoint every period, e.g.
 * utilities/balancing-position-control/script/blf-balancing-position-control.py:351-474, so that a multi-tick launch
 * needs no host traffic).  The set-point of `task` at time t becomes
 *     x_d(t) = x_0 + A sin(2 pi t / period + phase_i),   feed-forward += A (2 pi / period) cos(.),
 * x_0 being what blfik_set_setpoint stored; amplitude = [linear A (3), angular A (3, rotation vector about a fixed
 * world axis)], phase [N] per robot.  period <= 0 switches it off.  Time starts at 0 and advances by dt with every
 * integrating tick (blfik_set_time / blfik_get_time).                                                      */
int blfik_set_setpoint_trajectory(blfik_problem* problem, int32_t task, const double* amplitude,
                                  double period, const double* phase, int32_t mem);

/* Minimum-jerk point-to-point set-point producer on the device: what the reference's callers obtain from
 * Planners::SO3Planner<Right> for rotations (SO3Planner.tpp:96-105, 205-252 with zero boundary velocity and acceleration)
 * and from a two-knot QuinticSpline for positions, and then pass to setSetPoint every period.  From the stored set-point
 * x_0 (blfik_set_setpoint) to target [N][blfik_setpoint_len] over `duration` seconds starting at the problem's current
 * time, with s(tau) = 10 tau^3 - 15 tau^4 + 6 tau^5:
 *     p(t) = p_0 + s (p_1 - p_0),  R(t) = Exp(s Log(R_1 R_0^-1)) R_0,  feed-forward += sdot (p_1 - p_0) / sdot Log(R_1 R_0^-1);
 * before the start and after the end the time is clamped (SO3Planner::advance, :289-292).  Only the pose part of
 * `target` is read.  duration <= 0 switches it off; it replaces a sinusoidal generator on the same task and vice versa. */
int blfik_set_setpoint_minimum_jerk(blfik_problem* problem, int32_t task, const double* target,
                                    double duration, int32_t mem);

/* Multi-knot quintic spline as set-point producer on the device, for the linear part of an SE3 / CoM / R3 task: what the
 * reference's callers obtain from Math::QuinticSpline<Eigen::Vector3d> (src/Math/include/BipedalLocomotion/Math/Spline.h:
 * setKnots :323-383, setInitialConditions / setFinalConditions :385-436, evaluatePoint :594-644; QuinticSpline.h:266-415)
 * and pass to setSetPoint every period (e.g. Planners::SwingFootPlanner, the CoM trajectory of the balancing application).
 *   times      host [n_knots], strictly increasing, on the problem's clock (blfik_set_time / blfik_get_time), shared by
 *              the robots;
 *   positions  [N][n_knots][3], the knots of every robot;
 *   boundary   [N][4][3]: velocity and acceleration at the first knot, velocity and acceleration at the last one
 *              (NULL: all zero);
 * On a JointTrackingTask the spline lives in joint space (Math::QuinticSpline<Eigen::VectorXd>, whose output the reference's
 * callers hand to JointTrackingTask::setSetPoint(position, velocity), JointTrackingTask.cpp:162-185): positions
 * [N][n_knots][n_dofs], boundary [N][4][n_dofs]; the joint positions of the set-point are the spline, the joint velocities the
 * stored ones plus its derivative (coefficients: 48 n_dofs (n_knots - 1) bytes per robot of device memory);
 * `mem` tells where positions and boundary live.  The velocities and accelerations at the interior knots are those that
 * make the third and fourth derivative continuous (the reference's linear system, solved here on the device for all
 * robots at once).  At a tick with time t the position of the set-point is the spline, its feed-forward velocity the stored
 * one plus the spline's derivative; before the first knot and from the last knot on the end states hold, derivative
 * included, as Spline::evaluatePoint does.  The orientation of an SE3 task keeps what drives it: the stored set-point or
 * the minimum-jerk generator (swing foot: spline for the origin, SO3Planner for the orientation).  n_knots == 0 switches
 * the spline off; n_knots == 1 and equal consecutive times are refused as by the reference; it replaces a sinusoidal
 * generator on the same task and vice versa.                                                                     */
#define BLFIK_MAX_SPLINE_KNOTS 1024
int blfik_set_setpoint_spline(blfik_problem* problem, int32_t task, int32_t n_knots, const double* times,
                              const double* positions, const double* boundary, int32_t mem);
/* The same with Math::CubicSpline<Eigen::Vector3d> (CubicSpline.h:77-330, the default interpolation of
 * Planners::SwingFootPlanner): knot velocities from the continuity of the acceleration; boundary [N][2][3] = velocity at
 * the first and at the last knot (NULL: zero) -- the boundary accelerations a caller hands to the reference's class are
 * not used by it.  A task has one spline: this replaces a quintic one and vice versa; n_knots == 0 switches it off. */
int blfik_set_setpoint_cubic_spline(blfik_problem* problem, int32_t task, int32_t n_knots, const double* times,
                                    const double* positions, const double* boundary, int32_t mem);
/* The set-point of a task as a tick at `time` reads it: the stored set-point (blfik_set_setpoint) with the attached
 * generators applied, out [N][blfik_setpoint_len(task)].  What the reference's callers log next to the measured pose (the
 * output of their spline / planner, e.g. Spline::getOutput, Spline.h:248-254, before it goes to setSetPoint); no solver
 * state is touched.  Host-tick records (blfik_tick_packed) are not consulted: this reads what blfik_set_setpoint stored. */
int blfik_evaluate_setpoint(blfik_problem* problem, int32_t task, double time, double* out, int32_t mem);
int blfik_set_time(blfik_problem* problem, double t);
double blfik_get_time(const blfik_problem* problem);

/* replaces QPInverseKinematics::setTaskWeight (QPInverseKinematics.cpp:251-305); host pointer,
 * shared by the whole batch, takes effect on the next tick                             */
int blfik_set_weight(blfik_problem* problem, int32_t task, const double* weight);
/* replaces the kp_* keys / setTaskControllerMode after construction                     */
int blfik_set_gains(blfik_problem* problem, int32_t task, const double* kp_linear,
                    const double* kp_angular, const double* kp_joint);
int blfik_set_controller_mode(blfik_problem* problem, int32_t task, int32_t enabled);
/* replaces a MultiStateWeightProvider / FirstOrderSmoother on the task's weight port (MultiStateWeightProvider.cpp:22-134,
 * FirstOrderSmoother.cpp:60-115) for ticks that run without the host in between: from its current value the weight moves to
 * `target` [rows] along dx = -a x + a u, a = 3 / settling_time, one ForwardEuler step of dt at the start of every tick --
 * what provider->advance() followed by ik->advance() does in the reference's loop.  Inside a multi-tick launch the weight
 * therefore changes from tick to tick (two-launch kernel path; the other kernels take one tick per call with a schedule
 * in force).  settling_time <= 0 or target = NULL ends the schedule, the weight stays where it is; blfik_set_weight sets it
 * outright.  blfik_get_weight reads the current value.                                                              */
int blfik_set_weight_schedule(blfik_problem* problem, int32_t task, const double* target, double settling_time);
int blfik_get_weight(blfik_problem* problem, int32_t task, double* weight);
/* replaces the upper_limits / lower_limits / klim / sampling_time keys of JointLimitsTask (JointLimitsTask.cpp:136-198)
 * and JointVelocityLimitsTask (JointVelocityLimitsTask.cpp:111-124) after construction; NULL / 0 keeps a value        */
int blfik_set_limits(blfik_problem* problem, int32_t task, const double* upper_limits,
                     const double* lower_limits, const double* klim, double sampling_time);

/* ---- the tick ---------------------------------------------------------------------- */
/* replaces QPInverseKinematics::advance() (+ ForwardEuler::integrate(0, dt) when
 * BLFIK_TICK_INTEGRATE) repeated nticks times with the set-points held.  Asynchronous.
 * Status of a launch: the WORST tick of each robot (0 < 1 < 2), not the last one.  A tick that ends FAILED
 * (the reference's advance() returns false there, QPInverseKinematics.cpp:511-523, and its caller stops
 * integrating) ends that robot's rollout: its nu is zero, its state stays where the failure happened and the
 * remaining ticks of the launch leave it alone.  MAX_ITER ("solved inaccurate") is accepted like the reference
 * accepts it: the tick's velocity is reported and integrated.
 * Closed loops (two-launch kernel path, launches over the whole batch): like the warm start, two pieces of history
 * travel from tick to tick, inside a launch and across launches -- the robots that needed many iterations at the
 * previous tick are handed out first, and a robot that has been ending with MAX_ITER gets a quarter of the iteration
 * budget (the full one every sixteenth such tick; the first solved tick clears it).  Neither changes a robot that
 * solves every tick; blfik_set_state starts a new history.  A tick without history hands the robots out by how
 * narrow their joint boxes are, the narrowest first; the order changes no result.                             */
int blfik_tick(blfik_problem* problem, int32_t nticks, double dt, uint32_t flags);

/* The whole reference-facing tick for callers whose data live in HOST memory (what the reference's caller does
 * per period: setRobotState + setSetPoint + advance + getOutput, QPInverseKinematicsTest.cpp:459-484), pipelined:
 * the batch is cut into nchunks pieces and host->device copies, the tick and the device->host copies of
 * consecutive pieces overlap on three streams.  setpoints[t] is the host batch of task t, or NULL to keep what
 * the task already has (setpoints itself may be NULL); status may be NULL.  Returns when nu (and status) are
 * complete.  Copies only overlap if the host buffers are page-locked (cudaHostAlloc / cudaHostRegister /
 * torch pin_memory); with pageable memory the call is correct but serial.  One tick, cold or warm per flags.    */
int blfik_tick_host(blfik_problem* problem, const double* base_pos, const double* base_quat,
                    const double* s, const double* const* setpoints, double* nu, int32_t* status,
                    double dt, uint32_t flags, int32_t nchunks);

/* Same tick from ONE packed host record per robot, so that each pipeline chunk is a single host->device copy
 * (with separate arrays the per-copy latency of ~10 small copies per chunk eats the overlap).  A record is
 * blfik_packed_record_len() doubles: base position (3), base quaternion (4), joint positions (n), then the
 * set-point of every task that has one, in task order; blfik_packed_offset(task) gives each offset (task = -1:
 * the state, offset 0; tasks without a set-point: -1).  The state arrays of the problem are updated from the
 * records (and advanced when BLFIK_TICK_INTEGRATE is set); the per-task set-point buffers are not touched.
 * nchunks <= 0 selects zero-copy: if records, nu and status are page-locked and mapped (cudaHostAlloc /
 * cudaHostRegister / torch pin_memory), ONE launch reads the records and writes the results over PCIe itself, so
 * the transfers overlap the computation robot by robot; with pageable buffers it falls back to two staged chunks.  */
int32_t blfik_packed_record_len(const blfik_problem* problem);
int32_t blfik_packed_offset(const blfik_problem* problem, int32_t task);
/* BLFIK_TICK_NO_FEEDFORWARD in the flags of blfik_tick_packed selects the COMPACT record: every set-point shrinks to its
 * pose part -- SE3: p(3) q(4); CoM / R3: p(3); SO3: q(4); JointTracking: s_d(n); Distance 1; Gravity: direction(3);
 * AngularMomentum 3 -- followed by the task's exogenous feedback if it has one, and all feed-forward terms are zero.
 * That is the reference's setSetPoint(pose) with its default zero velocity argument (SE3Task.h, CoMTask.h,
 * JointTrackingTask::setSetPoint(jointPosition), JointTrackingTask.cpp:157-160); for ergoCub C3 a record shrinks from
 * 920 to 592 bytes, which matters when the host link is the bound (8 GPUs on one host).  The _ex queries describe
 * either layout (flags = 0 or BLFIK_TICK_NO_FEEDFORWARD); blfik_packed_part_len is the number of doubles of one task.   */
int32_t blfik_packed_record_len_ex(const blfik_problem* problem, uint32_t flags);
int32_t blfik_packed_offset_ex(const blfik_problem* problem, int32_t task, uint32_t flags);
int32_t blfik_packed_part_len(const blfik_problem* problem, int32_t task, uint32_t flags);
int blfik_tick_packed(blfik_problem* problem, const double* records, double* nu, int32_t* status,
                      double dt, uint32_t flags, int32_t nchunks);
/* A set-point that is the same for every robot of the batch (one record of blfik_setpoint_len doubles): what a caller
 * gets in the reference by handing the same vector to every instance's setSetPoint, e.g. the postural
 * JointTrackingTask::setSetPoint(jointPosition) of a regularisation task (JointTrackingTask.cpp:157-160).  The task then
 * reads that one record for all robots, and it leaves the packed host records (blfik_packed_offset returns -1 for it and
 * blfik_packed_record_len shrinks): for ergoCub C3 the postural set-point is 184 of the 592 bytes a compact record holds.
 * blfik_set_setpoint on the task goes back to one set-point per robot.  Not available for tasks with exogenous feedback. */
int blfik_set_setpoint_shared(blfik_problem* problem, int32_t task, const double* data, int32_t mem);

/* replaces QPInverseKinematics::getOutput() (QPInverseKinematics.cpp:553-556): nu [N][6+n] =
 * [base linear; base angular; joint velocity] per robot                                 */
int blfik_get_output(blfik_problem* problem, double* nu, int32_t mem);
/* per-instance blfik_status and ADMM iteration count (either may be NULL); host getters
 * synchronise.  blfik_count_unsolved is isOutputValid() for the batch.                  */
int blfik_get_status(blfik_problem* problem, int32_t* status, int32_t* iters, int32_t mem);
int blfik_count_unsolved(blfik_problem* problem, int64_t* count);

/* ---- stand-alone stages (K1, K4) ---------------------------------------------------- */
/* replaces kinDyn->getWorldTransform / getFrameFreeFloatingJacobian / getCenterOfMassPosition /
 * getCenterOfMassJacobian (call sites SE3Task.cpp:263-296, CoMTask.cpp:194-214) for the
 * current state.  Any output may be NULL.  frames: nframes indices.
 *   frame_pos [N][nframes][3], frame_rot [N][nframes][9], frame_jac [N][nframes][6][nv],
 *   com [N][3], com_jac [N][3][nv]                                                      */
int blfik_kinematics(blfik_problem* problem, const int32_t* frames, int32_t nframes,
                     double* frame_pos, double* frame_rot, double* frame_jac, double* com,
                     double* com_jac, int32_t mem);
/* No counterpart in the reference.  Same outputs as blfik_kinematics, for callers that must not
 * allocate device memory while they run: a hard-real-time control thread (cudaMalloc/cudaFree take
 * driver locks and synchronise the whole device, so one call can blow a 1 ms period), or a process
 * whose device memory is entirely owned by another allocator.  blfik_kinematics stages host-bound
 * results and the frame list in buffers it allocates and frees per call; this variant stages them in
 * a fixed BLFIK_NOALLOC_SCRATCH_BYTES array that lives inside the library image on every GPU that
 * loads it, and calls no cudaMalloc / cudaFree / allocating routine.
 *  - Limit: the frame list plus every requested output, each rounded up to 256 B, must fit in
 *    BLFIK_NOALLOC_SCRATCH_BYTES (ergoCub, two frames with Jacobians + CoM + CoM Jacobian: about
 *    1,100 robots).  Larger requests fail with BLFIK_ERR_UNSUPPORTED and a message that states the
 *    bytes needed and the maximum; nothing is launched.
 *  - The first call on each device looks up the device address of the array (with that device
 *    current).  Make that first call during set-up, before any CUDA stream capture begins.
 *  - All calls on one device share the one array, and it keeps whatever the previous call left in
 *    it.  Every byte a call reads back it has written itself in that call, so nothing needs
 *    clearing; but two calls that overlap on the GPU (two problems, two streams, two threads) race
 *    on it.  Serialise them.  The call returns after its results have reached the caller's buffers. */
#define BLFIK_NOALLOC_SCRATCH_BYTES (4u << 20)
int blfik_kinematics_noalloc(blfik_problem* problem, const int32_t* frames, int32_t nframes,
                             double* frame_pos, double* frame_rot, double* frame_jac, double* com,
                             double* com_jac, int32_t mem);
/* replaces dynamics->setControlInput({twist, sdot}) + integrator->integrate(0, dt)
 * (FloatingBaseSystemVelocityKinematics.cpp:80-84, FixedStepIntegrator.h:94-140);
 * nu == NULL integrates the last tick's output.  nsteps full Euler steps of dt.          */
int blfik_integrate(blfik_problem* problem, const double* nu, int32_t mem, double dt,
                    int32_t nsteps);

/* Stateless form of the same integrator for callers that own the state (the reference's
 * FloatingBaseSystemVelocityKinematics + ForwardEuler objects are independent of the IK,
 * ForwardEuler.h:96-124): advances base_pos [N][3], base_quat [N][4], s [N][n] in place by nsteps Euler
 * steps of dt under the held input nu [N][6+n].  mem = BLFIK_MEM_DEVICE works in place on `device`;
 * BLFIK_MEM_HOST stages through temporary device buffers and returns when the result is back.      */
int blfik_integrate_state(int32_t n_dofs, int64_t n_instances, double* base_pos, double* base_quat,
                          double* s, const double* nu, double dt, int32_t nsteps, int32_t mem,
                          int32_t device);

/* ---- introspection for parity tests -------------------------------------------------- */
typedef enum {
    BLFIK_DEBUG_H = 0,   /* [N][nv][nv]  Hessian  (QPInverseKinematics.cpp:442-454)      */
    BLFIK_DEBUG_G = 1,   /* [N][nv]      gradient                                        */
    BLFIK_DEBUG_CEQ = 2, /* [N][me][nv]  dense equality rows in addTask order            */
    BLFIK_DEBUG_BEQ = 3, /* [N][me]                                                      */
    BLFIK_DEBUG_LO = 4,  /* [N][n]  joint-velocity box = intersection of the +-identity
                            inequality rows (JointLimitsTask.cpp:227-231,
                            JointVelocityLimitsTask.cpp:140-141)                          */
    BLFIK_DEBUG_HI = 5   /* [N][n]                                                       */
} blfik_debug_what;
/* enable before a tick (allocates the dump buffers), then read after it                  */
int blfik_debug_enable(blfik_problem* problem, int32_t enabled);
int blfik_debug_get(blfik_problem* problem, int32_t what, double* out, int32_t mem);

/* ---- plumbing ------------------------------------------------------------------------ */
int32_t blfik_num_dofs(const blfik_problem* problem);
int32_t blfik_num_equality_rows(const blfik_problem* problem);
int64_t blfik_num_instances(const blfik_problem* problem);
int blfik_sync(blfik_problem* problem);
/* the problem's cudaStream_t (as void*); set_stream makes later work use the caller's stream */
void* blfik_get_stream(blfik_problem* problem);
int blfik_set_stream(blfik_problem* problem, void* cuda_stream);
/* Four code paths implement the tick.  For nv <= 32 velocity variables and <= 32 dense task rows (all ergoCub /
 * iCub-26 configurations): the two-launch path -- K1 + K2 one thread per robot, K3 + K4 one warp per robot, the QP of
 * every robot handed over through device memory (blfik_split.cuh; allocates N x ~5.5 KB on the first tick) -- and the
 * older fused register-resident kernel (one launch; also what the debug read-back and the zero-copy host tick use).  For
 * wider problems (<= 64 dense rows: the 32-joint iCubGazeboV2_5 configurations, with or without limit tasks) a kernel
 * that keeps the QP in factored form; and a generic shared-memory kernel for everything else.
 * path 0 = choose automatically (default), 1 = force the generic kernel, 2 = prefer the fused register-resident kernel
 * over the two-launch path (parity tests run all three).
 * blfik_get_kernel_path returns what the next tick runs: 4 two-launch path, 2 fused register-resident kernel,
 * 3 factored-form kernel, 1 generic kernel.
 * No reference counterpart (the reference has a single OSQP code path).                     */
int blfik_set_kernel_path(blfik_problem* problem, int32_t path);
int32_t blfik_get_kernel_path(const blfik_problem* problem);
/* Measurement utility (no reference counterpart): with timing enabled, the last full-batch blfik_tick on the two-launch
 * path is bracketed by CUDA events on the problem's stream; blfik_get_kernel_times waits for it and returns the duration
 * of the assemble kernel (K1 + K2) and of the solve kernel (K3 + K4) in milliseconds.  bench.py's roofline uses it.   */
int blfik_kernel_timing(blfik_problem* problem, int32_t enabled);
int blfik_get_kernel_times(blfik_problem* problem, double* ms_assemble, double* ms_solve);
/* ---- first pieces of the TSID path ---------------------------------------------------------------------
 * replaces FloatingBaseSystemAccelerationKinematics + ForwardEuler::integrate (FloatingBaseSystemAccelerationKinematics.cpp:20-75,
 * ForwardEuler.h:72-124) for N robots: state base_pos [N][3], base_quat [N][4], s [N][n], base linear velocity dp [N][3], base
 * angular velocity omega [N][3] (inertial frame), joint velocity ds [N][n]; input acc [N][6+n] = [base linear; base angular;
 * joint accelerations] (MIXED), held for nsteps Euler steps of dt.  Every component moves by its OLD derivative; the rotation
 * by the left-plus Exp(omega dt) R.  Host arrays in and out.                                                          */
int blfik_integrate_acceleration_state(int32_t n_dofs, int64_t n_instances, double* base_pos, double* base_quat, double* s,
                                       double* base_lin_vel, double* base_ang_vel, double* joint_vel, const double* acc,
                                       double dt, int32_t nsteps, int32_t device);
/* replaces kinDyn->getFreeFloatingMassMatrix / generalizedGravityForces (what the TSID dynamics tasks read,
 * src/TSID/src/BaseDynamicsTask.cpp, JointDynamicsTask.cpp) for the state of the problem: mass_matrix [N][6+n][6+n] in the MIXED
 * representation, gravity_forces [N][6+n] = M_total J_com' (0, 0, gravity_acceleration); either may be NULL.  Needs the link
 * inertias in the model (blfik_model_desc.inertia; NULL = point masses).  Coriolis / centrifugal forces are not computed.  */
int blfik_dynamics(blfik_problem* problem, double* mass_matrix, double* gravity_forces, double gravity_acceleration, int32_t mem);
/* replaces kinDyn->generalizedBiasForces (src/TSID/src/BaseDynamicsTask.cpp:178, JointDynamicsTask.cpp:233), getFrameBiasAcc
 * (src/TSID/src/SE3Task.cpp:254, SO3Task.cpp:189, R3Task.cpp:219) and getCenterOfMassBiasAcc (src/TSID/src/CoMTask.cpp:195) for the
 * state of the problem and the generalised velocity nu [N][6+n] (MIXED: base linear and angular velocity in world coordinates,
 * joint velocities): bias_forces [N][6+n] = C(q, nu) nu + g(q), the h of M nu' + h = S'tau + sum J'f; frame_bias_acc
 * [N][nframes][6] = Jdot nu of the listed frames (linear, angular); frame_vel [N][nframes][6] = J nu of the same frames
 * (kinDyn->getFrameVel, the controller state of the TSID SE3Task, SE3Task.cpp:256-263); com_bias_acc [N][3] = Jdot_com nu.
 * Any output may be NULL.  With nu = 0 the bias forces equal blfik_dynamics' gravity forces.                          */
int blfik_bias_forces(blfik_problem* problem, const double* nu, double gravity_acceleration, double* bias_forces, const int32_t* frames,
                      int32_t nframes, double* frame_bias_acc, double* frame_vel, double* com_bias_acc, int32_t mem);
/* Page-locked host memory, visible to every device of the process (cudaHostAllocPortable | cudaHostAllocMapped), for the
 * buffers of blfik_tick_packed / blfik_tick_host: the host classes keep their packed records, nu and status there so that the
 * library above the C ABI needs no CUDA headers.  No reference counterpart (Eigen vectors in pageable memory).            */
int blfik_host_alloc(size_t bytes, void** out);
void blfik_host_free(void* ptr);
/* number of kernel launches issued so far by this problem                                */
int64_t blfik_launch_count(const blfik_problem* problem);
const char* blfik_last_error(void);

/* Measurement utility (no counterpart in the reference): sustained FP64 FMA throughput of
 * `device` in TFLOP/s from a dependency-free DFMA loop over all SMs, timed with CUDA events
 * for about `millis` ms.  It is the denominator of the roofline fraction bench.py reports,
 * because MEASURED_PEAKS.json carries no FP64 figure.                                     */
int blfik_measure_fp64_peak(int32_t device, double millis, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* BLFIK_H */
