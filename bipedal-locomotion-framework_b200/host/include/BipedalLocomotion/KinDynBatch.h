/**
 * KinDynBatch.h -- stands in for std::shared_ptr<iDynTree::KinDynComputations> (IKLinearTask.h:40): ONE object,
 * shared by every task and mutated by the caller between ticks (setRobotState), here holding the state of a
 * whole batch of N robots.  Kinematic queries (getWorldTransform, getFrameFreeFloatingJacobian,
 * getCenterOfMassPosition / Jacobian; call sites SE3Task.cpp:263-296, CoMTask.cpp:194-214) run on the GPU through
 * blfik_kinematics -- there is no CPU evaluation in this class.  Velocities are always MIXED
 * (the tasks refuse anything else, SE3Task.cpp:92-102).
 */
#ifndef BLF_B200_KINDYN_BATCH_H
#define BLF_B200_KINDYN_BATCH_H

#include <BipedalLocomotion/System.h>
#include <blfik.h>

namespace BipedalLocomotion
{

class KinDynBatch
{
public:
    KinDynBatch() = default;
    ~KinDynBatch();
    KinDynBatch(const KinDynBatch&) = delete;
    KinDynBatch& operator=(const KinDynBatch&) = delete;

    /** Reads the plain-text reduced-model format written by blf_b200.models.KinematicModel.to_text()
     *  (the image has no URDF parser and no URDF). */
    bool loadRobotModelFromFile(const std::string& path);
    /** number of robots evaluated at once (default 1 = the reference's semantics) and the CUDA device */
    bool setBatchSize(std::size_t n, int device = 0);
    std::size_t getBatchSize() const { return m_N; }
    int getDevice() const { return m_device; }

    bool isValid() const { return m_model != nullptr; }
    std::size_t getNrOfDegreesOfFreedom() const { return m_n; }
    /** -1 if the frame does not exist (iDynTree::FRAME_INVALID_INDEX) */
    int getFrameIndex(const std::string& name) const;
    std::string getFrameName(int index) const;
    /** name of the floating-base link frame (iDynTree::KinDynComputations::getFloatingBase); link 0 of the model */
    const std::string& getFloatingBase() const { return m_floatingBase; }
    const VectorXd& jointLowerLimits() const { return m_qLower; }
    const VectorXd& jointUpperLimits() const { return m_qUpper; }
    const std::vector<std::string>& jointNames() const { return m_jointNames; }

    /** kinDyn->setRobotState(world_H_base, s, ...) for one robot of the batch */
    bool setRobotState(const Vector3d& basePosition, const SO3d& baseRotation, const VectorXd& jointPositions, std::size_t instance = 0);
    /** robot-major batch arrays: p [N][3], q [N][4] (x,y,z,w), s [N][n] */
    bool setRobotStateBatch(const double* p, const double* q, const double* s);
    bool getJointPos(VectorXd& s, std::size_t instance = 0) const;

    // ---- kinematic queries of one robot (evaluated for the whole batch on the GPU, cached per state)
    bool getWorldTransform(int frame, SE3d& pose, std::array<double, 9>& rotation, std::size_t instance = 0);
    bool getFrameFreeFloatingJacobian(int frame, System::MatrixXd& jacobian, std::size_t instance = 0);
    bool getCenterOfMassPosition(Vector3d& com, std::size_t instance = 0);
    bool getCenterOfMassJacobian(System::MatrixXd& jacobian, std::size_t instance = 0);

    // ---- dynamic quantities the TSID tasks read (iDynTree::KinDynComputations names), MIXED representation,
    // evaluated for the whole batch on the GPU (blfik_dynamics, blfik_bias_forces) and cached per state / velocity
    /** kinDyn->setRobotState(world_H_base, s, base_velocity, s_dot, world_gravity): the velocity part; baseVelocity = linear and
     *  angular velocity of the base in world coordinates.  The gravity vector is (0, 0, -gravityAcceleration). */
    bool setRobotVelocity(const std::array<double, 6>& baseVelocity, const VectorXd& jointVelocities, std::size_t instance = 0);
    /** [N][6 + n] robot-major */
    bool setRobotVelocityBatch(const double* nu);
    void setGravityAcceleration(double g) { m_gravity = g, m_dynVersion = 0, m_biasVersion = 0; }
    bool getFreeFloatingMassMatrix(System::MatrixXd& massMatrix, std::size_t instance = 0);
    bool generalizedGravityForces(VectorXd& forces, std::size_t instance = 0);
    /** C(q, nu) nu + g(q): the h of M nu' + h = S'tau + sum J'f (BaseDynamicsTask.cpp:178, JointDynamicsTask.cpp:233) */
    bool generalizedBiasForces(VectorXd& forces, std::size_t instance = 0);
    /** Jdot nu of a frame, linear then angular (TSID SE3Task.cpp:254, SO3Task.cpp:189, R3Task.cpp:219) */
    bool getFrameBiasAcc(int frame, std::array<double, 6>& biasAcc, std::size_t instance = 0);
    /** J nu of a frame, linear then angular (TSID SE3Task.cpp:256-263) */
    bool getFrameVel(int frame, std::array<double, 6>& velocity, std::size_t instance = 0);
    /** Jdot_com nu (TSID CoMTask.cpp:195) */
    bool getCenterOfMassBiasAcc(Vector3d& biasAcc, std::size_t instance = 0);

    // ---- used by the solver
    const blfik_model* model() const { return m_model; }
    const std::vector<double>& basePositions() const { return m_p; }
    const std::vector<double>& baseQuaternions() const { return m_q; }
    const std::vector<double>& jointPositions() const { return m_s; }
    unsigned long stateVersion() const { return m_version; }
    /** The solver's page-locked packed records as a second home of the state: setRobotState then also writes robot i's
     *  p(3) q(4) s(n) at base[i * stride], so that advance() has nothing to gather.  base = nullptr unbinds. */
    void bindRecords(double* base, std::size_t stride);

private:
    bool ensureKinProblem();
    bool evalFrame(int frame);
    bool evalCom();
    bool evalDynamics();
    bool evalBias(int frame);

    double* m_recBase{nullptr};
    std::size_t m_recStride{0};
    void mirrorState(std::size_t a, std::size_t b);
    blfik_model* m_model{nullptr};
    blfik_problem* m_kin{nullptr};  // kinematics-only problem (owns the device copy of the state for queries)
    std::size_t m_n{0}, m_N{1};
    int m_device{0};
    std::vector<std::string> m_jointNames, m_frameNames;
    std::string m_floatingBase;
    VectorXd m_qLower, m_qUpper;
    std::vector<double> m_p, m_q, m_s;
    unsigned long m_version{1}, m_kinVersion{0};
    // caches
    int m_cachedFrame{-1};
    unsigned long m_frameVersion{0}, m_comVersion{0};
    std::vector<double> m_framePos, m_frameRot, m_frameJac, m_com, m_comJac;
    // dynamics
    std::vector<double> m_nu, m_mass, m_gravityForces, m_bias, m_frameBias, m_frameVel, m_comBias;
    double m_gravity{9.80665};
    unsigned long m_velVersion{1}, m_dynVersion{0}, m_biasVersion{0}, m_biasVelVersion{0};
    int m_biasFrame{-1};
};

}  // namespace BipedalLocomotion

#endif
