/**
 * IK.h -- the IK:: classes of the reference with a batch dimension: IKLinearTask, SE3Task, CoMTask, SO3Task, R3Task, DistanceTask, GravityTask, AngularMomentumTask,
 * JointTrackingTask, JointLimitsTask, JointVelocityLimitsTask, IntegrationBasedIK{,State,Problem}, QPInverseKinematics.
 * Same class / method / parameter-key names and the same bool + log()->error convention as
 * the headers under src/IK/include/BipedalLocomotion/IK; N = 1 reproduces the reference semantics, N > 1 runs N robots per tick.
 * The numerical work happens in libblfik_cuda.so (include/blfik.h); tasks only describe themselves
 * (blfik_task_desc) and hold their batch of set-points.
 */
#ifndef BLF_B200_IK_H
#define BLF_B200_IK_H

#include <BipedalLocomotion/KinDynBatch.h>
#include <BipedalLocomotion/Math.h>

namespace BipedalLocomotion
{
namespace IK
{

/** IKLinearTask.h:31-43 */
class IKLinearTask : public System::LinearTask
{
public:
    virtual bool setKinDyn(std::shared_ptr<KinDynBatch> kinDyn)
    {
        if (!kinDyn || !kinDyn->isValid()) {
            log()->error("[{}::setKinDyn] Invalid kinDyn object.", className());
            return false;
        }
        m_kinDyn = kinDyn;
        return true;
    }
    bool setVariablesHandler(const System::VariablesHandler& variablesHandler) override
    {
        if (!m_isInitialized) {
            log()->error("[{}::setVariablesHandler] The task is not initialized. Please call initialize method.", className());
            return false;
        }
        if (!variablesHandler.getVariable(m_robotVelocityVariable.name, m_robotVelocityVariable)) {
            log()->error("[{}::setVariablesHandler] Unable to get the variable named {}.", className(), m_robotVelocityVariable.name);
            return false;
        }
        if (m_robotVelocityVariable.size != static_cast<std::ptrdiff_t>(m_kinDyn->getNrOfDegreesOfFreedom() + 6)) {
            log()->error("[{}::setVariablesHandler] The size of the robot velocity variable is different from the one expected. "
                         "Expected size: {}. Given size: {}.",
                         className(), m_kinDyn->getNrOfDegreesOfFreedom() + 6, m_robotVelocityVariable.size);
            return false;
        }
        m_nVars = variablesHandler.getNumberOfVariables();
        m_A.resize(size(), m_nVars);   // zero outside the robot-velocity column block (SE3Task.cpp:61-64)
        m_b.assign(size(), 0.0);
        return true;
    }
    bool isValid() const override { return m_isValid; }
    std::shared_ptr<KinDynBatch> kinDyn() const { return m_kinDyn; }

    // ---- hooks of the batched back-end (no counterpart in the reference)
    virtual const char* className() const = 0;
    /** describe this task to the C ABI; `keep` owns the arrays the descriptor points to */
    virtual bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const = 0;
    /** doubles per robot of the set-point record (0: the task takes none) */
    virtual std::size_t setPointLength() const { return 0; }
    const std::vector<double>& setPointBuffer() const { return m_sp; }
    unsigned long setPointVersion() const { return m_spVersion; }
    /** raw robot-major batch of set-points, [N][setPointLength()] */
    bool setSetPointBatch(const double* data)
    {
        if (!m_kinDyn || !data || setPointLength() == 0) return false;
        m_sp.assign(data, data + m_kinDyn->getBatchSize() * setPointLength());
        m_spVersion++;
        m_spUniform = false;
        m_isSetPointSetAtLeastOnce = true;
        mirrorSetPoint(-1);
        return true;
    }
    /** The solver's page-locked packed records ([N][stride] doubles) as the place where this task's set-point lives: from now
     *  on every setSetPoint / setFeedback also writes robot i's `spDoubles` leading doubles of the set-point at
     *  base[i * stride + offset] and its feedback behind them, so that advance() has nothing to gather.  base = nullptr unbinds. */
    void bindRecords(double* base, std::size_t stride, std::size_t offset, std::size_t spDoubles)
    {
        m_recBase = base, m_recStride = stride, m_recOffset = offset, m_recSp = spDoubles;
        if (base) mirrorSetPoint(-1), mirrorFeedback(-1);
    }
    /** the last setSetPoint addressed every robot at once (instance = -1) and none individually since: the solver hands such a
     *  set-point to the device once (blfik_set_setpoint_shared) instead of once per robot and period */
    bool setPointIsUniform() const { return m_spUniform && !m_sp.empty(); }
    /** robot whose A and b are materialised by update() (getA / getB); -1 = none (default when N > 1) */
    void setInspectedInstance(int i) { m_inspect = i; }
    unsigned long gainsVersion() const { return m_gainsVersion; }
    /** doubles per robot of the exogenous feedback (0: the task uses none); uploaded with blfik_set_feedback */
    virtual std::size_t feedbackLength() const { return 0; }
    const std::vector<double>& feedbackBuffer() const { return m_fb; }
    unsigned long feedbackVersion() const { return m_fbVersion; }
    /** SE3Task, CoMTask, R3Task: the origin of the set-point can follow a quintic spline evaluated on the device at the
     *  solver's time (QPInverseKinematics::setTime; blfik_set_setpoint_spline) instead of a Math::QuinticSpline evaluated
     *  on the host whose point goes to setSetPoint every period.  One spline per robot, all with the same knot times. */
    virtual bool acceptsSetPointSpline() const { return false; }
    /** coordinates of that spline: 3, or the number of joints for a JointTrackingTask (Math::QuinticSpline<VectorXd>) */
    virtual std::size_t splineDimension() const { return 3; }
    template <class SplineType> bool setSetPointSpline(const std::vector<SplineType>& splines)   // Math::QuinticSpline<T> or Math::CubicSpline<T>, T = Vector3d / VectorXd
    {
        const std::size_t N = m_kinDyn ? m_kinDyn->getBatchSize() : 0;
        if (!acceptsSetPointSpline()) {
            log()->error("[{}::setSetPointSpline] The task has no position set-point a spline could drive.", className());
            return false;
        }
        if (splines.size() != N) {
            log()->error("[{}::setSetPointSpline] One spline per robot is expected. Expected: {}. Given: {}.", className(), N, splines.size());
            return false;
        }
        const auto& first = splines.front().knots();
        if (first.size() < 2) {
            log()->error("[Spline::setKnots] The number of knots has to be at least 2.");
            return false;
        }
        const std::size_t dim = splineDimension();
        const std::size_t nbc = (SplineType::isCubic ? 2 : 4) * dim;   // the cubic spline takes the two end velocities only
        std::vector<double> times(first.size()), pos(N * first.size() * dim), bc(N * nbc, 0.0);
        for (std::size_t k = 0; k < first.size(); k++) times[k] = std::chrono::duration<double>(first[k].time).count();
        for (std::size_t i = 0; i < N; i++) {
            const auto& kn = splines[i].knots();
            if (kn.size() != first.size()) {
                log()->error("[{}::setSetPointSpline] The splines of the robots must share their knot times.", className());
                return false;
            }
            for (std::size_t k = 0; k < kn.size(); k++) {
                if (kn[k].time != first[k].time) {
                    log()->error("[{}::setSetPointSpline] The splines of the robots must share their knot times.", className());
                    return false;
                }
                if (kn[k].position.size() != dim) {
                    log()->error("[{}::setSetPointSpline] The size of the knot is different from the one expected. Expected: {} Retrieved {}.", className(), dim,
                                 kn[k].position.size());
                    return false;
                }
                for (std::size_t c = 0; c < dim; c++) pos[(i * kn.size() + k) * dim + c] = kn[k].position[c];
            }
            typename SplineType::DerivativeBoundaryConditions ic, fc;
            if (splines[i].boundaryConditions(ic, fc))
                for (std::size_t c = 0; c < dim; c++) {
                    if (SplineType::isCubic)
                        bc[i * nbc + c] = ic.velocity[c], bc[i * nbc + dim + c] = fc.velocity[c];
                    else
                        bc[i * nbc + c] = ic.velocity[c], bc[i * nbc + dim + c] = ic.acceleration[c], bc[i * nbc + 2 * dim + c] = fc.velocity[c],
                                     bc[i * nbc + 3 * dim + c] = fc.acceleration[c];
                }
        }
        for (std::size_t k = 0; k + 1 < times.size(); k++)
            if (!(times[k + 1] > times[k])) {
                log()->error("[Splines::computePhasesDuration] Two consecutive points have the same time coordinate, or the times decrease.");
                return false;
            }
        m_splineTimes = std::move(times), m_splinePositions = std::move(pos), m_splineBoundary = std::move(bc);
        m_splineCubic = SplineType::isCubic;
        m_splineVersion++;
        return true;
    }
    void clearSetPointSpline() { m_splineTimes.clear(), m_splineVersion++; }
    /** From the stored set-point to `targetRecords` ([N][setPointLength()], of which the pose part is read) along the
     *  minimum-jerk profile over `duration`, evaluated on the device from the solver's time at the next advance() on
     *  (blfik_set_setpoint_minimum_jerk; for the rotation this is Planners::SO3PlannerInertial with zero boundary conditions).
     *  SE3 / SO3 / CoM / R3 tasks; a non-positive duration switches it off. */
    virtual bool acceptsMinimumJerk() const { return acceptsSetPointSpline(); }
    bool setSetPointMinimumJerk(const std::vector<double>& targetRecords, std::chrono::nanoseconds duration)
    {
        if (!acceptsMinimumJerk()) {
            log()->error("[{}::setSetPointMinimumJerk] The task has no pose set-point such a generator could drive.", className());
            return false;
        }
        if (duration <= std::chrono::nanoseconds::zero()) {
            m_minJerkDuration = 0.0, m_minJerkVersion++;
            return true;
        }
        const std::size_t N = m_kinDyn ? m_kinDyn->getBatchSize() : 0;
        if (targetRecords.size() != N * setPointLength()) {
            log()->error("[{}::setSetPointMinimumJerk] One set-point record per robot is expected. Expected size: {}. Given size: {}.", className(),
                         N * setPointLength(), targetRecords.size());
            return false;
        }
        m_minJerkTarget = targetRecords, m_minJerkDuration = std::chrono::duration<double>(duration).count(), m_minJerkVersion++;
        return true;
    }
    const std::vector<double>& minimumJerkTarget() const { return m_minJerkTarget; }
    double minimumJerkDuration() const { return m_minJerkDuration; }
    unsigned long minimumJerkVersion() const { return m_minJerkVersion; }
    const std::vector<double>& splineTimes() const { return m_splineTimes; }
    const std::vector<double>& splinePositions() const { return m_splinePositions; }
    const std::vector<double>& splineBoundary() const { return m_splineBoundary; }
    bool splineIsCubic() const { return m_splineCubic; }
    unsigned long splineVersion() const { return m_splineVersion; }

protected:
    bool commonInitialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler, std::shared_ptr<const ParametersHandler::IParametersHandler>& ptr)
    {
        m_isInitialized = false;
        if (!m_kinDyn || !m_kinDyn->isValid()) {
            log()->error("[{}::initialize] KinDynComputations object is not valid.", className());
            return false;
        }
        ptr = paramHandler.lock();
        if (!ptr) {
            log()->error("[{}::initialize] The parameter handler is not valid.", className());
            return false;
        }
        if (!ptr->getParameter("robot_velocity_variable_name", m_robotVelocityVariable.name)) {
            log()->error("[{}::initialize] Error while retrieving the robot velocity variable.", className());
            return false;
        }
        return true;
    }
    void storeSetPoint(const double* rec, std::ptrdiff_t instance)
    {
        const std::size_t len = setPointLength(), N = m_kinDyn->getBatchSize();
        if (m_sp.size() != N * len) m_sp.assign(N * len, 0.0);
        const std::size_t a = instance < 0 ? 0 : static_cast<std::size_t>(instance), b = instance < 0 ? N : a + 1;
        for (std::size_t i = a; i < b && i < N; i++) std::copy_n(rec, len, m_sp.data() + i * len);
        m_spVersion++;
        m_spUniform = instance < 0;
        mirrorSetPoint(instance);
    }
    void mirrorSetPoint(std::ptrdiff_t instance)
    {
        if (!m_recBase || m_sp.empty() || m_recSp == 0) return;
        const std::size_t len = setPointLength(), N = m_kinDyn->getBatchSize();
        const std::size_t a = instance < 0 ? 0 : static_cast<std::size_t>(instance), b = instance < 0 ? N : a + 1;
        for (std::size_t i = a; i < b; i++) std::copy_n(m_sp.data() + i * len, m_recSp, m_recBase + i * m_recStride + m_recOffset);
    }
    void mirrorFeedback(std::ptrdiff_t instance)
    {
        const std::size_t len = feedbackLength();
        if (!m_recBase || m_fb.empty() || len == 0 || m_recSp == 0) return;
        const std::size_t N = m_kinDyn->getBatchSize();
        const std::size_t a = instance < 0 ? 0 : static_cast<std::size_t>(instance), b = instance < 0 ? N : a + 1;
        for (std::size_t i = a; i < b; i++) std::copy_n(m_fb.data() + i * len, len, m_recBase + i * m_recStride + m_recOffset + m_recSp);
    }
    /** writes `count` doubles at `first` of the feedback record of robot `instance` (-1: every robot); `init` is the
     *  record a fresh buffer starts from (the controllers' initial state) */
    void storeFeedback(const double* values, std::size_t first, std::size_t count, std::ptrdiff_t instance, const std::vector<double>& init)
    {
        const std::size_t len = feedbackLength(), N = m_kinDyn->getBatchSize();
        if (m_fb.size() != N * len) {
            m_fb.resize(N * len);
            for (std::size_t i = 0; i < N; i++)
                for (std::size_t c = 0; c < len; c++) m_fb[i * len + c] = init[c];
        }
        const std::size_t a = instance < 0 ? 0 : static_cast<std::size_t>(instance), b = instance < 0 ? N : a + 1;
        for (std::size_t i = a; i < b && i < N; i++) std::copy_n(values, count, m_fb.data() + i * len + first);
        m_fbVersion++;
        mirrorFeedback(instance);
    }
    int inspected() const { return m_inspect >= 0 ? m_inspect : ((m_kinDyn && m_kinDyn->getBatchSize() == 1) ? 0 : -1); }
    static bool readGain3(const ParametersHandler::IParametersHandler& h, const std::string& key, Vector3d& out)
    {
        double scalar;
        std::vector<double> vec;
        if (h.getParameter(key, scalar)) {   // SE3Task.cpp:135-168: a scalar or a 3-vector
            out = {scalar, scalar, scalar};
            return true;
        }
        if (h.getParameter(key, vec) && vec.size() == 3) {
            out = {vec[0], vec[1], vec[2]};
            return true;
        }
        return false;
    }
    static bool readMask(const ParametersHandler::IParametersHandler& h, std::array<bool, 3>& mask, std::size_t& dofs)
    {
        std::vector<bool> m;
        mask = {true, true, true};
        if (h.getParameter("mask", m)) {
            if (m.size() != 3) return false;
            mask = {m[0], m[1], m[2]};
        }
        dofs = static_cast<std::size_t>(mask[0]) + mask[1] + mask[2];
        return true;
    }

    std::shared_ptr<KinDynBatch> m_kinDyn;
    System::VariablesHandler::VariableDescription m_robotVelocityVariable;
    std::size_t m_nVars{0};
    bool m_isInitialized{false}, m_isValid{false}, m_isSetPointSetAtLeastOnce{false};
    std::vector<double> m_sp, m_fb, m_splineTimes, m_splinePositions, m_splineBoundary;
    unsigned long m_spVersion{0}, m_gainsVersion{0}, m_fbVersion{0}, m_splineVersion{0};
    int m_inspect{-1};
    bool m_splineCubic{false};
    std::vector<double> m_minJerkTarget;
    double m_minJerkDuration{0.0};
    unsigned long m_minJerkVersion{0};
    double* m_recBase{nullptr};
    std::size_t m_recStride{0}, m_recOffset{0}, m_recSp{0};
    bool m_spUniform{false};
};

using IKLinearTaskFactory = System::Factory<IKLinearTask>;
#define BLF_REGISTER_IK_TASK(_type) \
    static inline const std::string _type##_registered = ::BipedalLocomotion::IK::IKLinearTaskFactory::registerBuilder(#_type, []() -> std::shared_ptr<::BipedalLocomotion::IK::IKLinearTask> { return std::make_shared<_type>(); })

/** SE3Task.h:51-205, SE3Task.cpp:21-390 */
class SE3Task : public IKLinearTask
{
public:
    enum class Mode { Enable, Disable };
    const char* className() const override { return "SE3Task"; }
    bool acceptsSetPointSpline() const override { return true; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setSetPoint(const SE3d& I_H_F, const SE3d::Tangent& mixedVelocity = SE3d::Tangent::Zero(), std::ptrdiff_t instance = -1);
    /** SE3Task.cpp:338-366: measured pose for the controllers; each part is used only if its use_*_exogenous_feedback key is set */
    bool setFeedback(const SE3d& I_H_F, std::ptrdiff_t instance = -1);
    bool setFeedback(const Vector3d& I_p_F, std::ptrdiff_t instance = -1);
    bool setFeedback(const SO3d& I_R_F, std::ptrdiff_t instance = -1);
    std::size_t feedbackLength() const override { return (m_usePositionExogenousFeedback || m_useOrientationExogenousFeedback) ? 7 : 0; }
    std::size_t size() const override { return m_linearDoFs + 3; }
    Type type() const override { return Type::equality; }
    void setTaskControllerMode(Mode mode) { m_controllerMode = mode, m_gainsVersion++; }
    Mode getTaskControllerMode() const { return m_controllerMode; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return 13; }

private:
    int m_frameIndex{-1};
    std::string m_frameName;
    Vector3d m_kpLinear{0, 0, 0}, m_kpAngular{0, 0, 0};
    std::array<bool, 3> m_mask{true, true, true};
    std::size_t m_linearDoFs{3};
    Mode m_controllerMode{Mode::Enable};
    bool m_usePositionExogenousFeedback{false}, m_useOrientationExogenousFeedback{false};
    BLF_REGISTER_IK_TASK(SE3Task);
};

/** CoMTask.cpp:19-273 */
class CoMTask : public IKLinearTask
{
public:
    const char* className() const override { return "CoMTask"; }
    bool acceptsSetPointSpline() const override { return true; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setSetPoint(const Vector3d& position, const Vector3d& velocity = {0, 0, 0}, std::ptrdiff_t instance = -1);
    /** CoMTask.cpp:262-273: measured CoM for the controller; used only with use_exogenous_feedback */
    bool setFeedback(const Vector3d& feedback, std::ptrdiff_t instance = -1);
    std::size_t feedbackLength() const override { return m_useExogenousFeedback ? 3 : 0; }
    std::size_t size() const override { return m_DoFs; }
    Type type() const override { return Type::equality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return 6; }

private:
    Vector3d m_kpLinear{0, 0, 0};
    std::array<bool, 3> m_mask{true, true, true};
    std::size_t m_DoFs{3};
    bool m_useExogenousFeedback{false};
    BLF_REGISTER_IK_TASK(CoMTask);
};

/** SO3Task.cpp:67-206 */
class SO3Task : public IKLinearTask
{
public:
    const char* className() const override { return "SO3Task"; }
    bool acceptsMinimumJerk() const override { return true; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setSetPoint(const SO3d& I_R_F, const SO3Tangentd& angularVelocity = SO3Tangentd::Zero(), std::ptrdiff_t instance = -1);
    std::size_t size() const override { return 3; }
    Type type() const override { return Type::equality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return 7; }

private:
    int m_frameIndex{-1};
    Vector3d m_kpAngular{0, 0, 0};
    BLF_REGISTER_IK_TASK(SO3Task);
};

/** R3Task.cpp:60-256 */
class R3Task : public IKLinearTask
{
public:
    const char* className() const override { return "R3Task"; }
    bool acceptsSetPointSpline() const override { return true; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setSetPoint(const Vector3d& I_p_F, const Vector3d& velocity = {0, 0, 0}, std::ptrdiff_t instance = -1);
    std::size_t size() const override { return m_DoFs; }
    Type type() const override { return Type::equality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return 6; }

private:
    int m_frameIndex{-1};
    Vector3d m_kpLinear{0, 0, 0};
    std::array<bool, 3> m_mask{true, true, true};
    std::size_t m_DoFs{3};
    BLF_REGISTER_IK_TASK(R3Task);
};

/** DistanceTask.h / DistanceTask.cpp:95-265: one row keeping |p_target - p_reference| at a desired value
 *  (reference_frame_name "" = the world origin). */
class DistanceTask : public IKLinearTask
{
public:
    const char* className() const override { return "DistanceTask"; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setDesiredDistance(double desiredDistance, std::ptrdiff_t instance = -1);
    bool setSetPoint(double desiredDistance, std::ptrdiff_t instance = -1) { return setDesiredDistance(desiredDistance, instance); }
    /** distance of the inspected robot at the last update() (DistanceTask.cpp:251-254) */
    double getDistance() const { return m_computedDistance; }
    std::size_t size() const override { return 1; }
    Type type() const override { return Type::equality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return 1; }

private:
    int m_targetFrameIndex{-1}, m_referenceFrameIndex{-1};
    double m_kp{0.0}, m_distanceNumericThreshold{0.001}, m_computedDistance{0.0};
    BLF_REGISTER_IK_TASK(DistanceTask);
};

/** GravityTask.h / GravityTask.cpp:76-250: two rows aligning the z axis of a frame with a desired direction given in that frame. */
class GravityTask : public IKLinearTask
{
public:
    const char* className() const override { return "GravityTask"; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setDesiredGravityDirectionInTargetFrame(const Vector3d& desiredGravityDirection, std::ptrdiff_t instance = -1);
    bool setFeedForwardVelocityInTargetFrame(const Vector3d& feedforwardVelocity, std::ptrdiff_t instance = -1);
    bool setSetPoint(const Vector3d& desiredGravityDirection, const Vector3d& feedforwardVelocity = {0, 0, 0}, std::ptrdiff_t instance = -1);
    std::size_t size() const override { return 2; }
    Type type() const override { return Type::equality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return 6; }

private:
    void storePart(const Vector3d& v, int first, std::ptrdiff_t instance);
    int m_targetFrameIndex{-1};
    double m_kp{0.0};
    BLF_REGISTER_IK_TASK(GravityTask);
};

/** AngularMomentumTask.h / AngularMomentumTask.cpp:44-230: A = (masked) rows 3..5 of the centroidal momentum matrix,
 *  b = desired centroidal angular momentum; key: mask (optional).  The rows exist on the device only: update() checks
 *  the state of the task, getA() is not materialised for this task. */
class AngularMomentumTask : public IKLinearTask
{
public:
    const char* className() const override { return "AngularMomentumTask"; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setSetPoint(const Vector3d& desiredAngularMomentum, std::ptrdiff_t instance = -1);
    std::size_t size() const override { return m_DoFs; }
    Type type() const override { return Type::equality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return 3; }

private:
    std::array<bool, 3> m_mask{true, true, true};
    std::size_t m_DoFs{3};
    BLF_REGISTER_IK_TASK(AngularMomentumTask);
};

/** JointTrackingTask.cpp:29-210 */
class JointTrackingTask : public IKLinearTask
{
public:
    const char* className() const override { return "JointTrackingTask"; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    bool setSetPoint(const VectorXd& jointPosition, std::ptrdiff_t instance = -1);
    bool setSetPoint(const VectorXd& jointPosition, const VectorXd& jointVelocity, std::ptrdiff_t instance = -1);
    std::size_t size() const override;
    Type type() const override { return Type::equality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;
    std::size_t setPointLength() const override { return m_kinDyn ? 2 * m_kinDyn->getNrOfDegreesOfFreedom() : 0; }
    /** the joint positions of the set-point can follow a Math::QuinticSpline<VectorXd> / CubicSpline<VectorXd> evaluated on the
     *  device (position = the spline, velocity = the stored feed-forward + its derivative) */
    bool acceptsSetPointSpline() const override { return true; }
    std::size_t splineDimension() const override { return m_kinDyn ? m_kinDyn->getNrOfDegreesOfFreedom() : 0; }
    bool acceptsMinimumJerk() const override { return false; }

private:
    VectorXd m_kp;
    BLF_REGISTER_IK_TASK(JointTrackingTask);
};

/** JointLimitsTask.cpp:22-284 */
class JointLimitsTask : public IKLinearTask
{
public:
    const char* className() const override { return "JointLimitsTask"; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    std::size_t size() const override;
    Type type() const override { return Type::inequality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;

private:
    VectorXd m_klim, m_upperLimits, m_lowerLimits;
    std::chrono::nanoseconds m_samplingTime{0};
    BLF_REGISTER_IK_TASK(JointLimitsTask);
};

/** JointVelocityLimitsTask.cpp:22-171 */
class JointVelocityLimitsTask : public IKLinearTask
{
public:
    const char* className() const override { return "JointVelocityLimitsTask"; }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler) override;
    bool update() override;
    std::size_t size() const override;
    Type type() const override { return Type::inequality; }
    bool fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const override;

private:
    VectorXd m_upperLimits, m_lowerLimits;
    BLF_REGISTER_IK_TASK(JointVelocityLimitsTask);
};

/** IntegrationBasedIK.h:31-35 */
struct IntegrationBasedIKState {
    VectorXd jointVelocity;
    SE3d::Tangent baseVelocity;
};

class IntegrationBasedIK : public System::ILinearTaskSolver<IKLinearTask, IntegrationBasedIKState>
{
public:
    virtual ~IntegrationBasedIK() = default;
};

/** IntegrationBasedIK.h:83-117 (tuple-like: auto [vh, weights, ik] = build(...)) */
struct IntegrationBasedIKProblem {
    System::VariablesHandler variablesHandler;
    std::unordered_map<std::string, std::shared_ptr<System::WeightProvider>> weights;
    std::unique_ptr<IntegrationBasedIK> ik;
    bool isValid() const { return ik != nullptr; }
};

/** QPInverseKinematics.h:54-304, QPInverseKinematics.cpp:26-806 */
class QPInverseKinematics : public IntegrationBasedIK
{
    struct Impl;
    std::unique_ptr<Impl> m_pimpl;
    bool tick(double dt, uint32_t flags);

public:
    QPInverseKinematics();
    ~QPInverseKinematics();

    bool addTask(std::shared_ptr<Task> task, const std::string& taskName, std::size_t priority,
                 std::shared_ptr<const System::WeightProviderPort> weightProvider = nullptr) override;
    bool addTask(std::shared_ptr<Task> task, const std::string& taskName, std::size_t priority, const VectorXd& weight) override;
    bool setTaskWeight(const std::string& taskName, std::shared_ptr<const System::WeightProviderPort> weightProvider) override;
    bool setTaskWeight(const std::string& taskName, const VectorXd& weight) override;
    std::weak_ptr<const System::WeightProviderPort> getTaskWeightProvider(const std::string& taskName) const override;
    bool finalize(const System::VariablesHandler& handler) override;
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler) override;
    std::vector<std::string> getTaskNames() const override;
    bool isOutputValid() const override;
    bool advance() override;
    const State& getOutput() const override;
    std::weak_ptr<Task> getTask(const std::string& name) const override;
    std::string toString() const override;
    const VectorXd& getRawSolution() const override;
    static IntegrationBasedIKProblem build(std::weak_ptr<const ParametersHandler::IParametersHandler> handler, std::shared_ptr<KinDynBatch> kinDyn);

    // ---- batch extensions (no counterpart in the reference)
    /** [N][6+n] velocities of the last advance(), robot-major */
    const std::vector<double>& getBatchOutput() const;
    /** per-robot blfik_status of the last advance() */
    const std::vector<int32_t>& getBatchStatus() const;
    /** the same two arrays where the device left them (page-locked host memory for batches of 64 robots and more): no copy */
    const double* batchOutputData() const;
    const int32_t* batchStatusData() const;
    /** advance() and apply ForwardEuler on the device in the same launch.  syncHostState = true: the new state is read back
     *  into kinDyn (the reference's caller does integrator->getSolution() + kinDyn->setRobotState every period).  false: the
     *  state stays on the device, the next advance() / advanceAndIntegrate() starts from it (unless kinDyn->setRobotState is
     *  called in between), and syncState() fetches it when the caller wants it. */
    bool advanceAndIntegrate(std::chrono::nanoseconds dT, bool syncHostState = true);
    bool syncState();
    /** Spread the batch over several GPUs of this host: contiguous instance blocks, one blfik_problem and one host thread per
     *  device, no data-path collective.  Call before finalize(); default: the device of the KinDynBatch. */
    bool setDevices(const std::vector<int>& devices);
    /** The caller promises that every set-point carries zero feed-forward velocities (the reference's setSetPoint(pose) with
     *  its default zero velocity): the host records then hold the pose part only (BLFIK_TICK_NO_FEEDFORWARD).  Before finalize(). */
    void setCompactRecords(bool on);
    /** advance() starts every solve from the previous solution, as the reference warm-starts OSQP (QPInverseKinematics.cpp:494-509);
     *  false: every tick solves from scratch (what bench.py times) */
    void setWarmStart(bool on);
    /** true: a robot whose QP is not solved inside the iteration budget counts as failed (BLFIK_TICK_STRICT) -- what the reference
     *  does when OSQP proves the problem infeasible or hits its iteration limit (QPInverseKinematics.cpp:518-523); default false:
     *  the best iterate is reported like OSQP's "solved inaccurate". */
    void setStrictFailure(bool on);
    /** true if robot i's last solution is usable (status SOLVED or MAX_ITER, as the reference accepts Solved and
     *  SolvedInaccurate, QPInverseKinematics.cpp:518-523) */
    bool isOutputValid(std::size_t instance) const;
    using IntegrationBasedIK::isOutputValid;
    /** time at which the on-device set-point generators (IKLinearTask::setSetPointSpline) are evaluated by the next advance();
     *  advanceAndIntegrate moves it on by dT */
    bool setTime(std::chrono::nanoseconds t);
    blfik_problem* handle();
};

/** QPFixedBaseInverseKinematics.h:54-96, QPFixedBaseInverseKinematics.cpp:17-122: QPInverseKinematics plus a priority-0
 *  SE3Task on the floating-base link with the controller disabled and a zero feed-forward ("base_se3_task"), i.e. the
 *  base twist is constrained to zero. */
class QPFixedBaseInverseKinematics : public QPInverseKinematics
{
    std::string m_baseLink;
    std::shared_ptr<SE3Task> m_baseSE3Task;
    bool m_isKinDynSet{false};

public:
    QPFixedBaseInverseKinematics();
    ~QPFixedBaseInverseKinematics();
    /** required keys: "robot_velocity_variable_name" (+ "verbosity", optional) */
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler) override;
    bool setKinDyn(std::shared_ptr<KinDynBatch> kinDyn);
};

}  // namespace IK
}  // namespace BipedalLocomotion

#endif
