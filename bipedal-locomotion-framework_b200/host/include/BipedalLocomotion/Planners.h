/**
 * Planners.h -- Planners::SO3Planner, the minimum-jerk rotation trajectory the reference's callers evaluate every period and
 * pass to SO3Task / SE3Task::setSetPoint.
 *
 * Same names, argument meaning and bool + log()->error convention as
 *   src/Planners/include/BipedalLocomotion/Planners/SO3Planner.h   (:23-161)
 *   src/Planners/include/BipedalLocomotion/Planners/SO3Planner.tpp (:24-310)
 * with SO3d / SO3Tangentd of Core.h in place of manif's.  This is the host-side, one-trajectory class; for a batch of robots
 * blfik_set_setpoint_minimum_jerk evaluates SO3PlannerInertial with zero boundary conditions on the device.
 *
 *   Right trivialisation (inertial velocity):  R(t) = Exp(s(t) d) R_0,  d = Log(R_1 R_0^-1)
 *   Left trivialisation (body velocity):       R(t) = R_0 Exp(s(t) d),  d = Log(R_0^-1 R_1)
 * s: the quintic with s(0) = 0, s(T) = 1 that meets the boundary velocities / accelerations, which must be parallel to d.
 */
#ifndef BLF_B200_PLANNERS_H
#define BLF_B200_PLANNERS_H

#include <algorithm>
#include <chrono>
#include <cmath>

#include <memory>

#include <BipedalLocomotion/Contacts.h>
#include <BipedalLocomotion/Core.h>
#include <BipedalLocomotion/Math.h>
#include <BipedalLocomotion/System.h>

namespace BipedalLocomotion
{
namespace Planners
{

enum class LieGroupTrivialization { Left, Right };

/** SO3Planner.h:31-38 */
struct SO3PlannerState {
    SO3d rotation;
    SO3Tangentd velocity;
    SO3Tangentd acceleration;
};

template <LieGroupTrivialization trivialization> class SO3Planner : public System::Source<SO3PlannerState>
{
public:
    bool setRotations(const SO3d& initialRotation, const SO3d& finalRotation, const std::chrono::nanoseconds& duration)
    {
        if (duration <= std::chrono::nanoseconds::zero()) {
            log()->error("[SO3Planner::setRotation] The trajectory duration must be a strictly positive number.");
            return false;
        }
        m_initialRotation = initialRotation, m_T = duration;
        m_distance = computeDistance(initialRotation, finalRotation);
        m_advanceCurrentTime = std::chrono::nanoseconds::zero();
        m_state = SO3PlannerState{};
        m_state.rotation = initialRotation;
        m_areCoefficientsComputed = false;
        return true;
    }
    bool setInitialConditions(const SO3Tangentd& velocity, const SO3Tangentd& acceleration)
    {
        m_initial = {velocity.w, acceleration.w, true}, m_areCoefficientsComputed = false;
        return true;
    }
    bool setFinalConditions(const SO3Tangentd& velocity, const SO3Tangentd& acceleration)
    {
        m_final = {velocity.w, acceleration.w, true}, m_areCoefficientsComputed = false;
        return true;
    }
    bool evaluatePoint(const std::chrono::nanoseconds& time, SO3d& rotation, SO3Tangentd& velocity, SO3Tangentd& acceleration)
    {
        constexpr auto logPrefix = "[SO3Planner::evaluatePoint]";
        if (time < std::chrono::nanoseconds::zero() || time > m_T) {
            log()->error("{} The time has to be a real positive number between 0 and {}.", logPrefix, seconds(m_T));
            return false;
        }
        if (!computeCoefficients()) {
            log()->error("{} Unable to compute the coefficients.", logPrefix);
            return false;
        }
        const double t = seconds(time);
        const double* a = m_a;
        const double s = a[0] + t * (a[1] + t * (a[2] + t * (a[3] + t * (a[4] + t * a[5]))));
        const double ds = a[1] + t * (2.0 * a[2] + t * (3.0 * a[3] + t * (4.0 * a[4] + t * 5.0 * a[5])));
        const double dds = 2.0 * a[2] + t * (6.0 * a[3] + t * (12.0 * a[4] + t * 20.0 * a[5]));
        const SO3d step = SO3d::Exp({s * m_distance[0], s * m_distance[1], s * m_distance[2]});
        rotation = trivialization == LieGroupTrivialization::Right ? step * m_initialRotation : m_initialRotation * step;
        velocity.w = {ds * m_distance[0], ds * m_distance[1], ds * m_distance[2]};
        acceleration.w = {dds * m_distance[0], dds * m_distance[1], dds * m_distance[2]};
        return true;
    }
    bool evaluatePoint(const std::chrono::nanoseconds& time, SO3PlannerState& state)
    {
        return evaluatePoint(time, state.rotation, state.velocity, state.acceleration);
    }
    bool setAdvanceTimeStep(const std::chrono::nanoseconds& dt)
    {
        if (dt <= std::chrono::nanoseconds::zero()) {
            log()->error("[SO3Planner::setAdvanceTimeStep] The time step of the advance time has to be a strictly positive number.");
            return false;
        }
        m_advanceTimeStep = dt;
        return true;
    }
    const SO3PlannerState& getOutput() const final { return m_state; }
    bool isOutputValid() const final { return m_advanceTimeStep != std::chrono::nanoseconds::zero(); }
    bool advance() final
    {
        if (!isOutputValid()) {
            log()->error("[SO3Planner::advance] The advance capabilities cannot be used. Did you set the advance time step?");
            return false;
        }
        m_advanceCurrentTime = std::min(m_advanceTimeStep + m_advanceCurrentTime, m_T);
        return evaluatePoint(m_advanceCurrentTime, m_state);
    }
    /** SO3Planner.tpp:289-305 */
    static SO3Tangentd projectTangentVector(const SO3d& initialRotation, const SO3d& finalRotation, const SO3Tangentd& vector)
    {
        const Vector3d d = computeDistance(initialRotation, finalRotation);
        const double n = norm(d);
        SO3Tangentd out;
        if (n < 1e-4) return out;
        const double k = dot(d, vector.w) / n;
        out.w = {k * d[0], k * d[1], k * d[2]};
        return out;
    }

private:
    struct Boundary {
        Vector3d velocity{0, 0, 0}, acceleration{0, 0, 0};
        bool isSet{false};
    };
    static double seconds(const std::chrono::nanoseconds& d) { return std::chrono::duration<double>(d).count(); }
    static double dot(const Vector3d& a, const Vector3d& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
    static double norm(const Vector3d& a) { return std::sqrt(dot(a, a)); }
    static Vector3d computeDistance(const SO3d& R0, const SO3d& R1)
    {
        return trivialization == LieGroupTrivialization::Right ? (R1 * R0.inverse()).Log() : (R0.inverse() * R1).Log();
    }
    /** the scalar k with vector = k * distance; false when the two are not parallel (:51-86, 1e-3 on the cosine) */
    bool coefficient(const Vector3d& vector, double& k) const
    {
        k = 0.0;
        const double nd = norm(m_distance), nv = norm(vector);
        if (nd == 0.0 || nv == 0.0) return true;
        if (std::abs(std::abs(dot(m_distance, vector) / (nd * nv)) - 1.0) > 1e-3) return false;
        for (int i = 0; i < 3; i++) {
            const double r = vector[i] / m_distance[i];
            if (std::isfinite(r)) {
                k = r;
                return true;
            }
        }
        return false;
    }
    bool computeCoefficients()
    {
        constexpr auto logPrefix = "[SO3Planner::computeCoefficients]";
        if (m_areCoefficientsComputed) return true;
        if (!m_initial.isSet) {
            log()->error("{} The initial boundary condition is not set.", logPrefix);
            return false;
        }
        if (!m_final.isSet) {
            log()->error("{} The final boundary condition is not set.", logPrefix);
            return false;
        }
        double kv0, ka0, kv1, ka1;
        if (!coefficient(m_initial.velocity, kv0) || !coefficient(m_initial.acceleration, ka0) || !coefficient(m_final.velocity, kv1)
            || !coefficient(m_final.acceleration, ka1)) {
            log()->error("{} The velocity and the boundary acceleration must be a linear dependent to the vector representing the axis "
                         "between the two rotations.", logPrefix);
            return false;
        }
        const double T = seconds(m_T);
        m_a[0] = 0.0, m_a[1] = kv0, m_a[2] = 0.5 * ka0;
        m_a[3] = (20.0 - (8.0 * kv1 + 12.0 * kv0) * T - (3.0 * ka0 - ka1) * T * T) / (2.0 * T * T * T);
        m_a[4] = (-30.0 + (14.0 * kv1 + 16.0 * kv0) * T + (3.0 * ka0 - 2.0 * ka1) * T * T) / (2.0 * T * T * T * T);
        m_a[5] = (12.0 - 6.0 * (kv1 + kv0) * T - (ka0 - ka1) * T * T) / (2.0 * T * T * T * T * T);
        m_areCoefficientsComputed = true;
        return true;
    }

    SO3d m_initialRotation;
    Vector3d m_distance{0, 0, 0};
    std::chrono::nanoseconds m_T{0}, m_advanceTimeStep{0}, m_advanceCurrentTime{0};
    SO3PlannerState m_state;
    Boundary m_initial, m_final;
    bool m_areCoefficientsComputed{false};
    double m_a[6]{0, 0, 0, 0, 0, 0};
};

using SO3PlannerInertial = SO3Planner<LieGroupTrivialization::Right>;
using SO3PlannerBody = SO3Planner<LieGroupTrivialization::Left>;

/** SwingFootPlanner.h:31-38 */
struct SwingFootPlannerState {
    SE3d transform;                       /**< link pose in the inertial frame */
    SE3d::Tangent mixedVelocity;          /**< 6D velocity, mixed representation */
    SE3d::Tangent mixedAcceleration;      /**< 6D acceleration, mixed representation */
    bool isInContact{true};
    std::chrono::nanoseconds time{0};
};

/**
 * SwingFootPlanner.h:44-196, SwingFootPlanner.cpp:29-558: from a ContactList to the pose / velocity / acceleration of a foot,
 * one sample per advance().  In contact: the contact's pose, zero velocity.  In swing: the trajectory to the next contact is
 * re-created on every advance() from the previous output (pose, velocity, acceleration) -- SO3PlannerInertial for the
 * orientation, a two-knot spline for x, y and a spline through the apex ("step_height" above the higher of the two contacts at
 * "foot_apex_time" of the swing) for z; "interpolation_method" = "min_acceleration" (CubicSpline, the default) or "min_jerk"
 * (QuinticSpline) -- so that a new contact list can be accepted in mid swing (setContactList's three cases).  The output goes to
 * SE3Task::setSetPoint(transform, mixedVelocity) every period.
 * Batch note: because re-planning from a point of the trajectory with the trajectory's own derivatives gives the same
 * trajectory, a swing can also be handed to the device once, at take-off: planOnDevice() returns the knots of ONE 3D spline
 * (the x, y value at the apex time taken from the two-knot planar spline, which is that spline) and the landing rotation for
 * IKLinearTask::setSetPointSpline + setSetPointMinimumJerk, and the tick kernels evaluate it without the host.
 */
class SwingFootPlanner : public System::Source<SwingFootPlannerState>
{
public:
    using Spline1 = Math::Spline<VectorXd>;

    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler)
    {
        constexpr auto logPrefix = "[SwingFootPlanner::initialize]";
        auto ptr = handler.lock();
        if (ptr == nullptr) {
            log()->error("{} The handler is not correctly initialized.", logPrefix);
            return false;
        }
        if (!ptr->getParameter("sampling_time", m_dT)) {
            log()->error("{} Unable to initialize the sampling time.", logPrefix);
            return false;
        }
        if (!ptr->getParameter("step_height", m_stepHeight)) {
            log()->error("{} Unable to initialize the step height.", logPrefix);
            return false;
        }
        if (!ptr->getParameter("foot_apex_time", m_footApexTime)) {
            log()->error("{} Unable to initialize the foot apex time.", logPrefix);
            return false;
        }
        ptr->getParameter("foot_take_off_velocity", m_footTakeOffVelocity);          // optional keys keep their defaults (:56-95)
        ptr->getParameter("foot_take_off_acceleration", m_footTakeOffAcceleration);
        ptr->getParameter("foot_landing_velocity", m_footLandingVelocity);
        ptr->getParameter("foot_landing_acceleration", m_footLandingAcceleration);
        ptr->getParameter("position_tolerance", m_positionTolerance);
        ptr->getParameter("orientation_tolerance", m_orientationTolerance);
        if (!(m_dT > std::chrono::nanoseconds::zero() && m_footApexTime > 0 && m_footApexTime < 1)) {
            log()->error("{} The sampling time should be a positive number and the foot_apex_time must be strictly between 0 and 1.", logPrefix);
            return false;
        }
        std::string method{"min_acceleration"};
        ptr->getParameter("interpolation_method", method);
        if (method != "min_acceleration" && method != "min_jerk") {
            log()->warn("{} The parameter named 'interpolation_method' must be equal to 'min_acceleration' or 'min_jerk'. The "
                        "'min_acceleration' method will be used.", logPrefix);
            method = "min_acceleration";
        }
        m_minJerk = method == "min_jerk";
        if (m_minJerk) {
            m_planarPlanner = std::make_unique<Math::QuinticSpline<VectorXd>>();
            m_heightPlanner = std::make_unique<Math::QuinticSpline<VectorXd>>();
        } else {
            m_planarPlanner = std::make_unique<Math::CubicSpline<VectorXd>>();
            m_heightPlanner = std::make_unique<Math::CubicSpline<VectorXd>>();
        }
        return true;
    }

    /** SwingFootPlanner.cpp:142-292 */
    bool setContactList(const Contacts::ContactList& contactList)
    {
        constexpr auto logPrefix = "[SwingFootPlanner::setContactList]";
        if (contactList.size() == 0) {
            log()->error("{} The provided contact list is empty.", logPrefix);
            return false;
        }
        if (m_contactList.size() == 0 || !m_isOutputValid) {   // case 1: nothing planned yet
            m_contactList = contactList;
            setTime(m_contactList.firstContact()->activationTime);
            return true;
        }
        const auto activeNew = contactList.getActiveContact(m_currentTrajectoryTime);
        const auto active = m_contactList.getActiveContact(m_currentTrajectoryTime);
        if (active != m_contactList.cend() && activeNew != contactList.cend()) {   // case 2: in contact, the same contact must be active
            double dp = 0.0;
            for (int c = 0; c < 3; c++) dp = std::max(dp, std::abs(active->pose.p[c] - activeNew->pose.p[c]));
            const Vector3d dr = (active->pose.R.inverse() * activeNew->pose.R).Log();
            if (dp > m_positionTolerance || std::sqrt(dr[0] * dr[0] + dr[1] * dr[1] + dr[2] * dr[2]) > m_orientationTolerance) {
                log()->error("{} The pose of the contact in the new contact list is different from the pose of the contact in the original "
                             "contact list.", logPrefix);
                return false;
            }
            m_contactList = contactList;
            return true;
        }
        if (active == m_contactList.cend() && activeNew != contactList.cend()) {
            log()->error("{} The new contact list contains an active contact at the current time instant. The original contact list does "
                         "not contain an active contact at the current time instant.", logPrefix);
            return false;
        }
        if (active != m_contactList.cend() && activeNew == contactList.cend()) {
            log()->error("{} The new contact list does not contain an active contact at the current time instant. The original contact "
                         "list contains an active contact at the current time instant.", logPrefix);
            return false;
        }
        if (contactList.getNextContact(m_currentTrajectoryTime) == contactList.cend()) {   // case 3: in swing, a landing spot must exist
            log()->error("{} The new contact list does not contain a contact after the current time instant.", logPrefix);
            return false;
        }
        m_contactList = contactList;
        return true;
    }

    void setTime(const std::chrono::nanoseconds& time) { m_currentTrajectoryTime = time; }
    const SwingFootPlannerState& getOutput() const final { return m_state; }
    bool isOutputValid() const final { return m_isOutputValid; }

    /** SwingFootPlanner.cpp:484-558 */
    bool advance() final
    {
        constexpr auto logPrefix = "[SwingFootPlanner::advance]";
        m_isOutputValid = false;
        m_state.time = m_currentTrajectoryTime;
        if (m_contactList.size() == 0) {
            log()->error("{} Empty contact list. Please call SwingFootPlanner::setContactList()", logPrefix);
            return false;
        }
        if (m_contactList.lastContact()->deactivationTime <= m_currentTrajectoryTime) {
            log()->error("{} The current time is greater than the deactivation time of the last contact. Please call "
                         "SwingFootPlanner::setContactList()", logPrefix);
            return false;
        }
        const auto active = m_contactList.getActiveContact(m_currentTrajectoryTime);
        if (active != m_contactList.cend()) {
            m_state.transform = active->pose, m_state.isInContact = true;
            m_state.mixedVelocity = SE3d::Tangent::Zero(), m_state.mixedAcceleration = SE3d::Tangent::Zero();
            m_lastValidContact = *active;
            m_isOutputValid = true;
            m_currentTrajectoryTime += m_dT;
            return true;
        }
        if (!createSE3Traj()) {
            log()->error("{} Unable to create the new SE(3) trajectory.", logPrefix);
            return false;
        }
        if (!evaluateSE3Traj(m_state)) {
            log()->error("{} Unable to update the SE(3) trajectory.", logPrefix);
            return false;
        }
        m_state.isInContact = false;
        m_isOutputValid = true;
        m_currentTrajectoryTime += m_dT;
        return true;
    }

    // ---- hook of the batched back-end (no counterpart in the reference)
    struct DeviceSwing {
        std::vector<std::chrono::nanoseconds> times;   /**< knot times of the 3D origin spline: take-off, apex, landing */
        std::vector<Vector3d> positions;               /**< its knots */
        Vector3d initialVelocity{0, 0, 0}, initialAcceleration{0, 0, 0}, finalVelocity{0, 0, 0}, finalAcceleration{0, 0, 0};
        SO3d landingRotation;                          /**< target of the minimum-jerk rotation */
        std::chrono::nanoseconds duration{0};          /**< of the swing */
        bool minJerk{false};                           /**< quintic (true) or cubic spline */
    };
    /** The swing that starts when `takeOff` (a contact of the list in force) ends, as ONE 3D spline + a landing rotation for
     *  IKLinearTask::setSetPointSpline / setSetPointMinimumJerk (see the class comment).  false: `takeOff` is the last contact. */
    bool planOnDevice(const Contacts::PlannedContact& takeOff, DeviceSwing& out) const
    {
        const auto next = m_contactList.getNextContact(takeOff.activationTime);
        if (next == m_contactList.cend()) return false;
        // the first swing sample (time = deactivationTime) re-plans from the output of the sample before it, which was still in
        // contact: the curve starts one period before the contact ends (createSE3Traj: previousTime), the apex time does not move
        const std::chrono::nanoseconds t0 = takeOff.deactivationTime - m_dT, t1 = next->activationTime;
        const auto ta = takeOff.deactivationTime + std::chrono::duration_cast<std::chrono::nanoseconds>(m_footApexTime * (t1 - takeOff.deactivationTime));
        std::unique_ptr<Spline1> planar = m_minJerk ? std::unique_ptr<Spline1>(new Math::QuinticSpline<VectorXd>()) : std::unique_ptr<Spline1>(new Math::CubicSpline<VectorXd>());
        planar->setInitialConditions(VectorXd{0.0, 0.0}, VectorXd{0.0, 0.0});
        planar->setFinalConditions(VectorXd{0.0, 0.0}, VectorXd{0.0, 0.0});
        planar->setKnots({VectorXd{takeOff.pose.p[0], takeOff.pose.p[1]}, VectorXd{next->pose.p[0], next->pose.p[1]}}, {t0, t1});
        VectorXd xyApex;
        if (!planar->evaluatePoint(ta, xyApex)) return false;
        out.times = {t0, ta, t1};
        out.positions = {takeOff.pose.p, {xyApex[0], xyApex[1], std::max(takeOff.pose.p[2], next->pose.p[2]) + m_stepHeight}, next->pose.p};
        out.initialVelocity = {0.0, 0.0, m_footTakeOffVelocity}, out.initialAcceleration = {0.0, 0.0, m_footTakeOffAcceleration};
        out.finalVelocity = {0.0, 0.0, m_footLandingVelocity}, out.finalAcceleration = {0.0, 0.0, m_footLandingAcceleration};
        out.landingRotation = next->pose.R, out.duration = t1 - t0, out.minJerk = m_minJerk;
        return true;
    }

private:
    /** SwingFootPlanner.cpp:310-432, from the previous output */
    bool createSE3Traj()
    {
        using namespace std::chrono_literals;
        constexpr auto logPrefix = "[SwingFootPlanner::createSE3Traj]";
        const auto previousTime = m_currentTrajectoryTime - m_dT;
        if (previousTime < 0s) {
            log()->error("{} Invalid previous time.", logPrefix);
            return false;
        }
        const auto nextContact = m_contactList.getNextContact(previousTime);
        if (nextContact == m_contactList.cend()) {
            log()->error("{} Invalid next contact.", logPrefix);
            return false;
        }
        m_startingTimeOfCurrentSO3Traj = previousTime;
        SO3Tangentd w0, a0;
        w0.w = m_state.mixedVelocity.angular, a0.w = m_state.mixedAcceleration.angular;
        m_SO3Planner.setInitialConditions(w0, a0);
        m_SO3Planner.setFinalConditions(SO3Tangentd::Zero(), SO3Tangentd::Zero());
        if (!m_SO3Planner.setRotations(m_state.transform.R, nextContact->pose.R, nextContact->activationTime - previousTime)) {
            log()->error("{} Unable to set the initial and final rotations for the SO(3) planner.", logPrefix);
            return false;
        }
        const Vector3d &p = m_state.transform.p, &v = m_state.mixedVelocity.linear, &a = m_state.mixedAcceleration.linear;
        if (!m_planarPlanner->setInitialConditions(VectorXd{v[0], v[1]}, VectorXd{a[0], a[1]})
            || !m_planarPlanner->setFinalConditions(VectorXd{0.0, 0.0}, VectorXd{0.0, 0.0})
            || !m_planarPlanner->setKnots({VectorXd{p[0], p[1]}, VectorXd{nextContact->pose.p[0], nextContact->pose.p[1]}},
                                          {previousTime, nextContact->activationTime})) {
            log()->error("{} Unable to set the knots for the planar planner.", logPrefix);
            return false;
        }
        const double viaZ = std::max(m_lastValidContact.pose.p[2], nextContact->pose.p[2]) + m_stepHeight;
        const std::chrono::nanoseconds swingDuration = nextContact->activationTime - m_lastValidContact.deactivationTime;
        const std::chrono::nanoseconds viaTime
            = std::chrono::duration_cast<std::chrono::nanoseconds>(m_footApexTime * swingDuration + m_lastValidContact.deactivationTime);
        if (!m_heightPlanner->setInitialConditions(VectorXd{v[2]}, VectorXd{a[2]})
            || !m_heightPlanner->setFinalConditions(VectorXd{m_footLandingVelocity}, VectorXd{m_footLandingAcceleration})) {
            log()->error("{} Unable to set the boundary conditions for the height planner.", logPrefix);
            return false;
        }
        const bool ok = previousTime < viaTime
                            ? m_heightPlanner->setKnots({VectorXd{p[2]}, VectorXd{viaZ}, VectorXd{nextContact->pose.p[2]}},
                                                        {previousTime, viaTime, nextContact->activationTime})
                            : m_heightPlanner->setKnots({VectorXd{p[2]}, VectorXd{nextContact->pose.p[2]}}, {previousTime, nextContact->activationTime});
        if (!ok) log()->error("{} Unable to set the knots for the knots for the height planner.", logPrefix);
        return ok;
    }
    /** SwingFootPlanner.cpp:434-482 */
    bool evaluateSE3Traj(SwingFootPlannerState& state)
    {
        constexpr auto logPrefix = "[SwingFootPlanner::updateSE3Traj]";
        SO3d rotation;
        SO3Tangentd w, dw;
        if (!m_SO3Planner.evaluatePoint(m_currentTrajectoryTime - m_startingTimeOfCurrentSO3Traj, rotation, w, dw)) {
            log()->error("{} Unable to get the SO(3) trajectory.", logPrefix);
            return false;
        }
        VectorXd xy, vxy, axy, z, vz, az;
        if (!m_planarPlanner->evaluatePoint(m_currentTrajectoryTime, xy, vxy, axy)) {
            log()->error("{} Unable to get the planar trajectory.", logPrefix);
            return false;
        }
        if (!m_heightPlanner->evaluatePoint(m_currentTrajectoryTime, z, vz, az)) {
            log()->error("{} Unable to the height trajectory.", logPrefix);
            return false;
        }
        state.transform = SE3d({xy[0], xy[1], z[0]}, rotation);
        state.mixedVelocity.linear = {vxy[0], vxy[1], vz[0]}, state.mixedVelocity.angular = w.w;
        state.mixedAcceleration.linear = {axy[0], axy[1], az[0]}, state.mixedAcceleration.angular = dw.w;
        return true;
    }

    SwingFootPlannerState m_state;
    std::chrono::nanoseconds m_dT{0}, m_currentTrajectoryTime{0}, m_startingTimeOfCurrentSO3Traj{0};
    Contacts::ContactList m_contactList;
    Contacts::PlannedContact m_lastValidContact;
    SO3PlannerInertial m_SO3Planner;
    std::unique_ptr<Spline1> m_planarPlanner, m_heightPlanner;
    double m_stepHeight{0.0}, m_footApexTime{0.5}, m_positionTolerance{1e-6}, m_orientationTolerance{1e-6};
    double m_footLandingVelocity{0.0}, m_footLandingAcceleration{0.0}, m_footTakeOffVelocity{0.0}, m_footTakeOffAcceleration{0.0};
    bool m_isOutputValid{false}, m_minJerk{false};
};

} // namespace Planners
} // namespace BipedalLocomotion

#endif // BLF_B200_PLANNERS_H
