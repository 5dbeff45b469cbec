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

#include <BipedalLocomotion/Core.h>
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

} // namespace Planners
} // namespace BipedalLocomotion

#endif // BLF_B200_PLANNERS_H
