/**
 * Math.h -- the spline classes of src/Math that produce the set-points of the IK tick.
 *
 * Same class / method names, argument meaning and bool + log()->error convention as
 *   src/Math/include/BipedalLocomotion/Math/Spline.h          (TrajectoryPoint :31-38, Spline :46-268, methods :271-762)
 *   src/Math/include/BipedalLocomotion/Math/QuinticSpline.h   (:26-415)
 * T is Vector3d or VectorXd (Core.h) instead of an Eigen type.  This is the host-side, one-trajectory class; for a batch
 * of robots the same spline is evaluated on the GPU inside the tick (IKLinearTask::setSetPointSpline, which takes the
 * knots of objects of this class; blfik_set_setpoint_spline).
 *
 * The interior-knot velocities and accelerations are the ones that make the third and fourth derivative continuous
 * (QuinticSpline.h:297-415).  The reference assembles the sparse block-tridiagonal system and calls Eigen::SparseQR; here the
 * 2 x 2 blocks are eliminated directly, the same elimination the device kernel runs per robot and coordinate.
 */
#ifndef BLF_B200_MATH_H
#define BLF_B200_MATH_H

#include <algorithm>
#include <chrono>
#include <vector>

#include <BipedalLocomotion/Core.h>
#include <BipedalLocomotion/System.h>

namespace BipedalLocomotion
{
namespace Math
{

namespace detail
{
inline void resizeLike(Vector3d&, std::size_t) {}
inline void resizeLike(VectorXd& v, std::size_t n) { v.assign(n, 0.0); }
inline double seconds(const std::chrono::nanoseconds& t) { return std::chrono::duration<double>(t).count(); }
} // namespace detail

/** Spline.h:31-38 */
template <typename T> struct TrajectoryPoint {
    T position{};
    T velocity{};
    T acceleration{};
    std::chrono::nanoseconds time{0};
};

/** Spline.h:46-268 */
template <typename T> class Spline : public System::Source<TrajectoryPoint<T>>
{
public:
    struct DerivativeBoundaryConditions {
        T velocity{};
        T acceleration{};
    };
    virtual ~Spline() = default;

    bool setAdvanceTimeStep(const std::chrono::nanoseconds& dt)
    {
        if (dt <= std::chrono::nanoseconds::zero()) {
            log()->error("[Spline::setAdvanceTimeStep] The time step has to be greater than zero.");
            return false;
        }
        m_dT = dt;
        return true;
    }

    bool setKnots(const std::vector<T>& position, const std::vector<std::chrono::nanoseconds>& time)
    {
        constexpr auto logPrefix = "[Spline::setKnots]";
        if (position.size() < 2) {
            log()->error("{} The number of knots has to be at least 2.", logPrefix);
            return false;
        }
        if (time.size() != position.size()) {
            log()->error("{} The number of points has to be equal to the length of the time vector.", logPrefix);
            return false;
        }
        const std::size_t dim = position.front().size();
        for (const T& k : position)
            if (k.size() != dim) {
                log()->error("{} The size of the knot is different from the one expected. Expected: {} Retrieved {}.", logPrefix, dim, k.size());
                return false;
            }
        m_knots.resize(position.size());
        for (std::size_t i = 0; i < position.size(); i++) {
            m_knots[i].position = position[i], m_knots[i].time = time[i];
            detail::resizeLike(m_knots[i].velocity, dim), detail::resizeLike(m_knots[i].acceleration, dim);
        }
        m_current = m_knots.front();
        m_advanceCurrentTime = m_knots.front().time;
        m_areCoefficientsComputed = false;
        return true;
    }

    bool setInitialConditions(const DerivativeBoundaryConditions& c) { return setConditions(c, m_initialConditions, m_hasInitial, "[Spline::setInitialConditions]"); }
    bool setFinalConditions(const DerivativeBoundaryConditions& c) { return setConditions(c, m_finalConditions, m_hasFinal, "[Spline::setFinalConditions]"); }
    bool setInitialConditions(const T& velocity, const T& acceleration) { return setInitialConditions(DerivativeBoundaryConditions{velocity, acceleration}); }
    bool setFinalConditions(const T& velocity, const T& acceleration) { return setFinalConditions(DerivativeBoundaryConditions{velocity, acceleration}); }

    bool evaluatePoint(const std::chrono::nanoseconds& t, T& position, T& velocity, T& acceleration)
    {
        if (!computeCoefficients()) {
            log()->error("[Spline::evaluatePoint] Unable to compute the coefficients of the spline.");
            return false;
        }
        // before the first knot and from the last knot on the end states hold (:607-625)
        std::size_t seg = 0;
        double tl = 0.0;
        if (t >= m_knots.back().time) {
            seg = m_knots.size() - 2, tl = detail::seconds(m_knots.back().time - m_knots[seg].time);
        } else if (t >= m_knots.front().time) {
            const auto next = std::upper_bound(m_knots.begin(), m_knots.end(), t, [](const std::chrono::nanoseconds& v, const TrajectoryPoint<T>& k) { return v < k.time; });
            seg = static_cast<std::size_t>(next - m_knots.begin()) - 1;
            tl = detail::seconds(t - m_knots[seg].time);
        }
        const std::size_t dim = m_knots.front().position.size();
        detail::resizeLike(position, dim), detail::resizeLike(velocity, dim), detail::resizeLike(acceleration, dim);
        const std::vector<T>& c = m_coefficients[seg];
        for (std::size_t k = 0; k < dim; k++) {
            position[k] = c[0][k] + tl * (c[1][k] + tl * (c[2][k] + tl * (c[3][k] + tl * (c[4][k] + tl * c[5][k]))));
            velocity[k] = c[1][k] + tl * (2.0 * c[2][k] + tl * (3.0 * c[3][k] + tl * (4.0 * c[4][k] + tl * 5.0 * c[5][k])));
            acceleration[k] = 2.0 * c[2][k] + tl * (6.0 * c[3][k] + tl * (12.0 * c[4][k] + tl * 20.0 * c[5][k]));
        }
        return true;
    }
    bool evaluatePoint(const std::chrono::nanoseconds& t, T& position) { return evaluatePoint(t, position, m_dummy[0], m_dummy[1]); }
    bool evaluatePoint(const std::chrono::nanoseconds& t, T& position, T& velocity) { return evaluatePoint(t, position, velocity, m_dummy[1]); }
    bool evaluatePoint(const std::chrono::nanoseconds& t, TrajectoryPoint<T>& state)
    {
        state.time = t;
        return evaluatePoint(t, state.position, state.velocity, state.acceleration);
    }
    bool evaluateOrderedPoints(const std::vector<std::chrono::nanoseconds>& t, std::vector<T>& position, std::vector<T>& velocity, std::vector<T>& acceleration)
    {
        position.resize(t.size()), velocity.resize(t.size()), acceleration.resize(t.size());
        for (std::size_t i = 0; i < t.size(); i++)
            if (!evaluatePoint(t[i], position[i], velocity[i], acceleration[i])) {
                log()->error("[Spline::evaluateOrderedPoints] Unable to compute the coefficients of the spline.");
                return false;
            }
        return true;
    }

    bool isOutputValid() const final { return m_isOutputValid; }
    const TrajectoryPoint<T>& getOutput() const final { return m_current; }
    bool advance() final
    {
        m_isOutputValid = false;
        if (m_dT <= std::chrono::nanoseconds::zero() || !evaluatePoint(m_advanceCurrentTime, m_current)) {
            log()->error("[Spline::advance] Unable to evaluate the spline at the current time.");
            return false;
        }
        m_advanceCurrentTime = std::min(m_dT + m_advanceCurrentTime, m_knots.back().time);
        m_isOutputValid = true;
        return true;
    }

    // ---- hooks of the batched back-end (no counterpart in the reference)
    const std::vector<TrajectoryPoint<T>>& knots() const { return m_knots; }
    bool boundaryConditions(DerivativeBoundaryConditions& initial, DerivativeBoundaryConditions& final_) const
    {
        if (!m_hasInitial || !m_hasFinal) return false;
        initial = m_initialConditions, final_ = m_finalConditions;
        return true;
    }

protected:
    std::vector<TrajectoryPoint<T>> m_knots;
    std::vector<std::vector<T>> m_coefficients;   // per interval: c_0 .. c_5 in the local time of the interval
    /** velocities and accelerations of the interior knots (only called with more than two knots) */
    virtual void computeIntermediateQuantities() = 0;
    /** coefficients of interval i from the states of its two knots */
    virtual void updatePolynomialCoefficients(std::size_t i) = 0;

private:
    bool setConditions(const DerivativeBoundaryConditions& c, DerivativeBoundaryConditions& dst, bool& has, const char* prefix)
    {
        if (c.velocity.size() != c.acceleration.size()) {
            log()->error("{} The size of the velocity is different from the one of the acceleration. Expected: {} Retrieved: {}.", prefix,
                         c.velocity.size(), c.acceleration.size());
            return false;
        }
        m_areCoefficientsComputed = false;
        dst = c, has = true;
        return true;
    }
    bool computeCoefficients()
    {
        constexpr auto errorPrefix = "[Spline::computeCoefficients]";
        if (m_areCoefficientsComputed) return true;
        if (m_knots.empty()) {
            log()->error("{} Please set the knots before computing the coefficients.", errorPrefix);
            return false;
        }
        const std::size_t dim = m_knots.front().position.size();
        if (!m_hasInitial || !m_hasFinal || m_initialConditions.velocity.size() != dim || m_initialConditions.acceleration.size() != dim
            || m_finalConditions.velocity.size() != dim || m_finalConditions.acceleration.size() != dim) {
            log()->error("[Spline::setBoundaryVelocitiesAndAcceleration] The size of a boundary velocity or acceleration is different from the "
                         "expected one. Expected: {}.", dim);
            log()->error("{} Unable to set the boundary conditions related to the velocity and acceleration.", errorPrefix);
            return false;
        }
        m_knots.front().velocity = m_initialConditions.velocity, m_knots.front().acceleration = m_initialConditions.acceleration;
        m_knots.back().velocity = m_finalConditions.velocity, m_knots.back().acceleration = m_finalConditions.acceleration;
        for (std::size_t i = 0; i + 1 < m_knots.size(); i++)
            if (m_knots[i + 1].time == m_knots[i].time) {
                log()->error("[Splines::computePhasesDuration] Two consecutive points have the same time coordinate.");
                log()->error("{} Unable to compute the phase duration.", errorPrefix);
                return false;
            }
        if (m_knots.size() > 2) computeIntermediateQuantities();
        m_coefficients.resize(m_knots.size() - 1);
        for (std::size_t i = 0; i + 1 < m_knots.size(); i++) updatePolynomialCoefficients(i);
        m_areCoefficientsComputed = true;
        return true;
    }

    T m_dummy[2];
    TrajectoryPoint<T> m_current;
    bool m_isOutputValid{false}, m_areCoefficientsComputed{false}, m_hasInitial{false}, m_hasFinal{false};
    std::chrono::nanoseconds m_advanceCurrentTime{0}, m_dT{0};
    DerivativeBoundaryConditions m_initialConditions, m_finalConditions;
};

/** CubicSpline.h:26-330: the knot velocities make the acceleration continuous; the boundary accelerations handed to
 *  setInitialConditions / setFinalConditions are accepted and not used, as in the reference. */
template <typename T> class CubicSpline : public Spline<T>
{
public:
    static constexpr bool isCubic = true;

private:
    void updatePolynomialCoefficients(std::size_t i) final
    {
        const auto& k0 = this->m_knots[i];
        const auto& k1 = this->m_knots[i + 1];
        const double d = detail::seconds(k1.time - k0.time);
        const std::size_t dim = k0.position.size();
        std::vector<T>& c = this->m_coefficients[i];
        c.assign(6, k0.position);
        for (std::size_t k = 0; k < dim; k++) {
            const double dx = k1.position[k] - k0.position[k], v0 = k0.velocity[k], v1 = k1.velocity[k];
            c[1][k] = v0;
            c[2][k] = (3.0 * dx - (2.0 * v0 + v1) * d) / (d * d);
            c[3][k] = (-2.0 * dx + (v0 + v1) * d) / (d * d * d);
            c[4][k] = 0.0, c[5][k] = 0.0;
        }
    }
    void computeIntermediateQuantities() final
    {
        // 2/Tp v_{i-1} + 4 (1/Tp + 1/Tn) v_i + 2/Tn v_{i+1} = 6 (dp/Tp^2 + dn/Tn^2) at every interior knot i
        auto& kn = this->m_knots;
        const std::size_t m = kn.size() - 2, dim = kn.front().position.size();
        std::vector<double> di(m), up(m), r(m * dim);
        for (std::size_t i = 0; i < m; i++) {
            const double Tp = detail::seconds(kn[i + 1].time - kn[i].time), Tn = detail::seconds(kn[i + 2].time - kn[i + 1].time);
            const double lo = 2.0 / Tp;
            di[i] = 4.0 / Tp + 4.0 / Tn, up[i] = 2.0 / Tn;
            const double w = i > 0 ? lo / di[i - 1] : 0.0;
            if (i > 0) di[i] -= w * up[i - 1];
            for (std::size_t k = 0; k < dim; k++) {
                double v = 6.0 * ((kn[i + 1].position[k] - kn[i].position[k]) / (Tp * Tp) + (kn[i + 2].position[k] - kn[i + 1].position[k]) / (Tn * Tn));
                if (i == 0) v -= lo * kn[0].velocity[k];
                if (i == m - 1) v -= up[i] * kn[m + 1].velocity[k];
                if (i > 0) v -= w * r[(i - 1) * dim + k];
                r[i * dim + k] = v;
            }
        }
        for (std::size_t i = m; i-- > 0;)
            for (std::size_t k = 0; k < dim; k++) {
                double v = r[i * dim + k];
                if (i + 1 < m) v -= up[i] * kn[i + 2].velocity[k];
                kn[i + 1].velocity[k] = v / di[i];
            }
    }
};

/** QuinticSpline.h:26-415 */
template <typename T> class QuinticSpline : public Spline<T>
{
public:
    static constexpr bool isCubic = false;

private:
    void updatePolynomialCoefficients(std::size_t i) final
    {
        const auto& k0 = this->m_knots[i];
        const auto& k1 = this->m_knots[i + 1];
        const double d = detail::seconds(k1.time - k0.time);
        const std::size_t dim = k0.position.size();
        std::vector<T>& c = this->m_coefficients[i];
        c.assign(6, k0.position);
        for (std::size_t k = 0; k < dim; k++) {
            const double dx = k1.position[k] - k0.position[k], v0 = k0.velocity[k], v1 = k1.velocity[k], a0 = k0.acceleration[k], a1 = k1.acceleration[k];
            c[1][k] = v0, c[2][k] = 0.5 * a0;
            c[3][k] = (20.0 * dx - (8.0 * v1 + 12.0 * v0) * d - (3.0 * a0 - a1) * d * d) / (2.0 * d * d * d);
            c[4][k] = (-30.0 * dx + (14.0 * v1 + 16.0 * v0) * d + (3.0 * a0 - 2.0 * a1) * d * d) / (2.0 * d * d * d * d);
            c[5][k] = (12.0 * dx - 6.0 * (v1 + v0) * d - (a0 - a1) * d * d) / (2.0 * d * d * d * d * d);
        }
    }

    void computeIntermediateQuantities() final
    {
        // rows of knot i: jerk and snap of interval i-1 at its end = those of interval i at its start, unknowns (v, a) of the
        // knots i-1, i, i+1 through the 2 x 2 blocks L, D, U; forward elimination, back substitution, every coordinate at once
        auto& kn = this->m_knots;
        const std::size_t m = kn.size() - 2, dim = kn.front().position.size();
        struct B2 { double a, b, c, d; };
        auto mul = [](const B2& x, const B2& y) { return B2{x.a * y.a + x.b * y.c, x.a * y.b + x.b * y.d, x.c * y.a + x.d * y.c, x.c * y.b + x.d * y.d}; };
        auto inv = [](const B2& x) { const double det = x.a * x.d - x.b * x.c; return B2{x.d / det, -x.b / det, -x.c / det, x.a / det}; };
        std::vector<B2> Dinv(m), U(m), M(m);
        std::vector<double> r0(m * dim), r1(m * dim);
        for (std::size_t i = 0; i < m; i++) {
            const double Tp = detail::seconds(kn[i + 1].time - kn[i].time), Tn = detail::seconds(kn[i + 2].time - kn[i + 1].time);
            const B2 L{-24.0 / (Tp * Tp), -3.0 / Tp, -168.0 / (Tp * Tp * Tp), -24.0 / (Tp * Tp)};
            B2 D{-36.0 / (Tp * Tp) + 36.0 / (Tn * Tn), 9.0 / Tp + 9.0 / Tn, -192.0 / (Tp * Tp * Tp) - 192.0 / (Tn * Tn * Tn), 36.0 / (Tp * Tp) - 36.0 / (Tn * Tn)};
            U[i] = B2{24.0 / (Tn * Tn), -3.0 / Tn, -168.0 / (Tn * Tn * Tn), 24.0 / (Tn * Tn)};
            if (i > 0) {
                M[i] = mul(L, Dinv[i - 1]);
                const B2 MU = mul(M[i], U[i - 1]);
                D = B2{D.a - MU.a, D.b - MU.b, D.c - MU.c, D.d - MU.d};
            }
            Dinv[i] = inv(D);
            for (std::size_t k = 0; k < dim; k++) {
                const double dp = kn[i + 1].position[k] - kn[i].position[k], dn = kn[i + 2].position[k] - kn[i + 1].position[k];
                double a = 60.0 * (dn / (Tn * Tn * Tn) - dp / (Tp * Tp * Tp)), b = -360.0 * (dp / (Tp * Tp * Tp * Tp) + dn / (Tn * Tn * Tn * Tn));
                if (i == 0) a -= L.a * kn[0].velocity[k] + L.b * kn[0].acceleration[k], b -= L.c * kn[0].velocity[k] + L.d * kn[0].acceleration[k];
                if (i == m - 1) {
                    a -= U[i].a * kn[m + 1].velocity[k] + U[i].b * kn[m + 1].acceleration[k];
                    b -= U[i].c * kn[m + 1].velocity[k] + U[i].d * kn[m + 1].acceleration[k];
                }
                if (i > 0) {
                    const double pa = r0[(i - 1) * dim + k], pb = r1[(i - 1) * dim + k];
                    a -= M[i].a * pa + M[i].b * pb, b -= M[i].c * pa + M[i].d * pb;
                }
                r0[i * dim + k] = a, r1[i * dim + k] = b;
            }
        }
        for (std::size_t i = m; i-- > 0;)
            for (std::size_t k = 0; k < dim; k++) {
                double a = r0[i * dim + k], b = r1[i * dim + k];
                if (i + 1 < m) {
                    a -= U[i].a * kn[i + 2].velocity[k] + U[i].b * kn[i + 2].acceleration[k];
                    b -= U[i].c * kn[i + 2].velocity[k] + U[i].d * kn[i + 2].acceleration[k];
                }
                kn[i + 1].velocity[k] = Dinv[i].a * a + Dinv[i].b * b;
                kn[i + 1].acceleration[k] = Dinv[i].c * a + Dinv[i].d * b;
            }
    }
};

} // namespace Math
} // namespace BipedalLocomotion

#endif // BLF_B200_MATH_H
