/**
 * ContinuousDynamicalSystem.h -- FloatingBaseSystemVelocityKinematics and ForwardEuler with a batch dimension.
 *   reference: src/ContinuousDynamicalSystem/include/BipedalLocomotion/ContinuousDynamicalSystem/
 *              FloatingBaseSystemVelocityKinematics.h:34-128 (.cpp:14-84), ForwardEuler.h:49-124,
 *              FixedStepIntegrator.h:38-140, Integrator.h:41-129
 * State = (p in R3, R in SO(3), s in R^n); Input = (mixed twist in R6, sdot in R^n);
 * one Euler step: p += v dt, R <- exp(w dt) R (manif tangent + group = left-plus), s += sdot dt.
 * The step itself runs on the GPU (blfik_integrate_state); there is no CPU integrator for the robot states here.
 *
 * Round 2 (first slice): FloatingBaseSystemAccelerationKinematics (.h:30-150, .cpp:14-98) with the same batch
 * dimension, stepped on the GPU by blfik_integrate_acceleration_state; LinearTimeInvariantSystem (.h:33-118, .cpp:14-115) and
 * RK4 (RK4.h:37-175) for the R^n systems the reference integrates with it (a handful of doubles: host arithmetic, like
 * FirstOrderSmoother below).
 */
#ifndef BLF_B200_CONTINUOUS_DYNAMICAL_SYSTEM_H
#define BLF_B200_CONTINUOUS_DYNAMICAL_SYSTEM_H

#include <BipedalLocomotion/Core.h>
#include <BipedalLocomotion/System.h>
#include <blfik.h>

#include <map>

#include <tuple>

namespace BipedalLocomotion
{
namespace ContinuousDynamicalSystem
{

class FloatingBaseSystemVelocityKinematics
{
public:
    using State = std::tuple<Vector3d, SO3d, VectorXd>;
    using StateDerivative = std::tuple<Vector3d, SO3Tangentd, VectorXd>;
    using Input = std::tuple<std::array<double, 6>, VectorXd>;

    /** no mandatory parameter; `rho` is accepted and ignored with an info message (.cpp:14-31) */
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler)
    {
        auto ptr = handler.lock();
        double rho;
        if (ptr && ptr->getParameter("rho", rho))
            log()->info("[FloatingBaseSystemVelocityKinematics::initialize] The rho parameter is not used by this system.");
        return true;
    }
    /** number of robots integrated at once (default 1) */
    void setBatchSize(std::size_t N) { m_N = N; }
    std::size_t getBatchSize() const { return m_N; }

    bool setState(const State& state, std::ptrdiff_t instance = -1)
    {
        const VectorXd& s = std::get<2>(state);
        if (m_n == 0) m_n = s.size();
        if (s.size() != m_n) {
            log()->error("[FloatingBaseSystemVelocityKinematics::setState] Wrong size of the joint vector.");
            return false;
        }
        resize();
        const auto q = std::get<1>(state).quat().coeffs();
        for (std::size_t i = 0; i < m_N; i++) {
            if (instance >= 0 && static_cast<std::size_t>(instance) != i) continue;
            for (int c = 0; c < 3; c++) m_p[3 * i + c] = std::get<0>(state)[c];
            for (int c = 0; c < 4; c++) m_q[4 * i + c] = q[c];
            for (std::size_t j = 0; j < m_n; j++) m_s[m_n * i + j] = s[j];
        }
        return true;
    }
    bool setStateBatch(std::size_t n, const double* p, const double* q, const double* s)
    {
        m_n = n;
        resize();
        m_p.assign(p, p + 3 * m_N), m_q.assign(q, q + 4 * m_N), m_s.assign(s, s + n * m_N);
        return true;
    }
    State getState(std::size_t instance = 0) const
    {
        State st;
        if (m_p.empty()) return st;
        for (int c = 0; c < 3; c++) std::get<0>(st)[c] = m_p[3 * instance + c];
        std::get<1>(st) = SO3d(Quaterniond{m_q[4 * instance], m_q[4 * instance + 1], m_q[4 * instance + 2], m_q[4 * instance + 3]});
        std::get<2>(st).assign(m_s.begin() + m_n * instance, m_s.begin() + m_n * (instance + 1));
        return st;
    }
    bool setControlInput(const Input& input, std::ptrdiff_t instance = -1)
    {
        const VectorXd& ds = std::get<1>(input);
        if (m_n == 0) m_n = ds.size();
        resize();
        if (ds.size() != m_n) {  // .cpp:52-56
            log()->error("[FloatingBaseSystemVelocityKinematics::dynamics] Wrong size of the vectors.");
            return false;
        }
        const std::size_t nv = m_n + 6;
        for (std::size_t i = 0; i < m_N; i++) {
            if (instance >= 0 && static_cast<std::size_t>(instance) != i) continue;
            for (int c = 0; c < 6; c++) m_nu[nv * i + c] = std::get<0>(input)[c];
            for (std::size_t j = 0; j < m_n; j++) m_nu[nv * i + 6 + j] = ds[j];
        }
        m_inputSet = true;
        return true;
    }
    /** [N][6+n] robot-major (the layout QPInverseKinematics::getBatchOutput returns) */
    bool setControlInputBatch(const double* nu)
    {
        if (m_n == 0) return false;
        resize();
        m_nu.assign(nu, nu + (m_n + 6) * m_N);
        m_inputSet = true;
        return true;
    }
    /** dp = twist.head3, omega = twist.tail3, ds = sdot (.cpp:33-67) */
    bool dynamics(const std::chrono::nanoseconds&, StateDerivative& d, std::size_t instance = 0) const
    {
        if (!m_inputSet) return false;
        const std::size_t nv = m_n + 6;
        for (int c = 0; c < 3; c++) std::get<0>(d)[c] = m_nu[nv * instance + c], std::get<1>(d).w[c] = m_nu[nv * instance + 3 + c];
        std::get<2>(d).assign(m_nu.begin() + nv * instance + 6, m_nu.begin() + nv * (instance + 1));
        return true;
    }
    std::size_t dofs() const { return m_n; }
    std::vector<double>&basePositions() { return m_p; }
    std::vector<double>&baseQuaternions() { return m_q; }
    std::vector<double>&jointPositions() { return m_s; }
    const std::vector<double>& controlInput() const { return m_nu; }
    bool hasInput() const { return m_inputSet; }
    /** `nsteps` Euler steps of `dt` with the input held (ForwardEuler.h:102-121), on `device` */
    bool forwardEulerSteps(double dt, int nsteps, int device)
    {
        return blfik_integrate_state(static_cast<int32_t>(m_n), static_cast<int64_t>(m_N), m_p.data(), m_q.data(), m_s.data(), m_nu.data(), dt,
                                     nsteps, BLFIK_MEM_HOST, device) == BLFIK_OK;
    }

private:
    void resize()
    {
        if (m_p.size() != 3 * m_N) {
            m_p.assign(3 * m_N, 0.0), m_q.assign(4 * m_N, 0.0), m_s.assign(m_n * m_N, 0.0);
            for (std::size_t i = 0; i < m_N; i++) m_q[4 * i + 3] = 1.0;
        }
        if (m_nu.size() != (m_n + 6) * m_N) m_nu.assign((m_n + 6) * m_N, 0.0);
    }
    std::size_t m_N{1}, m_n{0};
    std::vector<double> m_p, m_q, m_s, m_nu;
    bool m_inputSet{false};
};

/** deprecated alias kept by the reference (FloatingBaseSystemVelocityKinematics.h:126-128) */
using FloatingBaseSystemKinematics = FloatingBaseSystemVelocityKinematics;

template <class _DynamicalSystem> class ForwardEuler
{
public:
    using DynamicalSystem = _DynamicalSystem;
    using State = typename _DynamicalSystem::State;

    /** only once (Integrator.h:94-99) */
    bool setDynamicalSystem(std::shared_ptr<DynamicalSystem> dynamicalSystem)
    {
        constexpr auto errorPrefix = "[Integrator::setDynamicalSystem]";
        if (m_dynamicalSystem != nullptr) {
            log()->error("{} The dynamical system has been already set.", errorPrefix);
            return false;
        }
        if (dynamicalSystem == nullptr) {
            log()->error("{} The dynamical system passed to the function is corrupted.", errorPrefix);
            return false;
        }
        m_dynamicalSystem = dynamicalSystem;
        return true;
    }
    std::weak_ptr<DynamicalSystem> dynamicalSystem() const { return m_dynamicalSystem; }
    /** FixedStepIntegrator.h:73-85 */
    bool setIntegrationStep(const std::chrono::nanoseconds& dT)
    {
        if (dT <= std::chrono::nanoseconds::zero()) {
            log()->error("[FixedStepIntegrator::setIntegrationStep] The integration must be a strictly positive number.");
            return false;
        }
        m_dT = dT;
        return true;
    }
    std::chrono::nanoseconds getIntegrationStep() const { return m_dT; }
    void setDevice(int device) { m_device = device; }

    /** FixedStepIntegrator.h:94-140: (iterations - 1) steps of dT, then one step that lands on finalTime.
     *  The reference's size_t underflow for finalTime - initialTime < dT is rejected instead of reproduced. */
    bool integrate(const std::chrono::nanoseconds& initialTime, const std::chrono::nanoseconds& finalTime)
    {
        constexpr auto errorPrefix = "[FixedStepIntegrator::integrate]";
        if (m_dynamicalSystem == nullptr) {
            log()->error("{} Please set the dynamical system before call this function.", errorPrefix);
            return false;
        }
        if (initialTime > finalTime) {
            log()->error("{} The final time has to be greater than the initial one.", errorPrefix);
            return false;
        }
        if (m_dT <= std::chrono::nanoseconds::zero()) {
            log()->error("{} Please set the integration step.", errorPrefix);
            return false;
        }
        const std::size_t iterations = static_cast<std::size_t>((finalTime - initialTime) / m_dT);
        if (iterations < 1) {
            log()->error("{} The integration interval is shorter than the integration step.", errorPrefix);
            return false;
        }
        auto& sys = *m_dynamicalSystem;
        if (!sys.hasInput()) {
            log()->error("[ForwardEuler::oneStepIntegration] Unable to compute the system dynamics.");
            return false;
        }
        const double dT = std::chrono::duration<double>(m_dT).count();
        const auto lastStep = finalTime - (initialTime + m_dT * static_cast<long long>(iterations - 1));
        if (iterations > 1 && !step(sys, dT, static_cast<int>(iterations - 1))) return false;
        return step(sys, std::chrono::duration<double>(lastStep).count(), 1);
    }
    /** Integrator.h:117-122 */
    State getSolution(std::size_t instance = 0) const { return m_dynamicalSystem ? m_dynamicalSystem->getState(instance) : State(); }

private:
    bool step(DynamicalSystem& sys, double dt, int nsteps)
    {
        if (!sys.forwardEulerSteps(dt, nsteps, m_device)) {
            log()->error("[ForwardEuler::oneStepIntegration] {}", blfik_last_error());
            return false;
        }
        return true;
    }
    std::shared_ptr<DynamicalSystem> m_dynamicalSystem;
    std::chrono::nanoseconds m_dT{0};
    int m_device{0};
};

/** FloatingBaseSystemAccelerationKinematics.h:30-150, .cpp:14-98.  State = (p, R, s, dp, omega, ds); Input = (dtwist in R6
 *  mixed, dds); derivative = (dp, omega, ds, ddp, domega, dds).  Under ForwardEuler every component advances with the derivative
 *  evaluated at the OLD state (ForwardEuler.h:102-121): p += dp dt, R <- exp(omega dt) R, s += ds dt, then dp += ddp dt, ... */
class FloatingBaseSystemAccelerationKinematics
{
public:
    using State = std::tuple<Vector3d, SO3d, VectorXd, Vector3d, SO3Tangentd, VectorXd>;
    using StateDerivative = std::tuple<Vector3d, SO3Tangentd, VectorXd, Vector3d, SO3Tangentd, VectorXd>;
    using Input = std::tuple<std::array<double, 6>, VectorXd>;

    /** no parameter (.cpp:14-18) */
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler>) { return true; }
    void setBatchSize(std::size_t N) { m_N = N; }
    std::size_t getBatchSize() const { return m_N; }

    bool setState(const State& state, std::ptrdiff_t instance = -1)
    {
        const VectorXd &s = std::get<2>(state), &ds = std::get<5>(state);
        if (m_n == 0) m_n = s.size();
        if (s.size() != m_n || ds.size() != m_n) {
            log()->error("[FloatingBaseSystemAccelerationKinematics::setState] Wrong size of the joint vectors.");
            return false;
        }
        resize();
        const auto q = std::get<1>(state).quat().coeffs();
        for (std::size_t i = 0; i < m_N; i++) {
            if (instance >= 0 && static_cast<std::size_t>(instance) != i) continue;
            for (int c = 0; c < 3; c++)
                m_p[3 * i + c] = std::get<0>(state)[c], m_dp[3 * i + c] = std::get<3>(state)[c], m_w[3 * i + c] = std::get<4>(state).w[c];
            for (int c = 0; c < 4; c++) m_q[4 * i + c] = q[c];
            for (std::size_t j = 0; j < m_n; j++) m_s[m_n * i + j] = s[j], m_ds[m_n * i + j] = ds[j];
        }
        return true;
    }
    /** robot-major arrays: p [N][3], q [N][4] (x, y, z, w), s [N][n], dp [N][3], omega [N][3], ds [N][n] */
    bool setStateBatch(std::size_t n, const double* p, const double* q, const double* s, const double* dp, const double* omega, const double* ds)
    {
        m_n = n;
        resize();
        m_p.assign(p, p + 3 * m_N), m_q.assign(q, q + 4 * m_N), m_s.assign(s, s + n * m_N);
        m_dp.assign(dp, dp + 3 * m_N), m_w.assign(omega, omega + 3 * m_N), m_ds.assign(ds, ds + n * m_N);
        return true;
    }
    State getState(std::size_t instance = 0) const
    {
        State st;
        if (m_p.empty()) return st;
        for (int c = 0; c < 3; c++)
            std::get<0>(st)[c] = m_p[3 * instance + c], std::get<3>(st)[c] = m_dp[3 * instance + c], std::get<4>(st).w[c] = m_w[3 * instance + c];
        std::get<1>(st) = SO3d(Quaterniond{m_q[4 * instance], m_q[4 * instance + 1], m_q[4 * instance + 2], m_q[4 * instance + 3]});
        std::get<2>(st).assign(m_s.begin() + m_n * instance, m_s.begin() + m_n * (instance + 1));
        std::get<5>(st).assign(m_ds.begin() + m_n * instance, m_ds.begin() + m_n * (instance + 1));
        return st;
    }
    bool setControlInput(const Input& input, std::ptrdiff_t instance = -1)
    {
        const VectorXd& dds = std::get<1>(input);
        if (m_n == 0) m_n = dds.size();
        resize();
        if (dds.size() != m_n) {   // .cpp:46-51
            log()->error("[FloatingBaseSystemAccelerationKinematics::dynamics] Wrong size of the vectors.");
            return false;
        }
        const std::size_t nv = m_n + 6;
        for (std::size_t i = 0; i < m_N; i++) {
            if (instance >= 0 && static_cast<std::size_t>(instance) != i) continue;
            for (int c = 0; c < 6; c++) m_acc[nv * i + c] = std::get<0>(input)[c];
            for (std::size_t j = 0; j < m_n; j++) m_acc[nv * i + 6 + j] = dds[j];
        }
        m_inputSet = true;
        return true;
    }
    /** [N][6+n] robot-major generalized accelerations */
    bool setControlInputBatch(const double* acc)
    {
        if (m_n == 0) return false;
        resize();
        m_acc.assign(acc, acc + (m_n + 6) * m_N);
        m_inputSet = true;
        return true;
    }
    /** .cpp:20-67 */
    bool dynamics(const std::chrono::nanoseconds&, StateDerivative& d, std::size_t instance = 0) const
    {
        if (!m_inputSet) return false;
        const std::size_t nv = m_n + 6, i = instance;
        for (int c = 0; c < 3; c++) {
            std::get<0>(d)[c] = m_dp[3 * i + c], std::get<1>(d).w[c] = m_w[3 * i + c];
            std::get<3>(d)[c] = m_acc[nv * i + c], std::get<4>(d).w[c] = m_acc[nv * i + 3 + c];
        }
        std::get<2>(d).assign(m_ds.begin() + m_n * i, m_ds.begin() + m_n * (i + 1));
        std::get<5>(d).assign(m_acc.begin() + nv * i + 6, m_acc.begin() + nv * (i + 1));
        return true;
    }
    std::size_t dofs() const { return m_n; }
    std::vector<double>& basePositions() { return m_p; }
    std::vector<double>& baseQuaternions() { return m_q; }
    std::vector<double>& jointPositions() { return m_s; }
    std::vector<double>& baseLinearVelocities() { return m_dp; }
    std::vector<double>& baseAngularVelocities() { return m_w; }
    std::vector<double>& jointVelocities() { return m_ds; }
    const std::vector<double>& controlInput() const { return m_acc; }
    bool hasInput() const { return m_inputSet; }
    bool forwardEulerSteps(double dt, int nsteps, int device)
    {
        return blfik_integrate_acceleration_state(static_cast<int32_t>(m_n), static_cast<int64_t>(m_N), m_p.data(), m_q.data(), m_s.data(),
                                                  m_dp.data(), m_w.data(), m_ds.data(), m_acc.data(), dt, nsteps, device) == BLFIK_OK;
    }

private:
    void resize()
    {
        if (m_p.size() != 3 * m_N || m_s.size() != m_n * m_N) {
            m_p.assign(3 * m_N, 0.0), m_q.assign(4 * m_N, 0.0), m_s.assign(m_n * m_N, 0.0);
            m_dp.assign(3 * m_N, 0.0), m_w.assign(3 * m_N, 0.0), m_ds.assign(m_n * m_N, 0.0);
            for (std::size_t i = 0; i < m_N; i++) m_q[4 * i + 3] = 1.0;
        }
        if (m_acc.size() != (m_n + 6) * m_N) m_acc.assign((m_n + 6) * m_N, 0.0);
    }
    std::size_t m_N{1}, m_n{0};
    std::vector<double> m_p, m_q, m_s, m_dp, m_w, m_ds, m_acc;
    bool m_inputSet{false};
};

/** LinearTimeInvariantSystem.h:33-118, .cpp:14-115: dx = A x + B u.  A and B are row-major (rows x cols). */
class LinearTimeInvariantSystem
{
public:
    using State = std::tuple<VectorXd>;
    using StateDerivative = std::tuple<VectorXd>;
    using Input = std::tuple<VectorXd>;

    bool setSystemMatrices(const std::vector<double>& A, std::size_t rowsA, std::size_t colsA, const std::vector<double>& B, std::size_t rowsB,
                           std::size_t colsB)
    {
        constexpr auto errorPrefix = "[LinearTimeInvariantSystem::setSystemMatrices]";
        if (rowsA != rowsB) {
            log()->error("{} A and B must have the same number of rows.", errorPrefix);
            return false;
        }
        if (rowsA != colsA) {
            log()->error("{} A must be a square matrix.", errorPrefix);
            return false;
        }
        if (A.size() != rowsA * colsA || B.size() != rowsB * colsB) {
            log()->error("{} The storage does not match the sizes.", errorPrefix);
            return false;
        }
        m_A = A, m_B = B, m_nx = rowsA, m_nu = colsB;
        m_isInitialized = true;
        return true;
    }
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler>) { return true; }
    bool setState(const State& state) { m_x = std::get<0>(state); return true; }
    const State& getStateTuple() const { m_state = State{m_x}; return m_state; }
    State getState(std::size_t = 0) const { return State{m_x}; }
    bool setControlInput(const Input& input) { m_u = std::get<0>(input); return true; }
    bool dynamics(const std::chrono::nanoseconds&, StateDerivative& derivative) const { return derivativeAt(m_x, std::get<0>(derivative)); }
    /** dx = A x + B u at an arbitrary x (what RK4 evaluates at its intermediate points, RK4.h:122-160) */
    bool derivativeAt(const VectorXd& x, VectorXd& dx) const
    {
        constexpr auto errorPrefix = "[LinearTimeInvariantSystem::dynamics]";
        if (!m_isInitialized) {
            log()->error("{} Please initialize the matrices. Call setSystemMatrices()", errorPrefix);
            return false;
        }
        if (x.size() != m_nx) {
            log()->error("{} The size of the vector 'state' is not coherent with the system matrices.", errorPrefix);
            return false;
        }
        if (m_u.size() != m_nu) {
            log()->error("{} The size of the vector 'control input' is not coherent with the system matrices.", errorPrefix);
            return false;
        }
        dx.assign(m_nx, 0.0);
        for (std::size_t r = 0; r < m_nx; r++) {
            double acc = 0.0;
            for (std::size_t c = 0; c < m_nx; c++) acc += m_A[r * m_nx + c] * x[c];
            for (std::size_t c = 0; c < m_nu; c++) acc += m_B[r * m_nu + c] * m_u[c];
            dx[r] = acc;
        }
        return true;
    }
    std::size_t getBatchSize() const { return 1; }
    bool hasInput() const { return m_isInitialized && m_u.size() == m_nu; }
    VectorXd& state() { return m_x; }
    bool forwardEulerSteps(double dt, int nsteps, int)
    {
        VectorXd dx;
        for (int k = 0; k < nsteps; k++) {
            if (!derivativeAt(m_x, dx)) return false;
            for (std::size_t i = 0; i < m_nx; i++) m_x[i] += dt * dx[i];
        }
        return true;
    }

private:
    std::vector<double> m_A, m_B;
    std::size_t m_nx{0}, m_nu{0};
    VectorXd m_x, m_u;
    mutable State m_state;
    bool m_isInitialized{false};
};

/** RK4.h:37-175: x+ = x + dT/6 (k1 + 2 k2 + 2 k3 + k4) with the input held over the step; like the reference, only for systems
 *  whose state is a vector of R^n (the tableau adds states and scales derivatives; RK4.h:23-35 says so for Lie-group states).
 *  The system offers `state()` and `derivativeAt(x, dx)`.  Stepping follows FixedStepIntegrator.h:94-140 like ForwardEuler. */
template <class _DynamicalSystem> class RK4
{
public:
    using DynamicalSystem = _DynamicalSystem;
    using State = typename _DynamicalSystem::State;

    bool setDynamicalSystem(std::shared_ptr<DynamicalSystem> dynamicalSystem)
    {
        constexpr auto errorPrefix = "[Integrator::setDynamicalSystem]";
        if (m_dynamicalSystem != nullptr) {
            log()->error("{} The dynamical system has been already set.", errorPrefix);
            return false;
        }
        if (dynamicalSystem == nullptr) {
            log()->error("{} The dynamical system passed to the function is corrupted.", errorPrefix);
            return false;
        }
        m_dynamicalSystem = dynamicalSystem;
        return true;
    }
    std::weak_ptr<DynamicalSystem> dynamicalSystem() const { return m_dynamicalSystem; }
    bool setIntegrationStep(const std::chrono::nanoseconds& dT)
    {
        if (dT <= std::chrono::nanoseconds::zero()) {
            log()->error("[FixedStepIntegrator::setIntegrationStep] The integration must be a strictly positive number.");
            return false;
        }
        m_dT = dT;
        return true;
    }
    std::chrono::nanoseconds getIntegrationStep() const { return m_dT; }
    bool integrate(const std::chrono::nanoseconds& initialTime, const std::chrono::nanoseconds& finalTime)
    {
        constexpr auto errorPrefix = "[FixedStepIntegrator::integrate]";
        if (m_dynamicalSystem == nullptr) {
            log()->error("{} Please set the dynamical system before call this function.", errorPrefix);
            return false;
        }
        if (initialTime > finalTime) {
            log()->error("{} The final time has to be greater than the initial one.", errorPrefix);
            return false;
        }
        if (m_dT <= std::chrono::nanoseconds::zero()) {
            log()->error("{} Please set the integration step.", errorPrefix);
            return false;
        }
        const std::size_t iterations = static_cast<std::size_t>((finalTime - initialTime) / m_dT);
        if (iterations < 1) {
            log()->error("{} The integration interval is shorter than the integration step.", errorPrefix);
            return false;
        }
        const double dT = std::chrono::duration<double>(m_dT).count();
        const auto lastStep = finalTime - (initialTime + m_dT * static_cast<long long>(iterations - 1));
        for (std::size_t k = 0; k + 1 < iterations; k++)
            if (!oneStep(dT)) return false;
        return oneStep(std::chrono::duration<double>(lastStep).count());
    }
    State getSolution(std::size_t instance = 0) const { return m_dynamicalSystem ? m_dynamicalSystem->getState(instance) : State(); }

private:
    bool oneStep(double dt)
    {
        auto& sys = *m_dynamicalSystem;
        VectorXd& x = sys.state();
        const std::size_t n = x.size();
        m_tmp.resize(n);
        if (!sys.derivativeAt(x, m_k1)) return fail();
        for (std::size_t i = 0; i < n; i++) m_tmp[i] = x[i] + 0.5 * dt * m_k1[i];
        if (!sys.derivativeAt(m_tmp, m_k2)) return fail();
        for (std::size_t i = 0; i < n; i++) m_tmp[i] = x[i] + 0.5 * dt * m_k2[i];
        if (!sys.derivativeAt(m_tmp, m_k3)) return fail();
        for (std::size_t i = 0; i < n; i++) m_tmp[i] = x[i] + dt * m_k3[i];
        if (!sys.derivativeAt(m_tmp, m_k4)) return fail();
        for (std::size_t i = 0; i < n; i++) x[i] += dt / 6.0 * (m_k1[i] + 2.0 * m_k2[i] + 2.0 * m_k3[i] + m_k4[i]);
        return true;
    }
    static bool fail()
    {
        log()->error("[RK4::oneStepIntegration] Unable to compute the system dynamics.");
        return false;
    }
    std::shared_ptr<DynamicalSystem> m_dynamicalSystem;
    std::chrono::nanoseconds m_dT{0};
    VectorXd m_k1, m_k2, m_k3, m_k4, m_tmp;
};

/** FirstOrderSmoother.h:32-94, FirstOrderSmoother.cpp:18-140: dx = -a x + a u with a = 3 / settling_time, propagated by one
 *  ForwardEuler step of `sampling_time` per advance().  Host-side: it smooths the few doubles of a task weight, which the
 *  solver re-reads on every tick (QPInverseKinematics.cpp:450-451) -- shared by all robots of the batch. */
class FirstOrderSmoother
{
public:
    /** keys: settling_time [s] (double), sampling_time (chrono nanoseconds or seconds as double) */
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler)
    {
        constexpr auto logPrefix = "[FirstOrderSmoother::initialize]";
        auto ptr = handler.lock();
        if (ptr == nullptr) {
            log()->error("{} Invalid parameter handler.", logPrefix);
            return false;
        }
        if (!ptr->getParameter("settling_time", m_settlingTime)) {
            log()->error("{} Unable to get the 'settling_time' parameter.", logPrefix);
            return false;
        }
        std::chrono::nanoseconds samplingTime;
        if (!ptr->getParameter("sampling_time", samplingTime)) {
            log()->error("{} Unable to get the 'sampling_time' parameter.", logPrefix);
            return false;
        }
        if (samplingTime <= std::chrono::nanoseconds::zero()) {   // FixedStepIntegrator::setIntegrationStep
            log()->error("{} Unable to set the integration time step.", logPrefix);
            return false;
        }
        m_dt = std::chrono::duration<double>(samplingTime).count();
        m_isInitialized = true;
        return true;
    }
    bool reset(const VectorXd& initialPoint)
    {
        m_isInitialStateSet = false;
        if (!m_isInitialized) {
            log()->error("[FirstOrderSmoother::reset] Please initialize the class before.");
            return false;
        }
        m_state = initialPoint, m_output = initialPoint;
        m_isInitialStateSet = true;
        return true;
    }
    bool setInput(const VectorXd& input)
    {
        if (!m_isInitialized || !m_isInitialStateSet) {
            log()->error("[FirstOrderSmoother::setInput] Please call initialize() and reset() before advance.");
            return false;
        }
        m_input = input;
        return true;
    }
    bool advance()
    {
        m_isOutputValid = false;
        if (!m_isInitialized || !m_isInitialStateSet) {
            log()->error("[FirstOrderSmoother::advance] Please call initialize() and reset() before advance.");
            return false;
        }
        if (m_input.size() != m_state.size()) {   // LinearTimeInvariantSystem.cpp:58-63: the input was never set
            log()->error("[FirstOrderSmoother::advance] Unable to propagate the dynamical system.");
            return false;
        }
        const double a = 3.0 / m_settlingTime;
        for (std::size_t i = 0; i < m_state.size(); i++) m_state[i] += m_dt * (a * m_input[i] - a * m_state[i]);
        m_output = m_state;
        m_isOutputValid = true;
        return true;
    }
    const VectorXd& getOutput() const { return m_output; }
    bool isOutputValid() const { return m_isOutputValid; }

private:
    double m_settlingTime{1.0}, m_dt{0.0};
    VectorXd m_state, m_input, m_output;
    bool m_isInitialized{false}, m_isInitialStateSet{false}, m_isOutputValid{false};
};

/** MultiStateWeightProvider.h:28-93, MultiStateWeightProvider.cpp:12-134: named weight vectors ("states") with a first-order
 *  transition between them; plug it into QPInverseKinematics::addTask / setTaskWeight as the weight provider. */
class MultiStateWeightProvider : public System::WeightProvider
{
public:
    /** keys: settling_time, sampling_time, states (vector<string> of group names); each group: name, weight */
    bool initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler)
    {
        constexpr auto logPrefix = "[MultiStateWeightProvider::initialize]";
        if (!m_smoother.initialize(handler)) {
            log()->error("{} Unable to initialize the smoother.", logPrefix);
            return false;
        }
        auto ptr = handler.lock();
        if (ptr == nullptr) {
            log()->error("{} Invalid parameter handler.", logPrefix);
            return false;
        }
        std::vector<std::string> states;
        if (!ptr->getParameter("states", states) || states.empty()) {
            log()->error("{} Unable to get the parameter named 'states'.", logPrefix);
            return false;
        }
        std::string name, firstState;
        for (const auto& state : states) {
            auto group = ptr->getGroup(state).lock();
            if (group == nullptr) {
                log()->error("{} Unable to get the group named '{}'.", logPrefix, state);
                return false;
            }
            if (!group->getParameter("name", name)) {
                log()->error("{} Unable to get the parameter named 'name' in the group named '{}'.", logPrefix, state);
                return false;
            }
            if (!group->getParameter("weight", m_states[name])) {
                log()->error("{} Unable to get the parameter named 'weight' in the group named '{}'.", logPrefix, state);
                return false;
            }
            if (firstState.empty()) firstState = name;
        }
        const std::size_t size = m_states.begin()->second.size();
        for (const auto& state : m_states)
            if (state.second.size() != size) {
                log()->error("{} The size of the weight associated to the state named '{}' does not match with the size of the weight "
                             "associated to the state named '{}'.", logPrefix, state.first, m_states.begin()->first);
                return false;
            }
        return this->reset(firstState);
    }
    bool reset(const std::string& state)
    {
        const auto weight = m_states.find(state);
        if (weight == m_states.end()) {
            log()->error("[MultiStateWeightProvider::reset] Unable to find the weight named '{}'", state);
            return false;
        }
        return m_smoother.reset(weight->second);
    }
    bool setState(const std::string& state)
    {
        const auto weight = m_states.find(state);
        if (weight == m_states.end()) {
            log()->error("[MultiStateWeightProvider::setState] Unable to find the weight named '{}'", state);
            return false;
        }
        return m_smoother.setInput(weight->second);
    }
    bool advance() { return m_smoother.advance(); }
    const VectorXd& getOutput() const override { return m_smoother.getOutput(); }
    /** like the reference (MultiStateWeightProvider.cpp:17-20): valid once advance() has run */
    bool isOutputValid() const override { return m_smoother.isOutputValid(); }

private:
    FirstOrderSmoother m_smoother;
    std::map<std::string, VectorXd> m_states;
};

}  // namespace ContinuousDynamicalSystem
}  // namespace BipedalLocomotion

#endif
