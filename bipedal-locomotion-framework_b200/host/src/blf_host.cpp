// blf_host.cpp -- implementation of the host classes (KinDynBatch, the IK tasks, QPInverseKinematics).
// Every numerical step goes through the C ABI of include/blfik.h; nothing here computes kinematics or solves QPs.
#include <BipedalLocomotion/ContinuousDynamicalSystem.h>
#include <BipedalLocomotion/IK.h>

#include <algorithm>
#include <thread>
#include <fstream>

namespace BipedalLocomotion
{

// =================================================================== KinDynBatch
KinDynBatch::~KinDynBatch()
{
    if (m_kin) blfik_problem_destroy(m_kin);
    if (m_model) blfik_model_destroy(m_model);
}

bool KinDynBatch::loadRobotModelFromFile(const std::string& path)
{
    constexpr auto logPrefix = "[KinDynBatch::loadRobotModelFromFile]";
    std::ifstream in(path);
    if (!in) {
        log()->error("{} Unable to open {}.", logPrefix, path);
        return false;
    }
    std::string tok, header;
    int version = 0, n = 0, nf = 0;
    in >> header >> version;
    if (header != "blfik_model") {
        log()->error("{} {} is not a blfik model file.", logPrefix, path);
        return false;
    }
    std::vector<int> parent, dof, frameLink;
    std::vector<double> Xr, Xp, axis, mass, com, Fr, Fp, inertia;
    std::vector<std::string> jointOfLink;
    std::map<std::string, std::array<double, 3>> limits;
    while (in >> tok) {
        if (tok == "name") {
            in >> tok;
        } else if (tok == "n") {
            in >> n;
        } else if (tok == "nf") {
            in >> nf;
        } else if (tok == "link") {
            std::string lname, jname;
            int p, d;
            in >> lname >> p >> d >> jname;
            parent.push_back(p), dof.push_back(d), jointOfLink.push_back(jname);
            double v;
            for (int i = 0; i < 9; i++) in >> v, Xr.push_back(v);
            for (int i = 0; i < 3; i++) in >> v, Xp.push_back(v);
            for (int i = 0; i < 3; i++) in >> v, axis.push_back(v);
            in >> v, mass.push_back(v);
            for (int i = 0; i < 3; i++) in >> v, com.push_back(v);
        } else if (tok == "frame") {
            std::string fname;
            int l;
            in >> fname >> l;
            m_frameNames.push_back(fname), frameLink.push_back(l);
            double v;
            for (int i = 0; i < 9; i++) in >> v, Fr.push_back(v);
            for (int i = 0; i < 3; i++) in >> v, Fp.push_back(v);
            if (l == 0 && m_floatingBase.empty()) {   // the link frame of link 0 (identity offset) names the floating base
                const double* r = &Fr[Fr.size() - 9];
                const double* t = &Fp[Fp.size() - 3];
                if (r[0] == 1 && r[4] == 1 && r[8] == 1 && r[1] == 0 && r[2] == 0 && r[3] == 0 && r[5] == 0 && r[6] == 0 && r[7] == 0 && t[0] == 0 && t[1] == 0 && t[2] == 0)
                    m_floatingBase = fname;
            }
        } else if (tok == "inertia") {
            std::string lname;
            int k;
            in >> lname >> k;
            if (inertia.size() < static_cast<std::size_t>(6 * (k + 1))) inertia.resize(6 * (k + 1), 0.0);
            for (int i = 0; i < 6; i++) in >> inertia[6 * k + i];
        } else if (tok == "limits") {
            std::string jname;
            std::array<double, 3> l;
            in >> jname >> l[0] >> l[1] >> l[2];
            limits[jname] = l;
        }
    }
    if (n < 1 || static_cast<int>(parent.size()) != n + 1 || static_cast<int>(frameLink.size()) != nf) {
        log()->error("{} Malformed model file {}.", logPrefix, path);
        return false;
    }
    m_n = static_cast<std::size_t>(n);
    m_jointNames.assign(n, "");
    for (int k = 1; k <= n; k++) m_jointNames[dof[k]] = jointOfLink[k];
    m_qLower.assign(n, -BLFIK_INFTY), m_qUpper.assign(n, BLFIK_INFTY);
    for (int j = 0; j < n; j++) {
        auto it = limits.find(m_jointNames[j]);
        if (it != limits.end()) m_qLower[j] = it->second[0], m_qUpper[j] = it->second[1];
    }
    blfik_model_desc d;
    d.n_dofs = n, d.n_frames = nf, d.parent = parent.data(), d.dof = dof.data(), d.X_rot = Xr.data(), d.X_pos = Xp.data(), d.axis = axis.data();
    d.mass = mass.data(), d.com = com.data(), d.frame_link = frameLink.data(), d.frame_rot = Fr.data(), d.frame_pos = Fp.data();
    d.inertia = inertia.size() == static_cast<std::size_t>(6 * (n + 1)) ? inertia.data() : nullptr;
    if (blfik_model_create(&d, &m_model) != BLFIK_OK) {
        log()->error("{} {}", logPrefix, blfik_last_error());
        return false;
    }
    return setBatchSize(m_N, m_device);
}

bool KinDynBatch::setBatchSize(std::size_t N, int device)
{
    if (N < 1) {
        log()->error("[KinDynBatch::setBatchSize] The batch size must be positive.");
        return false;
    }
    m_N = N, m_device = device;
    m_p.assign(3 * N, 0.0), m_q.assign(4 * N, 0.0), m_s.assign(m_n * N, 0.0);
    for (std::size_t i = 0; i < N; i++) m_q[4 * i + 3] = 1.0;
    m_version++;
    if (m_kin) {
        blfik_problem_destroy(m_kin);
        m_kin = nullptr;
    }
    return true;
}

int KinDynBatch::getFrameIndex(const std::string& name) const
{
    auto it = std::find(m_frameNames.begin(), m_frameNames.end(), name);
    return it == m_frameNames.end() ? -1 : static_cast<int>(it - m_frameNames.begin());
}

std::string KinDynBatch::getFrameName(int index) const
{
    return (index >= 0 && index < static_cast<int>(m_frameNames.size())) ? m_frameNames[index] : std::string();
}

bool KinDynBatch::setRobotState(const Vector3d& p, const SO3d& R, const VectorXd& s, std::size_t instance)
{
    if (!isValid() || s.size() != m_n || instance >= m_N) {
        log()->error("[KinDynBatch::setRobotState] Invalid model, joint vector size or instance index.");
        return false;
    }
    const auto q = R.quat().coeffs();
    for (int c = 0; c < 3; c++) m_p[3 * instance + c] = p[c];
    for (int c = 0; c < 4; c++) m_q[4 * instance + c] = q[c];
    std::copy(s.begin(), s.end(), m_s.begin() + m_n * instance);
    m_version++;
    mirrorState(instance, instance + 1);
    return true;
}

void KinDynBatch::bindRecords(double* base, std::size_t stride)
{
    m_recBase = base, m_recStride = stride;
    mirrorState(0, m_N);
}

void KinDynBatch::mirrorState(std::size_t a, std::size_t b)
{
    if (!m_recBase) return;
    for (std::size_t i = a; i < b; i++) {
        double* r = m_recBase + i * m_recStride;
        std::copy_n(m_p.data() + 3 * i, 3, r);
        std::copy_n(m_q.data() + 4 * i, 4, r + 3);
        std::copy_n(m_s.data() + m_n * i, m_n, r + 7);
    }
}

bool KinDynBatch::setRobotStateBatch(const double* p, const double* q, const double* s)
{
    if (!isValid() || !p || !q || !s) return false;
    m_p.assign(p, p + 3 * m_N), m_q.assign(q, q + 4 * m_N), m_s.assign(s, s + m_n * m_N);
    m_version++;
    mirrorState(0, m_N);
    return true;
}

bool KinDynBatch::getJointPos(VectorXd& s, std::size_t instance) const
{
    if (!isValid() || instance >= m_N) return false;
    s.assign(m_s.begin() + m_n * instance, m_s.begin() + m_n * (instance + 1));
    return true;
}

bool KinDynBatch::ensureKinProblem()
{
    if (!isValid()) return false;
    if (!m_kin) {
        // a problem with one trivial cost task: it only owns the device copy of the state for the K1 kernel
        std::vector<double> ones(m_n, 1.0);
        blfik_task_desc t{};
        t.type = BLFIK_TASK_JOINT_TRACKING, t.priority = 1, t.frame = -1, t.controller_enabled = 1;
        t.kp_joint = ones.data(), t.weight = ones.data();
        if (blfik_problem_create(m_model, &t, 1, static_cast<int64_t>(m_N), m_device, &m_kin) != BLFIK_OK) {
            log()->error("[KinDynBatch] {}", blfik_last_error());
            return false;
        }
        m_kinVersion = 0;
    }
    if (m_kinVersion != m_version) {
        if (blfik_set_state(m_kin, m_p.data(), m_q.data(), m_s.data(), BLFIK_MEM_HOST) != BLFIK_OK) {
            log()->error("[KinDynBatch] {}", blfik_last_error());
            return false;
        }
        m_kinVersion = m_version;
    }
    return true;
}

bool KinDynBatch::evalFrame(int frame)
{
    if (frame < 0 || frame >= static_cast<int>(m_frameNames.size())) {
        log()->error("[KinDynBatch] Frame index {} out of range.", frame);
        return false;
    }
    if (!ensureKinProblem()) return false;
    if (m_cachedFrame == frame && m_frameVersion == m_version) return true;
    const std::size_t nv = m_n + 6;
    m_framePos.resize(3 * m_N), m_frameRot.resize(9 * m_N), m_frameJac.resize(6 * nv * m_N);
    const int32_t f = frame;
    if (blfik_kinematics(m_kin, &f, 1, m_framePos.data(), m_frameRot.data(), m_frameJac.data(), nullptr, nullptr, BLFIK_MEM_HOST) != BLFIK_OK) {
        log()->error("[KinDynBatch] {}", blfik_last_error());
        return false;
    }
    m_cachedFrame = frame, m_frameVersion = m_version;
    return true;
}

bool KinDynBatch::evalCom()
{
    if (!ensureKinProblem()) return false;
    if (m_comVersion == m_version) return true;
    const std::size_t nv = m_n + 6;
    m_com.resize(3 * m_N), m_comJac.resize(3 * nv * m_N);
    if (blfik_kinematics(m_kin, nullptr, 0, nullptr, nullptr, nullptr, m_com.data(), m_comJac.data(), BLFIK_MEM_HOST) != BLFIK_OK) {
        log()->error("[KinDynBatch] {}", blfik_last_error());
        return false;
    }
    m_comVersion = m_version;
    return true;
}

// rotation matrix -> quaternion -> normalise (Conversions::toManifRot, ManifConversions.h:44-50)
static Quaterniond rotToQuat(const std::array<double, 9>& R)
{
    Quaterniond q;
    double t = R[0] + R[4] + R[8];
    if (t > 0) {
        t = std::sqrt(t + 1.0);
        q.w = 0.5 * t, t = 0.5 / t;
        q.x = (R[7] - R[5]) * t, q.y = (R[2] - R[6]) * t, q.z = (R[3] - R[1]) * t;
    } else {
        int i = 0;
        if (R[4] > R[0]) i = 1;
        if (R[8] > R[4 * i]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        double v[3];
        t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
        v[i] = 0.5 * t, t = 0.5 / t;
        q.w = (R[3 * k + j] - R[3 * j + k]) * t;
        v[j] = (R[3 * j + i] + R[3 * i + j]) * t, v[k] = (R[3 * k + i] + R[3 * i + k]) * t;
        q.x = v[0], q.y = v[1], q.z = v[2];
    }
    q.normalize();
    return q;
}

bool KinDynBatch::getWorldTransform(int frame, SE3d& pose, std::array<double, 9>& R, std::size_t instance)
{
    if (instance >= m_N || !evalFrame(frame)) return false;
    for (int c = 0; c < 9; c++) R[c] = m_frameRot[9 * instance + c];
    pose.p = {m_framePos[3 * instance], m_framePos[3 * instance + 1], m_framePos[3 * instance + 2]};
    pose.R = SO3d(rotToQuat(R));
    return true;
}

bool KinDynBatch::getFrameFreeFloatingJacobian(int frame, System::MatrixXd& J, std::size_t instance)
{
    if (instance >= m_N || !evalFrame(frame)) return false;
    const std::size_t nv = m_n + 6;
    J.resize(6, nv);
    std::copy(m_frameJac.begin() + 6 * nv * instance, m_frameJac.begin() + 6 * nv * (instance + 1), J.data.begin());
    return true;
}

bool KinDynBatch::getCenterOfMassPosition(Vector3d& com, std::size_t instance)
{
    if (instance >= m_N || !evalCom()) return false;
    com = {m_com[3 * instance], m_com[3 * instance + 1], m_com[3 * instance + 2]};
    return true;
}

bool KinDynBatch::getCenterOfMassJacobian(System::MatrixXd& J, std::size_t instance)
{
    if (instance >= m_N || !evalCom()) return false;
    const std::size_t nv = m_n + 6;
    J.resize(3, nv);
    std::copy(m_comJac.begin() + 3 * nv * instance, m_comJac.begin() + 3 * nv * (instance + 1), J.data.begin());
    return true;
}

bool KinDynBatch::setRobotVelocity(const std::array<double, 6>& baseVelocity, const VectorXd& jointVelocities, std::size_t instance)
{
    if (!isValid() || jointVelocities.size() != m_n || instance >= m_N) {
        log()->error("[KinDynBatch::setRobotVelocity] Invalid model, joint vector size or instance index.");
        return false;
    }
    const std::size_t nv = m_n + 6;
    if (m_nu.size() != nv * m_N) m_nu.assign(nv * m_N, 0.0);
    std::copy(baseVelocity.begin(), baseVelocity.end(), m_nu.begin() + nv * instance);
    std::copy(jointVelocities.begin(), jointVelocities.end(), m_nu.begin() + nv * instance + 6);
    m_velVersion++;
    return true;
}

bool KinDynBatch::setRobotVelocityBatch(const double* nu)
{
    if (!isValid() || !nu) return false;
    m_nu.assign(nu, nu + (m_n + 6) * m_N);
    m_velVersion++;
    return true;
}

bool KinDynBatch::evalDynamics()
{
    if (!ensureKinProblem()) return false;
    if (m_dynVersion == m_version) return true;
    const std::size_t nv = m_n + 6;
    m_mass.resize(nv * nv * m_N), m_gravityForces.resize(nv * m_N);
    if (blfik_dynamics(m_kin, m_mass.data(), m_gravityForces.data(), m_gravity, BLFIK_MEM_HOST) != BLFIK_OK) {
        log()->error("[KinDynBatch] {}", blfik_last_error());
        return false;
    }
    m_dynVersion = m_version;
    return true;
}

bool KinDynBatch::evalBias(int frame)
{
    if (frame >= static_cast<int>(m_frameNames.size())) {
        log()->error("[KinDynBatch] Frame index {} out of range.", frame);
        return false;
    }
    if (!ensureKinProblem()) return false;
    const std::size_t nv = m_n + 6;
    if (m_nu.size() != nv * m_N) m_nu.assign(nv * m_N, 0.0);   // a velocity never set is zero, as in iDynTree
    if (m_biasVersion == m_version && m_biasVelVersion == m_velVersion && (frame < 0 || frame == m_biasFrame)) return true;
    if (frame < 0) frame = m_biasFrame;                       // keep the frame of the last query in the cache
    m_bias.resize(nv * m_N), m_comBias.resize(3 * m_N), m_frameBias.resize(6 * m_N), m_frameVel.resize(6 * m_N);
    const int32_t f = frame;
    if (blfik_bias_forces(m_kin, m_nu.data(), m_gravity, m_bias.data(), frame >= 0 ? &f : nullptr, frame >= 0 ? 1 : 0,
                          frame >= 0 ? m_frameBias.data() : nullptr, frame >= 0 ? m_frameVel.data() : nullptr, m_comBias.data(), BLFIK_MEM_HOST) != BLFIK_OK) {
        log()->error("[KinDynBatch] {}", blfik_last_error());
        return false;
    }
    m_biasVersion = m_version, m_biasVelVersion = m_velVersion, m_biasFrame = frame;
    return true;
}

bool KinDynBatch::getFreeFloatingMassMatrix(System::MatrixXd& M, std::size_t instance)
{
    if (instance >= m_N || !evalDynamics()) return false;
    const std::size_t nv = m_n + 6;
    M.resize(nv, nv);
    std::copy(m_mass.begin() + nv * nv * instance, m_mass.begin() + nv * nv * (instance + 1), M.data.begin());
    return true;
}

bool KinDynBatch::generalizedGravityForces(VectorXd& g, std::size_t instance)
{
    if (instance >= m_N || !evalDynamics()) return false;
    const std::size_t nv = m_n + 6;
    g.assign(m_gravityForces.begin() + nv * instance, m_gravityForces.begin() + nv * (instance + 1));
    return true;
}

bool KinDynBatch::generalizedBiasForces(VectorXd& h, std::size_t instance)
{
    if (instance >= m_N || !evalBias(-1)) return false;
    const std::size_t nv = m_n + 6;
    h.assign(m_bias.begin() + nv * instance, m_bias.begin() + nv * (instance + 1));
    return true;
}

bool KinDynBatch::getFrameBiasAcc(int frame, std::array<double, 6>& a, std::size_t instance)
{
    if (instance >= m_N || frame < 0 || !evalBias(frame)) return false;
    std::copy_n(m_frameBias.begin() + 6 * instance, 6, a.begin());
    return true;
}

bool KinDynBatch::getFrameVel(int frame, std::array<double, 6>& v, std::size_t instance)
{
    if (instance >= m_N || frame < 0 || !evalBias(frame)) return false;
    std::copy_n(m_frameVel.begin() + 6 * instance, 6, v.begin());
    return true;
}

bool KinDynBatch::getCenterOfMassBiasAcc(Vector3d& a, std::size_t instance)
{
    if (instance >= m_N || !evalBias(-1)) return false;
    a = {m_comBias[3 * instance], m_comBias[3 * instance + 1], m_comBias[3 * instance + 2]};
    return true;
}

namespace IK
{

// =================================================================== helpers for the inspected A / b
static Quaterniond quatMul(const Quaterniond& a, const Quaterniond& b)
{
    return {a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x,
            a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w, a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z};
}

// LieGroupControllers P-law on SO(3): Kp o Log(R_d R^-1) + feed-forward (SE3Task.cpp:261-281)
static Vector3d so3Control(const Quaterniond& qd, const Quaterniond& qc, const Vector3d& ff, const Vector3d& kp, bool enabled)
{
    const Quaterniond e = quatMul(qd, Quaterniond{-qc.x, -qc.y, -qc.z, qc.w});
    const double n2 = e.x * e.x + e.y * e.y + e.z * e.z;
    double scale;
    if (n2 > 1e-20) {
        const double n = std::sqrt(n2);
        scale = 2.0 * (e.w < 0 ? std::atan2(-n, -e.w) : std::atan2(n, e.w)) / n;
    } else {
        scale = 2.0 / e.w;
    }
    const Vector3d err{scale * e.x, scale * e.y, scale * e.z};
    Vector3d out;
    for (int c = 0; c < 3; c++) out[c] = enabled ? ff[c] + kp[c] * err[c] : ff[c];
    return out;
}

static void setRow(System::MatrixXd& A, std::size_t row, std::ptrdiff_t offset, const System::MatrixXd& J, std::size_t srcRow)
{
    for (std::size_t c = 0; c < J.cols; c++) A(row, offset + c) = J(srcRow, c);
}

static const double* keepVector(std::vector<std::vector<double>>& keep, const VectorXd& v)
{
    keep.push_back(v);
    return keep.back().data();
}

// =================================================================== SE3Task
bool SE3Task::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[SE3Task::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    if (!ptr->getParameter("frame_name", m_frameName) || (m_frameIndex = m_kinDyn->getFrameIndex(m_frameName)) < 0) {
        log()->error("{} Error while retrieving the frame that should be controlled.", errorPrefix);
        return false;
    }
    if (!readGain3(*ptr, "kp_linear", m_kpLinear)) {
        log()->error("{} Unable to get the kp_linear gain.", errorPrefix);
        return false;
    }
    if (!readGain3(*ptr, "kp_angular", m_kpAngular)) {
        log()->error("{} Unable to get the kp_angular gain.", errorPrefix);
        return false;
    }
    if (!readMask(*ptr, m_mask, m_linearDoFs)) {
        log()->error("{} The size of the mask must be 3.", errorPrefix);
        return false;
    }
    // optional, default false (SE3Task.cpp:196-214)
    m_useOrientationExogenousFeedback = m_usePositionExogenousFeedback = false;
    ptr->getParameter("use_orientation_exogenous_feedback", m_useOrientationExogenousFeedback);
    ptr->getParameter("use_position_exogenous_feedback", m_usePositionExogenousFeedback);
    m_description = "SE3Task Optimal Control Element - Frame name: " + m_frameName + ".";
    m_isInitialized = true;
    return true;
}

static const std::vector<double> kIdentityPoseFeedback{0, 0, 0, 0, 0, 0, 1};   // controllers' initial state: p = 0, q = identity

bool SE3Task::setFeedback(const SE3d& I_H_F, std::ptrdiff_t instance)
{
    return setFeedback(I_H_F.p, instance) && setFeedback(I_H_F.R, instance);
}

bool SE3Task::setFeedback(const Vector3d& I_p_F, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    if (m_usePositionExogenousFeedback) storeFeedback(I_p_F.data(), 0, 3, instance, kIdentityPoseFeedback);   // ignored otherwise (:345-353)
    return true;
}

bool SE3Task::setFeedback(const SO3d& I_R_F, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    if (m_useOrientationExogenousFeedback) {
        const auto q = I_R_F.quat().coeffs();
        const double qq[4] = {q[0], q[1], q[2], q[3]};
        storeFeedback(qq, 3, 4, instance, kIdentityPoseFeedback);
    }
    return true;
}

bool SE3Task::setSetPoint(const SE3d& I_H_F, const SE3d::Tangent& v, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    const auto q = I_H_F.quat().coeffs();
    const double rec[13] = {I_H_F.p[0], I_H_F.p[1], I_H_F.p[2], q[0], q[1], q[2], q[3], v.linear[0], v.linear[1], v.linear[2], v.angular[0], v.angular[1], v.angular[2]};
    storeSetPoint(rec, instance);
    m_isSetPointSetAtLeastOnce = true;
    return true;
}

bool SE3Task::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[SE3Task::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("[SE3Task::update] The set point has not been set at least once.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        SE3d pose;
        std::array<double, 9> R;
        System::MatrixXd J;
        if (!m_kinDyn->getWorldTransform(m_frameIndex, pose, R, i) || !m_kinDyn->getFrameFreeFloatingJacobian(m_frameIndex, J, i)) {
            log()->error("[SE3Task::update] Unable to get the jacobian.");
            return m_isValid;
        }
        const double* sp = m_sp.data() + 13 * i;
        const bool on = m_controllerMode == Mode::Enable;
        // state of the controllers: the model's pose, or the caller's measurement (SE3Task.cpp:261-268)
        const double* fb = m_fb.size() >= 7 * static_cast<std::size_t>(i + 1) ? m_fb.data() + 7 * i : kIdentityPoseFeedback.data();
        Vector3d pc = pose.p;
        Quaterniond qc = pose.R.quat();
        if (m_usePositionExogenousFeedback) pc = {fb[0], fb[1], fb[2]};
        if (m_useOrientationExogenousFeedback) qc = Quaterniond{fb[3], fb[4], fb[5], fb[6]};
        std::size_t idx = 0;
        for (int c = 0; c < 3; c++)
            if (m_mask[c]) {
                m_b[idx] = on ? sp[7 + c] + m_kpLinear[c] * (sp[c] - pc[c]) : sp[7 + c];
                setRow(m_A, idx++, m_robotVelocityVariable.offset, J, c);
            }
        const Vector3d w = so3Control(Quaterniond{sp[3], sp[4], sp[5], sp[6]}, qc, {sp[10], sp[11], sp[12]}, m_kpAngular, on);
        for (int c = 0; c < 3; c++) {
            m_b[idx + c] = w[c];
            setRow(m_A, idx + c, m_robotVelocityVariable.offset, J, 3 + c);
        }
    }
    m_isValid = true;
    return m_isValid;
}

bool SE3Task::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>&) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_SE3, d.frame = m_frameIndex, d.controller_enabled = m_controllerMode == Mode::Enable;
    d.use_position_exogenous_feedback = m_usePositionExogenousFeedback, d.use_orientation_exogenous_feedback = m_useOrientationExogenousFeedback;
    for (int c = 0; c < 3; c++) d.mask[c] = m_mask[c], d.kp_linear[c] = m_kpLinear[c], d.kp_angular[c] = m_kpAngular[c];
    return m_isInitialized;
}

// =================================================================== CoMTask
bool CoMTask::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[CoMTask::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    if (!readGain3(*ptr, "kp_linear", m_kpLinear)) {
        log()->error("{} Unable to get the kp_linear gain.", errorPrefix);
        return false;
    }
    if (!readMask(*ptr, m_mask, m_DoFs)) {
        log()->error("{} The size of the mask must be 3.", errorPrefix);
        return false;
    }
    m_useExogenousFeedback = false;   // optional (CoMTask.cpp:116-123)
    ptr->getParameter("use_exogenous_feedback", m_useExogenousFeedback);
    m_description = "CoM Task";
    m_isInitialized = true;
    return true;
}

bool CoMTask::setFeedback(const Vector3d& feedback, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    static const std::vector<double> zero{0, 0, 0};
    if (m_useExogenousFeedback) storeFeedback(feedback.data(), 0, 3, instance, zero);   // ignored otherwise (:267-270)
    return true;
}

bool CoMTask::setSetPoint(const Vector3d& p, const Vector3d& v, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    const double rec[6] = {p[0], p[1], p[2], v[0], v[1], v[2]};
    storeSetPoint(rec, instance);
    m_isSetPointSetAtLeastOnce = true;
    return true;
}

bool CoMTask::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[CoMTask::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("[CoMTask::update] The set point has not been set at least once.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        Vector3d com;
        System::MatrixXd J;
        if (!m_kinDyn->getCenterOfMassPosition(com, i) || !m_kinDyn->getCenterOfMassJacobian(J, i)) {
            log()->error("[CoMTask::update] Unable to get the jacobian.");
            return m_isValid;
        }
        const double* sp = m_sp.data() + 6 * i;
        if (m_useExogenousFeedback)   // CoMTask.cpp:192-195: the measured CoM (zero until the first setFeedback)
            for (int c = 0; c < 3; c++) com[c] = m_fb.size() >= 3 * static_cast<std::size_t>(i + 1) ? m_fb[3 * i + c] : 0.0;
        std::size_t idx = 0;
        for (int c = 0; c < 3; c++)
            if (m_mask[c]) {
                m_b[idx] = sp[3 + c] + m_kpLinear[c] * (sp[c] - com[c]);
                setRow(m_A, idx++, m_robotVelocityVariable.offset, J, c);
            }
    }
    m_isValid = true;
    return m_isValid;
}

bool CoMTask::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>&) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_COM, d.frame = -1, d.controller_enabled = 1;
    d.use_position_exogenous_feedback = m_useExogenousFeedback;
    for (int c = 0; c < 3; c++) d.mask[c] = m_mask[c], d.kp_linear[c] = m_kpLinear[c];
    return m_isInitialized;
}

// =================================================================== SO3Task
bool SO3Task::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[SO3Task::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    std::string frameName;
    if (!ptr->getParameter("frame_name", frameName) || (m_frameIndex = m_kinDyn->getFrameIndex(frameName)) < 0) {
        log()->error("{} Error while retrieving the frame that should be controlled.", errorPrefix);
        return false;
    }
    if (!readGain3(*ptr, "kp_angular", m_kpAngular)) {
        log()->error("{} Unable to get the kp_angular gain.", errorPrefix);
        return false;
    }
    m_description = "SO3Task Optimal Control Element - Frame name: " + frameName + ".";
    m_isInitialized = true;
    return true;
}

bool SO3Task::setSetPoint(const SO3d& I_R_F, const SO3Tangentd& w, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    const auto q = I_R_F.quat().coeffs();
    const double rec[7] = {q[0], q[1], q[2], q[3], w.w[0], w.w[1], w.w[2]};
    storeSetPoint(rec, instance);
    m_isSetPointSetAtLeastOnce = true;
    return true;
}

bool SO3Task::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[SO3Task::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("[SO3Task::update] The set point has not been set at least once.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        SE3d pose;
        std::array<double, 9> R;
        System::MatrixXd J;
        if (!m_kinDyn->getWorldTransform(m_frameIndex, pose, R, i) || !m_kinDyn->getFrameFreeFloatingJacobian(m_frameIndex, J, i)) {
            log()->error("[SO3Task::update] Unable to get the jacobian.");
            return m_isValid;
        }
        const double* sp = m_sp.data() + 7 * i;
        const Vector3d w = so3Control(Quaterniond{sp[0], sp[1], sp[2], sp[3]}, pose.R.quat(), {sp[4], sp[5], sp[6]}, m_kpAngular, true);
        for (int c = 0; c < 3; c++) {
            m_b[c] = w[c];
            setRow(m_A, c, m_robotVelocityVariable.offset, J, 3 + c);   // A = J.bottomRows<3>() (SO3Task.cpp:187)
        }
    }
    m_isValid = true;
    return m_isValid;
}

bool SO3Task::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>&) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_SO3, d.frame = m_frameIndex, d.controller_enabled = 1;
    for (int c = 0; c < 3; c++) d.mask[c] = 1, d.kp_angular[c] = m_kpAngular[c];
    return m_isInitialized;
}

// =================================================================== R3Task
bool R3Task::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[R3Task::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    std::string frameName;
    if (!ptr->getParameter("frame_name", frameName) || (m_frameIndex = m_kinDyn->getFrameIndex(frameName)) < 0) {
        log()->error("{} Error while retrieving the frame that should be controlled.", errorPrefix);
        return false;
    }
    if (!readGain3(*ptr, "kp_linear", m_kpLinear)) {
        log()->error("{} Unable to get the kp_linear gain.", errorPrefix);
        return false;
    }
    if (!readMask(*ptr, m_mask, m_DoFs)) {
        log()->error("{} The size of the mask must be 3.", errorPrefix);
        return false;
    }
    m_description = "R3Task Optimal Control Element - Frame name: " + frameName + ".";
    m_isInitialized = true;
    return true;
}

bool R3Task::setSetPoint(const Vector3d& p, const Vector3d& v, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    const double rec[6] = {p[0], p[1], p[2], v[0], v[1], v[2]};
    storeSetPoint(rec, instance);
    m_isSetPointSetAtLeastOnce = true;
    return true;
}

bool R3Task::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[R3Task::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("[R3Task::update] The set-point has not been set at least once.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        SE3d pose;
        std::array<double, 9> R;
        System::MatrixXd J;
        if (!m_kinDyn->getWorldTransform(m_frameIndex, pose, R, i) || !m_kinDyn->getFrameFreeFloatingJacobian(m_frameIndex, J, i)) {
            log()->error("[R3Task::update] Unable to get the jacobian.");
            return m_isValid;
        }
        const double* sp = m_sp.data() + 6 * i;
        std::size_t idx = 0;
        for (int c = 0; c < 3; c++)
            if (m_mask[c]) {
                m_b[idx] = sp[3 + c] + m_kpLinear[c] * (sp[c] - pose.p[c]);
                setRow(m_A, idx++, m_robotVelocityVariable.offset, J, c);
            }
    }
    m_isValid = true;
    return m_isValid;
}

bool R3Task::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>&) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_R3, d.frame = m_frameIndex, d.controller_enabled = 1;
    for (int c = 0; c < 3; c++) d.mask[c] = m_mask[c], d.kp_linear[c] = m_kpLinear[c];
    return m_isInitialized;
}

// =================================================================== DistanceTask
bool DistanceTask::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[DistanceTask::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    if (!ptr->getParameter("kp", m_kp)) {
        log()->error("{} Failed to get the proportional linear gain.", errorPrefix);
        return false;
    }
    std::string referenceName, targetName;
    ptr->getParameter("reference_frame_name", referenceName);   // optional, "" = the world origin (DistanceTask.cpp:131-137)
    if (!ptr->getParameter("target_frame_name", targetName)) {
        log()->error("{} Failed to get the end effector frame name.", errorPrefix);
        return false;
    }
    ptr->getParameter("distance_numeric_threshold", m_distanceNumericThreshold);   // optional (DistanceTask.cpp:149-156)
    m_referenceFrameIndex = -1;
    if (!referenceName.empty() && (m_referenceFrameIndex = m_kinDyn->getFrameIndex(referenceName)) < 0) {
        log()->error("{} The specified base name ({}) does not seem to exist.", errorPrefix, referenceName);
        return false;
    }
    if ((m_targetFrameIndex = m_kinDyn->getFrameIndex(targetName)) < 0) {
        log()->error("{} The specified target name ({}) does not seem to exist.", errorPrefix, targetName);
        return false;
    }
    m_description = "Distance Task - target frame: " + targetName + ".";
    m_isInitialized = true;
    return true;
}

bool DistanceTask::setDesiredDistance(double desiredDistance, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    const double rec[1] = {std::abs(desiredDistance)};   // DistanceTask.cpp:240
    storeSetPoint(rec, instance);
    m_isSetPointSetAtLeastOnce = true;
    return true;
}

bool DistanceTask::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[DistanceTask::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("[DistanceTask::update] The set point has not been set at least once.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        SE3d pose, poseRef;
        std::array<double, 9> R;
        System::MatrixXd J, Jref;
        if (!m_kinDyn->getWorldTransform(m_targetFrameIndex, pose, R, i) || !m_kinDyn->getFrameFreeFloatingJacobian(m_targetFrameIndex, J, i)) {
            log()->error("[DistanceTask::update] Unable to get the jacobian.");
            return m_isValid;
        }
        Vector3d d = pose.p;
        if (m_referenceFrameIndex >= 0) {
            if (!m_kinDyn->getWorldTransform(m_referenceFrameIndex, poseRef, R, i) || !m_kinDyn->getFrameFreeFloatingJacobian(m_referenceFrameIndex, Jref, i)) {
                log()->error("[DistanceTask::update] Unable to get the relative jacobian.");
                return m_isValid;
            }
            for (int c = 0; c < 3; c++) d[c] -= poseRef.p[c];
        }
        m_computedDistance = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        const double den = std::max(m_distanceNumericThreshold, m_computedDistance);
        for (std::size_t c = 0; c < J.cols; c++) {
            double a = 0.0;
            for (int r = 0; r < 3; r++) a += d[r] * (J(r, c) - (m_referenceFrameIndex >= 0 ? Jref(r, c) : 0.0));
            m_A(0, m_robotVelocityVariable.offset + c) = a / den;
        }
        m_b[0] = m_kp * (m_sp[i] - m_computedDistance);
    }
    m_isValid = true;
    return m_isValid;
}

bool DistanceTask::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>&) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_DISTANCE, d.frame = m_targetFrameIndex, d.ref_frame = m_referenceFrameIndex, d.controller_enabled = 1;
    d.distance_threshold = m_distanceNumericThreshold;
    for (int c = 0; c < 3; c++) d.mask[c] = 1, d.kp_linear[c] = m_kp;
    return m_isInitialized;
}

// =================================================================== GravityTask
bool GravityTask::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[GravityTask::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    if (!ptr->getParameter("kp", m_kp)) {
        log()->error("{} Failed to get the proportional linear gain.", errorPrefix);
        return false;
    }
    std::string targetName;
    if (!ptr->getParameter("target_frame_name", targetName) || (m_targetFrameIndex = m_kinDyn->getFrameIndex(targetName)) < 0) {
        log()->error("{} Failed to get the target frame name.", errorPrefix);
        return false;
    }
    m_description = "Gravity Task - target frame: " + targetName + ".";
    m_isInitialized = true;
    return true;
}

void GravityTask::storePart(const Vector3d& v, int first, std::ptrdiff_t instance)
{
    const std::size_t N = m_kinDyn->getBatchSize();
    if (m_sp.size() != 6 * N) {
        m_sp.assign(6 * N, 0.0);
        for (std::size_t i = 0; i < N; i++) m_sp[6 * i + 2] = 1.0;   // m_desiredZDirectionBody starts at e_z (GravityTask.cpp:71)
    }
    for (std::size_t i = 0; i < N; i++)
        if (instance < 0 || static_cast<std::size_t>(instance) == i)
            for (int c = 0; c < 3; c++) m_sp[6 * i + first + c] = v[c];
    m_spVersion++;
    m_isSetPointSetAtLeastOnce = true;
}

bool GravityTask::setDesiredGravityDirectionInTargetFrame(const Vector3d& g, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    const double nrm = std::sqrt(g[0] * g[0] + g[1] * g[1] + g[2] * g[2]);   // normalised as GravityTask.cpp:204-206
    storePart(nrm > 0 ? Vector3d{g[0] / nrm, g[1] / nrm, g[2] / nrm} : g, 0, instance);
    return true;
}

bool GravityTask::setFeedForwardVelocityInTargetFrame(const Vector3d& ff, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    storePart(ff, 3, instance);
    return true;
}

bool GravityTask::setSetPoint(const Vector3d& g, const Vector3d& ff, std::ptrdiff_t instance)
{
    return setDesiredGravityDirectionInTargetFrame(g, instance) && setFeedForwardVelocityInTargetFrame(ff, instance);
}

bool GravityTask::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[GravityTask::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("[GravityTask::update] The set point has not been set at least once.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        SE3d pose;
        std::array<double, 9> R;
        System::MatrixXd J;
        if (!m_kinDyn->getWorldTransform(m_targetFrameIndex, pose, R, i) || !m_kinDyn->getFrameFreeFloatingJacobian(m_targetFrameIndex, J, i)) {
            log()->error("[GravityTask::update] Unable to get the frame jacobian.");
            return m_isValid;
        }
        const double* sp = m_sp.data() + 6 * i;
        double z[3], ff[3];
        for (int r = 0; r < 3; r++) {
            z[r] = R[3 * r] * sp[0] + R[3 * r + 1] * sp[1] + R[3 * r + 2] * sp[2];
            ff[r] = R[3 * r] * sp[3] + R[3 * r + 1] * sp[4] + R[3 * r + 2] * sp[5];
        }
        // A = [1 0 0; 0 1 0] J_angular, b = ff_xy - kp [0 -1 0; 1 0 0] z  (GravityTask.cpp:56-62, 182-185)
        setRow(m_A, 0, m_robotVelocityVariable.offset, J, 3);
        setRow(m_A, 1, m_robotVelocityVariable.offset, J, 4);
        m_b[0] = ff[0] + m_kp * z[1];
        m_b[1] = ff[1] - m_kp * z[0];
    }
    m_isValid = true;
    return m_isValid;
}

bool GravityTask::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>&) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_GRAVITY, d.frame = m_targetFrameIndex, d.ref_frame = -1, d.controller_enabled = 1;
    for (int c = 0; c < 3; c++) d.mask[c] = 1, d.kp_linear[c] = m_kp;
    return m_isInitialized;
}

// =================================================================== AngularMomentumTask
bool AngularMomentumTask::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    std::vector<bool> mask;
    if (ptr->getParameter("mask", mask) && mask.size() == 3) {   // otherwise the default (all true) is used, :111-131
        m_mask = {mask[0], mask[1], mask[2]};
        m_DoFs = static_cast<std::size_t>(m_mask[0]) + m_mask[1] + m_mask[2];
    } else {
        log()->info("[AngularMomentumTask::initialize] Unable to find the mask parameter. The default value is used.");
    }
    m_description = "Angular momentum task.";
    m_isInitialized = true;
    return true;
}

bool AngularMomentumTask::setSetPoint(const Vector3d& L, std::ptrdiff_t instance)
{
    if (!m_kinDyn) return false;
    const double rec[3] = {L[0], L[1], L[2]};
    storeSetPoint(rec, instance);
    m_isSetPointSetAtLeastOnce = true;
    return true;
}

bool AngularMomentumTask::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[AngularMomentumTask::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("[AngularMomentumTask::update] The set point has not been set at least once.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_b.size() == size()) {   // b = masked desired momentum (:187-210)
        std::size_t idx = 0;
        for (int c = 0; c < 3; c++)
            if (m_mask[c]) m_b[idx++] = m_sp[3 * i + c];
    }
    m_isValid = true;
    return m_isValid;
}

bool AngularMomentumTask::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>&) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_ANGULAR_MOMENTUM, d.frame = -1, d.ref_frame = -1, d.controller_enabled = 1;
    for (int c = 0; c < 3; c++) d.mask[c] = m_mask[c];
    return m_isInitialized;
}

// =================================================================== JointTrackingTask
bool JointTrackingTask::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[JointTrackingTask::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    if (!ptr->getParameter("kp", m_kp)) {
        log()->error("{} Error while retrieving the proportional gain.", errorPrefix);
        return false;
    }
    if (m_kp.size() != m_kinDyn->getNrOfDegreesOfFreedom()) {  // JointTrackingTask.cpp:100-110
        log()->error("{} The size of the kp gain does not match with the number of actuated joints. Expected: {}. Given: {}.", errorPrefix,
                     m_kinDyn->getNrOfDegreesOfFreedom(), m_kp.size());
        return false;
    }
    m_description = "Joint tracking task";
    m_isInitialized = true;
    return true;
}

std::size_t JointTrackingTask::size() const
{
    if (!m_kinDyn) {
        log()->warn("[JointTrackingTask::size] Please call setKinDyn method before. A size equal to zero will be returned.");
        return 0;
    }
    return m_kinDyn->getNrOfDegreesOfFreedom();
}

bool JointTrackingTask::setSetPoint(const VectorXd& jointPosition, std::ptrdiff_t instance)
{
    return setSetPoint(jointPosition, VectorXd(jointPosition.size(), 0.0), instance);
}

bool JointTrackingTask::setSetPoint(const VectorXd& jointPosition, const VectorXd& jointVelocity, std::ptrdiff_t instance)
{
    constexpr auto errorPrefix = "[JointTrackingTask::setSetPoint]";
    if (!m_kinDyn) return false;
    const std::size_t n = m_kinDyn->getNrOfDegreesOfFreedom();
    if (jointPosition.size() != n || jointVelocity.size() != n) {
        log()->error("{} Wrong size of the desired reference setpoint. Expected size: {}. Given joint position size: {}, given joint velocity size: {}.",
                     errorPrefix, n, jointPosition.size(), jointVelocity.size());
        return false;
    }
    std::vector<double> rec(2 * n);
    std::copy(jointPosition.begin(), jointPosition.end(), rec.begin());
    std::copy(jointVelocity.begin(), jointVelocity.end(), rec.begin() + n);
    storeSetPoint(rec.data(), instance);
    m_isSetPointSetAtLeastOnce = true;
    return true;
}

bool JointTrackingTask::update()
{
    constexpr auto errorPrefix = "[JointTrackingTask::update]";
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("{} The task is not initialized. Please call initialize method.", errorPrefix);
        return m_isValid;
    }
    if (!m_isSetPointSetAtLeastOnce) {
        log()->error("{} The set point has not been set at least once.", errorPrefix);
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        const std::size_t n = size();
        VectorXd s;
        m_kinDyn->getJointPos(s, i);
        const double* sp = m_sp.data() + 2 * n * i;
        for (std::size_t j = 0; j < n; j++) {
            m_A(j, m_robotVelocityVariable.offset + 6 + j) = 1.0;   // A = [0 I] (JointTrackingTask.cpp:65-67)
            m_b[j] = sp[n + j] + m_kp[j] * (sp[j] - s[j]);          // :151
        }
    }
    m_isValid = true;
    return m_isValid;
}

bool JointTrackingTask::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_JOINT_TRACKING, d.frame = -1, d.controller_enabled = 1;
    d.kp_joint = keepVector(keep, m_kp);
    return m_isInitialized;
}

// =================================================================== JointLimitsTask
bool JointLimitsTask::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[JointLimitsTask::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    const std::size_t n = m_kinDyn->getNrOfDegreesOfFreedom();
    if (!ptr->getParameter("klim", m_klim) || m_klim.size() != n) {
        log()->error("{} Unable to get the klim parameter or its size is different from the number of DoFs.", errorPrefix);
        return false;
    }
    for (double k : m_klim)
        if (k < 0 || k > 1) {  // JointLimitsTask.cpp:150-166
            log()->error("{} The klim parameter must be a vector whose elements belong to [0, 1].", errorPrefix);
            return false;
        }
    if (!ptr->getParameter("sampling_time", m_samplingTime) || m_samplingTime <= std::chrono::nanoseconds::zero()) {
        log()->error("{} Unable to get a strictly positive sampling_time.", errorPrefix);
        return false;
    }
    bool useModelLimits = false;
    if (!ptr->getParameter("use_model_limits", useModelLimits)) {   // mandatory (JointLimitsTask.cpp:178-184)
        log()->error("{} Unable to get the use_model_limits parameter.", errorPrefix);
        return false;
    }
    if (useModelLimits) {
        m_upperLimits = m_kinDyn->jointUpperLimits(), m_lowerLimits = m_kinDyn->jointLowerLimits();
    } else if (!ptr->getParameter("upper_limits", m_upperLimits) || !ptr->getParameter("lower_limits", m_lowerLimits) || m_upperLimits.size() != n
               || m_lowerLimits.size() != n) {
        log()->error("{} Unable to get upper_limits / lower_limits of size {}.", errorPrefix, n);
        return false;
    }
    for (std::size_t j = 0; j < n; j++)
        if (m_upperLimits[j] >= BLFIK_INFTY || m_lowerLimits[j] <= -BLFIK_INFTY) {
            // the reference's partial-limits branch (JointLimitsTask.cpp:232-255) is not reproduced
            log()->error("{} Every joint must have finite limits.", errorPrefix);
            return false;
        }
    m_description = "Joint limits task";
    m_isInitialized = true;
    return true;
}

std::size_t JointLimitsTask::size() const { return m_kinDyn ? 2 * m_kinDyn->getNrOfDegreesOfFreedom() : 0; }

bool JointLimitsTask::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[JointLimitsTask::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        const std::size_t n = size() / 2;
        const double dT = std::chrono::duration<double>(m_samplingTime).count();
        VectorXd s;
        m_kinDyn->getJointPos(s, i);
        for (std::size_t j = 0; j < n; j++) {
            m_A(j, m_robotVelocityVariable.offset + 6 + j) = 1.0;        // [0 I; 0 -I] (JointLimitsTask.cpp:110-113)
            m_A(n + j, m_robotVelocityVariable.offset + 6 + j) = -1.0;
            m_b[j] = m_klim[j] * (m_upperLimits[j] - s[j]) / dT;         // :227-231
            m_b[n + j] = m_klim[j] * (-m_lowerLimits[j] + s[j]) / dT;
        }
    }
    m_isValid = true;
    return m_isValid;
}

bool JointLimitsTask::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_JOINT_LIMITS, d.frame = -1, d.controller_enabled = 1;
    d.klim = keepVector(keep, m_klim), d.upper_limits = keepVector(keep, m_upperLimits), d.lower_limits = keepVector(keep, m_lowerLimits);
    d.sampling_time = std::chrono::duration<double>(m_samplingTime).count();
    return m_isInitialized;
}

// =================================================================== JointVelocityLimitsTask
bool JointVelocityLimitsTask::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> paramHandler)
{
    constexpr auto errorPrefix = "[JointVelocityLimitsTask::initialize]";
    std::shared_ptr<const ParametersHandler::IParametersHandler> ptr;
    if (!commonInitialize(paramHandler, ptr)) return false;
    const std::size_t n = m_kinDyn->getNrOfDegreesOfFreedom();
    if (!ptr->getParameter("upper_limits", m_upperLimits) || !ptr->getParameter("lower_limits", m_lowerLimits) || m_upperLimits.size() != n
        || m_lowerLimits.size() != n) {
        log()->error("{} Unable to get upper_limits / lower_limits of size {}.", errorPrefix, n);
        return false;
    }
    m_description = "Joint velocity limits task";
    m_isInitialized = true;
    return true;
}

std::size_t JointVelocityLimitsTask::size() const { return m_kinDyn ? 2 * m_kinDyn->getNrOfDegreesOfFreedom() : 0; }

bool JointVelocityLimitsTask::update()
{
    m_isValid = false;
    if (!m_isInitialized) {
        log()->error("[JointVelocityLimitsTask::update] The task is not initialized. Please call initialize method.");
        return m_isValid;
    }
    const int i = inspected();
    if (i >= 0 && m_A.rows == size()) {
        const std::size_t n = size() / 2;
        for (std::size_t j = 0; j < n; j++) {
            m_A(j, m_robotVelocityVariable.offset + 6 + j) = 1.0;
            m_A(n + j, m_robotVelocityVariable.offset + 6 + j) = -1.0;
            m_b[j] = m_upperLimits[j];            // JointVelocityLimitsTask.cpp:140-141
            m_b[n + j] = -m_lowerLimits[j];
        }
    }
    m_isValid = true;
    return m_isValid;
}

bool JointVelocityLimitsTask::fillDescriptor(blfik_task_desc& d, std::vector<std::vector<double>>& keep) const
{
    d = blfik_task_desc{};
    d.type = BLFIK_TASK_JOINT_VELOCITY_LIMITS, d.frame = -1, d.controller_enabled = 1;
    d.upper_limits = keepVector(keep, m_upperLimits), d.lower_limits = keepVector(keep, m_lowerLimits);
    return m_isInitialized;
}

// =================================================================== QPInverseKinematics
struct QPInverseKinematics::Impl {
    struct TaskEntry {
        std::shared_ptr<IKLinearTask> task;
        std::string name;
        std::size_t priority{0};
        std::shared_ptr<const System::WeightProviderPort> weightProvider;
        VectorXd uploadedWeight;
        unsigned long uploadedSetPoint{~0ul}, uploadedGains{~0ul}, uploadedFeedback{0}, uploadedSpline{0}, uploadedMinJerk{0};
    };
    std::vector<TaskEntry> tasks;   // insertion order = row order of the constraint matrix (QPInverseKinematics.cpp:201,236)
    std::unordered_map<std::string, std::size_t> index;
    std::shared_ptr<KinDynBatch> kinDyn;
    System::VariablesHandler::VariableDescription robotVelocityVariable;
    std::size_t nVars{0};
    bool isInitialized{false}, isFinalized{false}, isValid{false}, isVerbose{false};
    blfik_problem* problem{nullptr};   // the first shard's problem (the only one unless setDevices spread the batch)
    State solution;
    VectorXd rawSolution;
    std::vector<double> nu;
    std::vector<int32_t> status;
    // ---- the batch path: contiguous instance blocks, one problem per device, page-locked packed records as the native storage
    // of state and set-points (KinDynBatch and the tasks write straight into them), one blfik_tick_packed per device and tick
    struct Shard {
        int device{0};
        std::size_t first{0}, count{0};
        blfik_problem* prob{nullptr};
    };
    static constexpr std::size_t kPackedMin = 64;   // smaller batches keep the per-array calls (lower latency for a handful of robots)
    std::vector<int> devices;
    std::vector<Shard> shards;
    bool compact{false}, packedPath{false}, warmStart{true}, strict{false};
    double* hRec{nullptr};
    double* hNu{nullptr};
    int32_t* hStatus{nullptr};
    std::size_t recLen{0}, recCap{0};
    std::vector<char> sharedNow;
    std::vector<unsigned long> sharedVersion;
    bool nuStale{false}, statusStale{false};
    bool deviceStateAhead{false};
    unsigned long stateVersionAtTick{0};
    int64_t unsolved{0};

    uint32_t recordFlags() const { return compact ? BLFIK_TICK_NO_FEEDFORWARD : 0u; }
    void destroyProblems()
    {
        for (auto& sh : shards)
            if (sh.prob) blfik_problem_destroy(sh.prob);
        shards.clear();
        problem = nullptr;
    }
    void freeHost()
    {
        if (kinDyn) kinDyn->bindRecords(nullptr, 0);
        for (auto& e : tasks) e.task->bindRecords(nullptr, 0, 0, 0);
        if (hRec) blfik_host_free(hRec);
        if (hNu) blfik_host_free(hNu);
        if (hStatus) blfik_host_free(hStatus);
        hRec = hNu = nullptr, hStatus = nullptr;
    }
    /** which set-points travel as one batch-shared record, where everything else sits in a packed record; binds KinDynBatch
     *  and the tasks to the records */
    bool layoutRecords()
    {
        const std::size_t N = kinDyn->getBatchSize();
        for (std::size_t t = 0; t < tasks.size(); t++) {
            auto& task = *tasks[t].task;
            const bool want = task.setPointLength() > 0 && task.feedbackLength() == 0 && task.setPointIsUniform();
            if (want == static_cast<bool>(sharedNow[t]) && (!want || sharedVersion[t] == task.setPointVersion())) continue;
            for (auto& sh : shards) {
                const int rc = want ? blfik_set_setpoint_shared(sh.prob, static_cast<int32_t>(t), task.setPointBuffer().data(), BLFIK_MEM_HOST)
                                    : blfik_set_setpoint(sh.prob, static_cast<int32_t>(t), task.setPointBuffer().data() + sh.first * task.setPointLength(), BLFIK_MEM_HOST);
                if (rc != BLFIK_OK) return false;
            }
            sharedNow[t] = want, sharedVersion[t] = task.setPointVersion();
        }
        const std::size_t len = static_cast<std::size_t>(blfik_packed_record_len_ex(shards[0].prob, recordFlags()));
        if (len == recLen && hRec) return true;
        recLen = len;
        if (recLen > recCap) return false;
        kinDyn->bindRecords(hRec, recLen);
        for (std::size_t t = 0; t < tasks.size(); t++) {
            auto& task = *tasks[t].task;
            const int off = blfik_packed_offset_ex(shards[0].prob, static_cast<int32_t>(t), recordFlags());
            if (off < 0) {
                task.bindRecords(nullptr, 0, 0, 0);
                continue;
            }
            const std::size_t part = static_cast<std::size_t>(blfik_packed_part_len(shards[0].prob, static_cast<int32_t>(t), recordFlags()));
            task.bindRecords(hRec, recLen, static_cast<std::size_t>(off), part - task.feedbackLength());
        }
        (void)N;
        return true;
    }

    ~Impl()
    {
        freeHost();
        destroyProblems();
    }
};

QPInverseKinematics::QPInverseKinematics() : m_pimpl(std::make_unique<Impl>()) {}
QPInverseKinematics::~QPInverseKinematics() = default;

bool QPInverseKinematics::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler)
{
    constexpr auto logPrefix = "[QPInverseKinematics::initialize]";
    auto ptr = handler.lock();
    if (ptr == nullptr) {
        log()->error("{} The parameter handler is not valid.", logPrefix);
        return false;
    }
    if (!ptr->getParameter("robot_velocity_variable_name", m_pimpl->robotVelocityVariable.name)) {   // QPInverseKinematics.cpp:586-591
        log()->error("{} Unable to get the parameter 'robot_velocity_variable_name'.", logPrefix);
        return false;
    }
    // batch extensions, both optional: the GPUs the batch is spread over and the compact host records (no reference counterpart)
    std::vector<int> devs;
    if (ptr->getParameter("cuda_devices", devs)) m_pimpl->devices = devs;
    bool compactRecords = false;
    if (ptr->getParameter("compact_records", compactRecords)) m_pimpl->compact = compactRecords;
    if (!ptr->getParameter("verbosity", m_pimpl->isVerbose))
        log()->info("{} 'verbosity' not found. The following parameter will be used '{}'.", logPrefix, m_pimpl->isVerbose);
    m_pimpl->isInitialized = true;
    return true;
}

bool QPInverseKinematics::addTask(std::shared_ptr<Task> task, const std::string& taskName, std::size_t priority,
                                  std::shared_ptr<const System::WeightProviderPort> weightProvider)
{
    constexpr auto logPrefix = "[QPInverseKinematics::addTask]";
    if (!m_pimpl->isInitialized) {   // :173-177
        log()->error("{} Please call initialize() before addTask().", logPrefix);
        return false;
    }
    if (!task) {
        log()->error("{} The task named {} is not valid.", logPrefix, taskName);
        return false;
    }
    if (m_pimpl->index.count(taskName)) {   // :180-184
        log()->error("{} The task named {} already exist.", logPrefix, taskName);
        return false;
    }
    if (priority != 0 && priority != 1) {   // :186-194
        log()->error("{} - [Task name: '{}'] For the time being we support only priority equal to 0 or 1.", logPrefix, taskName);
        return false;
    }
    if (task->type() == System::LinearTask::Type::inequality && priority == 1) {   // :196-203
        log()->error("{} - [Task name: '{}'] This implementation of the inverse kinematics cannot handle inequality tasks with priority equal to 1.",
                     logPrefix, taskName);
        return false;
    }
    if (priority == 1) {
        if (weightProvider == nullptr) {   // :208-214
            log()->error("{} - [Task name: '{}'] Invalid weight provider. It is required by a task with priority equal to 1.", logPrefix, taskName);
            return false;
        }
        if (weightProvider->getOutput().size() != task->size()) {   // :216-229
            log()->error("{} - [Task name: '{}'] The size of the weight is not coherent with the size of the task. Expected: {}. Given: {}.", logPrefix,
                         taskName, task->size(), weightProvider->getOutput().size());
            return false;
        }
    }
    Impl::TaskEntry e;
    e.task = task, e.name = taskName, e.priority = priority, e.weightProvider = priority == 1 ? weightProvider : nullptr;
    if (!m_pimpl->kinDyn) m_pimpl->kinDyn = task->kinDyn();   // the one KinDyn shared by every task (IKLinearTask.h:40)
    m_pimpl->index[taskName] = m_pimpl->tasks.size();
    m_pimpl->tasks.push_back(std::move(e));
    m_pimpl->isFinalized = false;   // :245-246
    return true;
}

bool QPInverseKinematics::addTask(std::shared_ptr<Task> task, const std::string& taskName, std::size_t priority, const VectorXd& weight)
{
    return addTask(task, taskName, priority, std::make_shared<System::ConstantWeightProvider>(weight));
}

bool QPInverseKinematics::setTaskWeight(const std::string& taskName, std::shared_ptr<const System::WeightProviderPort> weightProvider)
{
    constexpr auto logPrefix = "[QPInverseKinematics::setTaskWeight]";
    auto it = m_pimpl->index.find(taskName);
    if (it == m_pimpl->index.end()) {   // :259-264
        log()->error("{} The task named {} does not exist.", logPrefix, taskName);
        return false;
    }
    auto& e = m_pimpl->tasks[it->second];
    if (e.priority != 1) {   // :268-276
        log()->error("{} - [Task name: '{}'] The weight can be set only to a task with priority equal to 1.", logPrefix, taskName);
        return false;
    }
    if (weightProvider == nullptr) {
        log()->error("{} - [Task name: '{}'] The weight provider is not valid.", logPrefix, taskName);
        return false;
    }
    if (weightProvider->getOutput().size() != e.task->size()) {   // :284-297
        log()->error("{} - [Task name: '{}'] The size of the weight is not coherent with the size of the task. Expected: {}. Given: {}.", logPrefix,
                     taskName, e.task->size(), weightProvider->getOutput().size());
        return false;
    }
    e.weightProvider = weightProvider;   // takes effect on the next advance() (:450-451)
    return true;
}

bool QPInverseKinematics::setTaskWeight(const std::string& taskName, const VectorXd& weight)
{
    return setTaskWeight(taskName, std::make_shared<System::ConstantWeightProvider>(weight));
}

std::weak_ptr<const System::WeightProviderPort> QPInverseKinematics::getTaskWeightProvider(const std::string& taskName) const
{
    auto it = m_pimpl->index.find(taskName);
    if (it == m_pimpl->index.end() || m_pimpl->tasks[it->second].priority != 1) {
        log()->error("[QPInverseKinematics::getTaskWeightProvider] The task named {} does not exist or is a hard constraint.", taskName);
        return std::weak_ptr<const System::WeightProviderPort>();
    }
    return m_pimpl->tasks[it->second].weightProvider;
}

bool QPInverseKinematics::finalize(const System::VariablesHandler& handler)
{
    constexpr auto logPrefix = "[QPInverseKinematics::finalize]";
    if (!m_pimpl->isInitialized) {   // :327-331
        log()->error("{} Please call initialize() before finalize().", logPrefix);
        return false;
    }
    if (m_pimpl->tasks.empty()) {
        log()->error("{} Please add at least one task.", logPrefix);
        return false;
    }
    if (!handler.getVariable(m_pimpl->robotVelocityVariable.name, m_pimpl->robotVelocityVariable)) {
        log()->error("{} Unable to find the variable named {} in the handler.", logPrefix, m_pimpl->robotVelocityVariable.name);
        return false;
    }
    m_pimpl->nVars = handler.getNumberOfVariables();
    for (auto& e : m_pimpl->tasks) {   // :343-360
        if (!e.task->setVariablesHandler(handler)) {
            log()->error("{} Error while setting the variable handler in the task named {}.", logPrefix, e.name);
            return false;
        }
    }
    // one KinDyn is shared by all the tasks (IKLinearTask.h:40); it also fixes the batch size
    std::vector<blfik_task_desc> descs(m_pimpl->tasks.size());
    std::vector<std::vector<double>> keep;
    keep.reserve(8 * m_pimpl->tasks.size());
    for (std::size_t t = 0; t < m_pimpl->tasks.size(); t++) {
        auto& e = m_pimpl->tasks[t];
        if (!e.task->fillDescriptor(descs[t], keep)) {
            log()->error("{} The task named {} is not initialized.", logPrefix, e.name);
            return false;
        }
        descs[t].priority = static_cast<int32_t>(e.priority);
        if (e.priority == 1) {
            e.uploadedWeight = e.weightProvider->getOutput();
            descs[t].weight = keepVector(keep, e.uploadedWeight);
        }
        e.uploadedSetPoint = ~0ul, e.uploadedGains = e.task->gainsVersion(), e.uploadedFeedback = 0, e.uploadedSpline = 0, e.uploadedMinJerk = 0;
    }
    if (!m_pimpl->kinDyn) {
        log()->error("{} No KinDyn object: build the solver with QPInverseKinematics::build or call setKinDyn through a task added by addTask.", logPrefix);
        return false;
    }
    m_pimpl->freeHost();
    m_pimpl->destroyProblems();
    const std::size_t N = m_pimpl->kinDyn->getBatchSize();
    {
        std::vector<int> devs = m_pimpl->devices;
        if (devs.empty() || N < Impl::kPackedMin) devs = {m_pimpl->kinDyn->getDevice()};
        if (devs.size() > N) devs.resize(N);
        for (std::size_t d = 0; d < devs.size(); d++) {
            Impl::Shard sh;
            sh.device = devs[d], sh.first = N * d / devs.size(), sh.count = N * (d + 1) / devs.size() - sh.first;
            if (blfik_problem_create(m_pimpl->kinDyn->model(), descs.data(), static_cast<int32_t>(descs.size()), static_cast<int64_t>(sh.count), sh.device,
                                     &sh.prob) != BLFIK_OK) {
                log()->error("{} {}", logPrefix, blfik_last_error());
                m_pimpl->destroyProblems();
                return false;
            }
            m_pimpl->shards.push_back(sh);
        }
        m_pimpl->problem = m_pimpl->shards[0].prob;
    }
    m_pimpl->packedPath = N >= Impl::kPackedMin;
    m_pimpl->sharedNow.assign(m_pimpl->tasks.size(), 0), m_pimpl->sharedVersion.assign(m_pimpl->tasks.size(), 0);
    m_pimpl->recLen = 0, m_pimpl->deviceStateAhead = false;
    if (m_pimpl->packedPath) {
        const std::size_t nvv = m_pimpl->kinDyn->getNrOfDegreesOfFreedom() + 6;
        m_pimpl->recCap = static_cast<std::size_t>(blfik_packed_record_len_ex(m_pimpl->problem, 0u));   // the full layout is the longest
        if (blfik_host_alloc(N * m_pimpl->recCap * sizeof(double), reinterpret_cast<void**>(&m_pimpl->hRec)) != BLFIK_OK ||
            blfik_host_alloc(N * nvv * sizeof(double), reinterpret_cast<void**>(&m_pimpl->hNu)) != BLFIK_OK ||
            blfik_host_alloc(N * sizeof(int32_t), reinterpret_cast<void**>(&m_pimpl->hStatus)) != BLFIK_OK) {
            log()->error("{} {}", logPrefix, blfik_last_error());
            return false;
        }
        std::fill(m_pimpl->hRec, m_pimpl->hRec + N * m_pimpl->recCap, 0.0);
        if (!m_pimpl->layoutRecords()) {
            log()->error("{} Unable to lay out the packed records: {}", logPrefix, blfik_last_error());
            return false;
        }
    }
    const std::size_t nv = m_pimpl->kinDyn->getNrOfDegreesOfFreedom() + 6;
    m_pimpl->nu.assign(N * nv, 0.0), m_pimpl->status.assign(N, 2);
    m_pimpl->solution.jointVelocity.assign(nv - 6, 0.0);
    m_pimpl->rawSolution.assign(m_pimpl->nVars, 0.0);
    m_pimpl->isFinalized = true;
    return true;
}

std::vector<std::string> QPInverseKinematics::getTaskNames() const
{
    std::vector<std::string> names;
    for (const auto& e : m_pimpl->tasks) names.push_back(e.name);
    return names;
}

bool QPInverseKinematics::isOutputValid() const { return m_pimpl->isValid; }

bool QPInverseKinematics::isOutputValid(std::size_t instance) const
{
    if (!m_pimpl->isValid && m_pimpl->status.size() == 1) return false;
    const int32_t* st = m_pimpl->packedPath ? m_pimpl->hStatus : m_pimpl->status.data();
    const std::size_t N = m_pimpl->kinDyn ? m_pimpl->kinDyn->getBatchSize() : 0;
    return st && instance < N && st[instance] != BLFIK_STATUS_FAILED;
}

bool QPInverseKinematics::setDevices(const std::vector<int>& devices)
{
    if (m_pimpl->isFinalized) {
        log()->error("[QPInverseKinematics::setDevices] Please call setDevices() before finalize().");
        return false;
    }
    m_pimpl->devices = devices;
    return true;
}

void QPInverseKinematics::setCompactRecords(bool on) { m_pimpl->compact = on; }
void QPInverseKinematics::setWarmStart(bool on) { m_pimpl->warmStart = on; }
void QPInverseKinematics::setStrictFailure(bool on) { m_pimpl->strict = on; }

bool QPInverseKinematics::advance()
{
    return tick(0.0, (m_pimpl->warmStart ? BLFIK_TICK_WARM_START : 0u) | (m_pimpl->strict ? BLFIK_TICK_STRICT : 0u));
}

bool QPInverseKinematics::syncState()
{
    Impl* pimpl = m_pimpl.get();
    if (!pimpl->isFinalized) return false;
    auto& kd = *pimpl->kinDyn;
    std::vector<double> p(kd.basePositions().size()), q(kd.baseQuaternions().size()), s(kd.jointPositions().size());
    const std::size_t n = kd.getNrOfDegreesOfFreedom();
    for (auto& sh : pimpl->shards)
        if (blfik_get_state(sh.prob, p.data() + 3 * sh.first, q.data() + 4 * sh.first, s.data() + n * sh.first, BLFIK_MEM_HOST) != BLFIK_OK) {
            log()->error("[QPInverseKinematics::syncState] {}", blfik_last_error());
            return false;
        }
    if (!kd.setRobotStateBatch(p.data(), q.data(), s.data())) return false;
    pimpl->deviceStateAhead = false;
    return true;
}

bool QPInverseKinematics::advanceAndIntegrate(std::chrono::nanoseconds dT, bool syncHostState)
{
    const double dt = std::chrono::duration<double>(dT).count();
    if (!tick(dt, (m_pimpl->warmStart ? BLFIK_TICK_WARM_START : 0u) | (m_pimpl->strict ? BLFIK_TICK_STRICT : 0u) | BLFIK_TICK_INTEGRATE)) return false;
    if (syncHostState) return syncState();
    m_pimpl->deviceStateAhead = true;
    m_pimpl->stateVersionAtTick = m_pimpl->kinDyn->stateVersion();
    return true;
}

bool QPInverseKinematics::tick(double dt, uint32_t flags)
{
    constexpr auto logPrefix = "[QPInverseKinematics::advance]";
    Impl* pimpl = m_pimpl.get();
    pimpl->isValid = false;   // when advance is called the previous solution is no more valid (:414)
    if (!pimpl->isInitialized) {
        log()->error("{} Please call initialize() before advance().", logPrefix);
        return false;
    }
    if (!pimpl->isFinalized) {
        log()->error("{} Please call finalize() before advance().", logPrefix);
        return false;
    }
    for (auto& e : pimpl->tasks) {   // :429-439
        if (!e.task->update()) {
            log()->error("{} Unable to update the task named {}.", logPrefix, e.name);
            return false;
        }
    }
    auto fail = [&](const char* what) {
        log()->error("{} {}: {}", logPrefix, what, blfik_last_error());
        return false;
    };
    auto& kd = *pimpl->kinDyn;
    const std::size_t n = kd.getNrOfDegreesOfFreedom(), nv = n + 6, N = kd.getBatchSize();
    // the state the device integrated itself stays in force until the caller sets a new one
    const bool keepState = pimpl->deviceStateAhead && kd.stateVersion() == pimpl->stateVersionAtTick;
    for (auto& sh : pimpl->shards) {
        blfik_problem* prob = sh.prob;
        for (std::size_t t = 0; t < pimpl->tasks.size(); t++) {
            auto& e = pimpl->tasks[t];
            const bool last = &sh == &pimpl->shards.back();
            if (e.priority == 1) {   // the weight is read from its provider on every tick (:450-451)
                const VectorXd& w = e.weightProvider->getOutput();
                if (w.size() != e.task->size()) {
                    log()->error("{} The weight of the task named {} has the wrong size.", logPrefix, e.name);
                    return false;
                }
                if (w != e.uploadedWeight) {
                    if (blfik_set_weight(prob, static_cast<int32_t>(t), w.data()) != BLFIK_OK) return fail("Unable to update the weight");
                    if (last) e.uploadedWeight = w;
                }
            }
            if (e.task->gainsVersion() != e.uploadedGains) {
                blfik_task_desc d;
                std::vector<std::vector<double>> keep;
                e.task->fillDescriptor(d, keep);
                if (blfik_set_controller_mode(prob, static_cast<int32_t>(t), d.controller_enabled) != BLFIK_OK) return fail("Unable to update the controller mode");
                if (last) e.uploadedGains = e.task->gainsVersion();
            }
            const std::size_t spl = e.task->setPointLength();
            if (!pimpl->packedPath && spl > 0 && e.task->setPointVersion() != e.uploadedSetPoint) {
                if (blfik_set_setpoint(prob, static_cast<int32_t>(t), e.task->setPointBuffer().data(), BLFIK_MEM_HOST) != BLFIK_OK) return fail("Unable to set the set-point");
                e.uploadedSetPoint = e.task->setPointVersion();
            }
            // A generator is attached to a stored set-point.  On the packed path the set-points live in the records and the device
            // copy was never written: it is seeded once, so that the attach calls below find it (the ticks keep reading the records).
            if (pimpl->packedPath && spl > 0 && !(t < pimpl->sharedNow.size() && pimpl->sharedNow[t]) &&
                ((e.task->acceptsMinimumJerk() && e.task->minimumJerkVersion() != e.uploadedMinJerk && e.task->minimumJerkDuration() > 0) ||
                 (e.task->acceptsSetPointSpline() && e.task->splineVersion() != e.uploadedSpline && !e.task->splineTimes().empty())) &&
                blfik_set_setpoint(prob, static_cast<int32_t>(t), e.task->setPointBuffer().data() + sh.first * spl, BLFIK_MEM_HOST) != BLFIK_OK)
                return fail("Unable to set the set-point");
            if (e.task->acceptsMinimumJerk() && e.task->minimumJerkVersion() != e.uploadedMinJerk) {   // setSetPointMinimumJerk(...) since the last tick
                const double T = e.task->minimumJerkDuration();
                if (blfik_set_setpoint_minimum_jerk(prob, static_cast<int32_t>(t), T > 0 ? e.task->minimumJerkTarget().data() + sh.first * spl : nullptr, T, BLFIK_MEM_HOST) != BLFIK_OK)
                    return fail("Unable to set the minimum-jerk generator");
                if (last) e.uploadedMinJerk = e.task->minimumJerkVersion();
            }
            if (e.task->acceptsSetPointSpline() && e.task->splineVersion() != e.uploadedSpline) {   // setSetPointSpline(...) since the last tick
                const auto& tm = e.task->splineTimes();
                const auto set = e.task->splineIsCubic() ? blfik_set_setpoint_cubic_spline : blfik_set_setpoint_spline;
                const std::size_t dim = e.task->splineDimension(), nbc = (e.task->splineIsCubic() ? 2 : 4) * dim;
                const int rc = tm.empty() ? set(prob, static_cast<int32_t>(t), 0, nullptr, nullptr, nullptr, BLFIK_MEM_HOST)
                                          : set(prob, static_cast<int32_t>(t), static_cast<int32_t>(tm.size()), tm.data(),
                                                e.task->splinePositions().data() + sh.first * tm.size() * dim, e.task->splineBoundary().data() + sh.first * nbc, BLFIK_MEM_HOST);
                if (rc != BLFIK_OK) return fail("Unable to set the set-point spline");
                if (last) e.uploadedSpline = e.task->splineVersion();
            }
            if (!pimpl->packedPath && e.task->feedbackLength() > 0 && e.task->feedbackVersion() != e.uploadedFeedback) {   // setFeedback(...) since the last tick
                if (blfik_set_feedback(prob, static_cast<int32_t>(t), e.task->feedbackBuffer().data(), BLFIK_MEM_HOST) != BLFIK_OK) return fail("Unable to set the feedback");
                e.uploadedFeedback = e.task->feedbackVersion();
            }
        }
    }
    if (pimpl->packedPath) {
        // State and set-points already sit in the page-locked records (KinDynBatch::setRobotState and the tasks' setSetPoint
        // wrote them there): one packed host tick per device, each from its own host thread.
        if (!pimpl->layoutRecords()) return fail("Unable to lay out the packed records");
        const uint32_t fl = flags | pimpl->recordFlags() | (keepState ? BLFIK_TICK_KEEP_STATE : 0u);
        std::vector<int> rcs(pimpl->shards.size(), BLFIK_OK);
        std::vector<std::string> errs(pimpl->shards.size());
        auto run = [&](std::size_t k) {
            const auto& sh = pimpl->shards[k];
            rcs[k] = blfik_tick_packed(sh.prob, pimpl->hRec + sh.first * pimpl->recLen, pimpl->hNu + sh.first * nv, pimpl->hStatus + sh.first, dt, fl, 8);
            if (rcs[k] != BLFIK_OK) errs[k] = blfik_last_error();   // the message is per thread
        };
        std::vector<std::thread> workers;
        for (std::size_t k = 1; k < pimpl->shards.size(); k++) workers.emplace_back(run, k);
        run(0);
        for (auto& w : workers) w.join();
        for (std::size_t k = 0; k < rcs.size(); k++)
            if (rcs[k] != BLFIK_OK) {
                log()->error("{} Unable to run the tick on device {}: {}", logPrefix, pimpl->shards[k].device, errs[k]);
                return false;
            }
        pimpl->nuStale = pimpl->statusStale = true;
        pimpl->unsolved = 0;
        int64_t failed = 0;
        for (std::size_t i = 0; i < N; i++) pimpl->unsolved += pimpl->hStatus[i] != BLFIK_STATUS_SOLVED, failed += pimpl->hStatus[i] == BLFIK_STATUS_FAILED;
        for (int c = 0; c < 3; c++) pimpl->solution.baseVelocity.linear[c] = pimpl->hNu[c], pimpl->solution.baseVelocity.angular[c] = pimpl->hNu[3 + c];
        pimpl->solution.jointVelocity.assign(pimpl->hNu + 6, pimpl->hNu + nv);
        std::fill(pimpl->rawSolution.begin(), pimpl->rawSolution.end(), 0.0);
        std::copy(pimpl->hNu, pimpl->hNu + nv, pimpl->rawSolution.begin() + pimpl->robotVelocityVariable.offset);
        // A batch is not all-or-nothing: the call succeeded, each robot answers for itself (isOutputValid(i), getBatchStatus()).
        if (failed != 0) log()->warn("{} osqp was not able to find a feasible solution for {} of {} robots.", logPrefix, failed, N);
        pimpl->isValid = true;
        return true;
    }
    blfik_problem* prob = pimpl->problem;
    if (!keepState && blfik_set_state(prob, kd.basePositions().data(), kd.baseQuaternions().data(), kd.jointPositions().data(), BLFIK_MEM_HOST) != BLFIK_OK)
        return fail("Unable to set the robot state");
    if (blfik_tick(prob, 1, dt, flags) != BLFIK_OK) return fail("Unable to run the tick");
    if (blfik_get_output(prob, pimpl->nu.data(), BLFIK_MEM_HOST) != BLFIK_OK) return fail("Unable to get the solution");
    if (blfik_get_status(prob, pimpl->status.data(), nullptr, BLFIK_MEM_HOST) != BLFIK_OK) return fail("Unable to get the status");
    int64_t failed = 0;
    pimpl->unsolved = 0;
    for (int32_t st : pimpl->status) pimpl->unsolved += st != BLFIK_STATUS_SOLVED, failed += st == BLFIK_STATUS_FAILED;
    // retrieve the solution of robot 0 (:531-541); the whole batch is in getBatchOutput()
    for (int c = 0; c < 3; c++) pimpl->solution.baseVelocity.linear[c] = pimpl->nu[c], pimpl->solution.baseVelocity.angular[c] = pimpl->nu[3 + c];
    pimpl->solution.jointVelocity.assign(pimpl->nu.begin() + 6, pimpl->nu.begin() + nv);
    std::fill(pimpl->rawSolution.begin(), pimpl->rawSolution.end(), 0.0);
    std::copy(pimpl->nu.begin(), pimpl->nu.begin() + nv, pimpl->rawSolution.begin() + pimpl->robotVelocityVariable.offset);
    if (N == 1 && failed != 0) {   // :518-523: Solved and SolvedInaccurate are accepted, anything else fails the call
        log()->error("{} osqp was not able to find a feasible solution.", logPrefix);
        return false;
    }
    if (failed != 0) log()->warn("{} osqp was not able to find a feasible solution for {} of {} robots.", logPrefix, failed, N);
    pimpl->isValid = true;
    return true;
}

const QPInverseKinematics::State& QPInverseKinematics::getOutput() const { return m_pimpl->solution; }
const VectorXd& QPInverseKinematics::getRawSolution() const { return m_pimpl->rawSolution; }
const std::vector<double>& QPInverseKinematics::getBatchOutput() const
{
    if (m_pimpl->packedPath && m_pimpl->nuStale) {   // the results live in page-locked memory; the vector is filled when somebody asks
        m_pimpl->nu.assign(m_pimpl->hNu, m_pimpl->hNu + m_pimpl->nu.size());
        m_pimpl->nuStale = false;
    }
    return m_pimpl->nu;
}
const std::vector<int32_t>& QPInverseKinematics::getBatchStatus() const
{
    if (m_pimpl->packedPath && m_pimpl->statusStale) {
        m_pimpl->status.assign(m_pimpl->hStatus, m_pimpl->hStatus + m_pimpl->status.size());
        m_pimpl->statusStale = false;
    }
    return m_pimpl->status;
}
const double* QPInverseKinematics::batchOutputData() const { return m_pimpl->packedPath ? m_pimpl->hNu : m_pimpl->nu.data(); }
const int32_t* QPInverseKinematics::batchStatusData() const { return m_pimpl->packedPath ? m_pimpl->hStatus : m_pimpl->status.data(); }
blfik_problem* QPInverseKinematics::handle() { return m_pimpl->problem; }
bool QPInverseKinematics::setTime(std::chrono::nanoseconds t)
{
    if (!m_pimpl->problem) {
        log()->error("[QPInverseKinematics::setTime] Please call finalize() first.");
        return false;
    }
    return blfik_set_time(m_pimpl->problem, std::chrono::duration<double>(t).count()) == BLFIK_OK;
}

std::weak_ptr<QPInverseKinematics::Task> QPInverseKinematics::getTask(const std::string& name) const
{
    auto it = m_pimpl->index.find(name);
    if (it == m_pimpl->index.end()) {   // :558-572
        log()->warn("[QPInverseKinematics::getTask] The task named {} does not exist. A nullptr will be returned.", name);
        return std::shared_ptr<Task>(nullptr);
    }
    return m_pimpl->tasks[it->second].task;
}

std::string QPInverseKinematics::toString() const
{
    std::ostringstream oss;
    oss << "====== QPInverseKinematics class ======\nThe optimization problem is composed by the following tasks:\n";
    for (const auto& e : m_pimpl->tasks) oss << "\t - " << e.name << ": " << e.task->getDescription() << " Priority: " << e.priority << ".\n";
    oss << "Note: The lower is the integer associated to the priority, the higher is the priority.\n==================================\n";
    return oss.str();
}

/** QPInverseKinematics::build, QPInverseKinematics.cpp:627-806 */
IntegrationBasedIKProblem QPInverseKinematics::build(std::weak_ptr<const ParametersHandler::IParametersHandler> handler, std::shared_ptr<KinDynBatch> kinDyn)
{
    constexpr auto logPrefix = "[QPInverseKinematics::build]";
    IntegrationBasedIKProblem problem;
    auto ptr = handler.lock();
    if (ptr == nullptr) {
        log()->error("{} Invalid parameter handler.", logPrefix);
        return IntegrationBasedIKProblem();
    }
    if (!kinDyn || !kinDyn->isValid()) {
        log()->error("{} Invalid kinDyn object.", logPrefix);
        return IntegrationBasedIKProblem();
    }
    auto solver = std::make_unique<QPInverseKinematics>();
    auto ikGroup = ptr->getGroup("IK").lock();
    if (!solver->initialize(ikGroup ? ikGroup : ptr)) {   // :649
        log()->error("{} Unable to initialize the IK.", logPrefix);
        return IntegrationBasedIKProblem();
    }
    solver->m_pimpl->kinDyn = kinDyn;
    std::vector<std::string> tasks;
    if (!ptr->getParameter("tasks", tasks)) {
        log()->error("{} Unable to find the parameter 'tasks'.", logPrefix);
        return IntegrationBasedIKProblem();
    }
    const std::string& variableName = solver->m_pimpl->robotVelocityVariable.name;
    if (!problem.variablesHandler.addVariable(variableName, kinDyn->getNrOfDegreesOfFreedom() + 6)) {   // :655-663
        log()->error("{} Unable to add the variable {}.", logPrefix, variableName);
        return IntegrationBasedIKProblem();
    }
    for (const auto& taskGroupName : tasks) {   // :665-792
        auto taskGroup = ptr->getGroup(taskGroupName).lock();
        if (taskGroup == nullptr) {
            log()->error("{} Unable to get the group named '{}'.", logPrefix, taskGroupName);
            return IntegrationBasedIKProblem();
        }
        auto group = taskGroup->clone();
        group->setParameter("robot_velocity_variable_name", variableName);   // :682-684
        std::string taskType;
        if (!group->getParameter("type", taskType)) {
            log()->error("{} Unable to find the parameter 'type' in the group named '{}'.", logPrefix, taskGroupName);
            return IntegrationBasedIKProblem();
        }
        std::shared_ptr<IKLinearTask> task = IKLinearTaskFactory::createInstance(taskType);   // :696
        if (task == nullptr) {
            log()->error("{} The task type '{}' has not been registered.", logPrefix, taskType);
            return IntegrationBasedIKProblem();
        }
        if (!task->setKinDyn(kinDyn) || !task->initialize(group)) {
            log()->error("{} Unable to initialize the task named '{}'.", logPrefix, taskGroupName);
            return IntegrationBasedIKProblem();
        }
        int priority = 0;
        if (!group->getParameter("priority", priority) || (priority != 0 && priority != 1)) {   // :718-735
            log()->error("{} Unable to find a valid 'priority' (0 or 1) in the group named '{}'.", logPrefix, taskGroupName);
            return IntegrationBasedIKProblem();
        }
        bool ok;
        if (priority == 1) {   // :737-780
            std::string providerType = "ConstantWeightProvider";
            group->getParameter("weight_provider_type", providerType);
            std::shared_ptr<System::WeightProvider> provider;
            bool providerOk = false;
            if (providerType == "ConstantWeightProvider") {
                auto constant = std::make_shared<System::ConstantWeightProvider>();
                providerOk = constant->initialize(group);
                provider = constant;
            } else if (providerType == "MultiStateWeightProvider") {   // the other provider of the reference's factory
                auto multi = std::make_shared<ContinuousDynamicalSystem::MultiStateWeightProvider>();
                providerOk = multi->initialize(group);
                provider = multi;
            } else {
                log()->error("{} The weight provider '{}' is not available.", logPrefix, providerType);
                return IntegrationBasedIKProblem();
            }
            if (!providerOk) {
                log()->error("{} Unable to initialize the weight provider of the task '{}'.", logPrefix, taskGroupName);
                return IntegrationBasedIKProblem();
            }
            problem.weights[taskGroupName] = provider;
            ok = solver->addTask(task, taskGroupName, 1, provider);
        } else {
            ok = solver->addTask(task, taskGroupName, 0);
        }
        if (!ok) {
            log()->error("{} Unable to add the task named '{}'.", logPrefix, taskGroupName);
            return IntegrationBasedIKProblem();
        }
    }
    if (!solver->finalize(problem.variablesHandler)) {   // :794
        log()->error("{} Unable to finalize the IK.", logPrefix);
        return IntegrationBasedIKProblem();
    }
    problem.ik = std::move(solver);
    return problem;
}

// =================================================================== QPFixedBaseInverseKinematics
QPFixedBaseInverseKinematics::QPFixedBaseInverseKinematics() : m_baseSE3Task(std::make_shared<SE3Task>()) {}
QPFixedBaseInverseKinematics::~QPFixedBaseInverseKinematics() = default;

bool QPFixedBaseInverseKinematics::setKinDyn(std::shared_ptr<KinDynBatch> kinDyn)
{
    constexpr auto logPrefix = "[QPFixedBaseInverseKinematics::setKinDyn]";
    if (kinDyn == nullptr) {
        log()->error("{} The kinDyn object is not valid.", logPrefix);
        return false;
    }
    if (!m_baseSE3Task->setKinDyn(kinDyn)) {
        log()->error("{} Unable to set the kinDyn object for SE3Task on the base.", logPrefix);
        return false;
    }
    m_baseLink = kinDyn->getFloatingBase();   // QPFixedBaseInverseKinematics.cpp:84
    m_isKinDynSet = true;
    return true;
}

bool QPFixedBaseInverseKinematics::initialize(std::weak_ptr<const ParametersHandler::IParametersHandler> handler)
{
    constexpr auto logPrefix = "[QPFixedBaseInverseKinematics::initialize]";
    if (!m_isKinDynSet) {
        log()->error("{} The object kinDyn is not valid.", logPrefix);
        return false;
    }
    auto ptr = handler.lock();
    if (ptr == nullptr) {
        log()->error("{} The parameter handler is not valid.", logPrefix);
        return false;
    }
    // createBaseSE3Task (QPFixedBaseInverseKinematics.cpp:24-58): zero gains, controller disabled, zero feed-forward
    std::string robotVelocityVariableName;
    if (!ptr->getParameter("robot_velocity_variable_name", robotVelocityVariableName)) {
        log()->error("{} Unable to get the parameter 'robot_velocity_variable_name'", logPrefix);
        return false;
    }
    auto baseHandler = std::make_shared<ParametersHandler::StdImplementation>();
    baseHandler->setParameter("robot_velocity_variable_name", robotVelocityVariableName);
    baseHandler->setParameter("kp_linear", 0.0);
    baseHandler->setParameter("kp_angular", 0.0);
    baseHandler->setParameter("frame_name", m_baseLink);
    if (!m_baseSE3Task->initialize(baseHandler)) {
        log()->error("{} Error creating SE3task for the base.", logPrefix);
        return false;
    }
    m_baseSE3Task->setTaskControllerMode(SE3Task::Mode::Disable);
    m_baseSE3Task->setSetPoint(SE3d(), SE3d::Tangent::Zero());
    if (!this->QPInverseKinematics::initialize(ptr)) {
        log()->error("{} Unable to initialize QPInverseKinematics class.", logPrefix);
        return false;
    }
    constexpr std::size_t highPriority = 0;
    return this->addTask(m_baseSE3Task, "base_se3_task", highPriority);
}

}  // namespace IK
}  // namespace BipedalLocomotion
