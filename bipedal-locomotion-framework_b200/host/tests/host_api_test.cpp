// host_api_test.cpp -- tests of the C++ host layer, written the way the reference's Catch2 tests are
// (src/IK/tests/*TaskTest.cpp, QPInverseKinematicsTest.cpp, IntegratorFloatingBaseSystemVelocityKinematics.cpp).
//   host_api_test cpu <model.txt>    checks that need no GPU (parameters, variables, addTask rules, error paths)
//   host_api_test gpu <model.txt>    checks that run the CUDA path (task A/b identities, closed loop, batch, integrator)
#include <BipedalLocomotion/ContinuousDynamicalSystem.h>
#include <BipedalLocomotion/IK.h>
#include <BipedalLocomotion/Planners.h>

#include <cstdlib>
#include <cstring>
#include <random>

using namespace BipedalLocomotion;
using namespace BipedalLocomotion::IK;
using namespace BipedalLocomotion::System;
using namespace BipedalLocomotion::ContinuousDynamicalSystem;
using namespace BipedalLocomotion::ParametersHandler;
using namespace std::chrono_literals;

static int g_failures = 0, g_checks = 0;
#define REQUIRE(cond)                                                              \
    do {                                                                           \
        g_checks++;                                                                \
        if (!(cond)) {                                                             \
            g_failures++;                                                          \
            std::fprintf(stderr, "REQUIRE failed %s:%d: %s\n", __FILE__, __LINE__, #cond); \
        }                                                                          \
    } while (0)
#define REQUIRE_FALSE(cond) REQUIRE(!(cond))

constexpr auto robotVelocity = "robotVelocity";

static std::shared_ptr<KinDynBatch> loadKinDyn(const char* path, std::size_t N = 1)
{
    auto kd = std::make_shared<KinDynBatch>();
    if (!kd->loadRobotModelFromFile(path)) return nullptr;
    kd->setBatchSize(N);
    return kd;
}

// the ergoCub IK of utilities/balancing-position-control/config/robots/ergoCubGazeboV1/blf_balancing_position_control/ik.ini
static std::shared_ptr<IParametersHandler> ergoCubHandler(std::size_t n, bool withLimits, const std::vector<int>& devices = {}, bool compactRecords = false)
{
    auto h = std::make_shared<StdImplementation>();
    std::vector<std::string> tasks{"COM_TASK", "RIGHT_FOOT_TASK", "LEFT_FOOT_TASK", "TORSO_TASK", "JOINT_REGULARIZATION_TASK"};
    if (withLimits) tasks.push_back("JOINT_LIMITS_TASK"), tasks.push_back("JOINT_VELOCITY_LIMITS_TASK");
    h->setParameter("tasks", tasks);
    auto ik = std::make_shared<StdImplementation>();
    ik->setParameter("robot_velocity_variable_name", robotVelocity);
    if (!devices.empty()) ik->setParameter("cuda_devices", devices);
    if (compactRecords) ik->setParameter("compact_records", true);
    h->setGroup("IK", ik);
    auto com = std::make_shared<StdImplementation>();
    com->setParameter("type", "CoMTask"), com->setParameter("kp_linear", 2.0), com->setParameter("priority", 0);
    com->setParameter("mask", std::vector<bool>{true, true, true});
    h->setGroup("COM_TASK", com);
    for (auto side : {std::make_pair("RIGHT_FOOT_TASK", "r_sole"), std::make_pair("LEFT_FOOT_TASK", "l_sole")}) {
        auto foot = std::make_shared<StdImplementation>();
        foot->setParameter("type", "SE3Task"), foot->setParameter("frame_name", side.second);
        foot->setParameter("kp_linear", 7.0), foot->setParameter("kp_angular", 5.0), foot->setParameter("priority", 0);
        h->setGroup(side.first, foot);
    }
    auto torso = std::make_shared<StdImplementation>();
    torso->setParameter("type", "SO3Task"), torso->setParameter("frame_name", "chest"), torso->setParameter("kp_angular", 5.0);
    torso->setParameter("priority", 1), torso->setParameter("weight", std::vector<double>{5.0, 5.0, 5.0});
    h->setGroup("TORSO_TASK", torso);
    auto reg = std::make_shared<StdImplementation>();
    reg->setParameter("type", "JointTrackingTask"), reg->setParameter("kp", std::vector<double>(n, 5.0));
    reg->setParameter("priority", 1), reg->setParameter("weight", std::vector<double>(n, 1.0));
    h->setGroup("JOINT_REGULARIZATION_TASK", reg);
    auto lim = std::make_shared<StdImplementation>();
    lim->setParameter("type", "JointLimitsTask"), lim->setParameter("klim", std::vector<double>(n, 0.5));
    lim->setParameter("sampling_time", 0.01), lim->setParameter("use_model_limits", true), lim->setParameter("priority", 0);
    h->setGroup("JOINT_LIMITS_TASK", lim);
    auto vel = std::make_shared<StdImplementation>();
    vel->setParameter("type", "JointVelocityLimitsTask"), vel->setParameter("upper_limits", std::vector<double>(n, 2.0));
    vel->setParameter("lower_limits", std::vector<double>(n, -2.0)), vel->setParameter("priority", 0);
    h->setGroup("JOINT_VELOCITY_LIMITS_TASK", vel);
    return h;
}

// ------------------------------------------------------------------------------------------ CPU-only
static void testParametersAndVariables()
{
    auto h = std::make_shared<StdImplementation>();
    h->setParameter("answer", 42), h->setParameter("pi", 3.14), h->setParameter("name", "blf"), h->setParameter("flag", true);
    h->setParameter("vec", std::vector<double>{1, 2, 3});
    int i = 0;
    double d = 0;
    std::string s;
    bool b = false;
    std::vector<double> v;
    REQUIRE(h->getParameter("answer", i) && i == 42);
    REQUIRE(h->getParameter("pi", d) && d == 3.14);
    REQUIRE(h->getParameter("answer", d) && d == 42.0);
    REQUIRE(h->getParameter("name", s) && s == "blf");
    REQUIRE(h->getParameter("flag", b) && b);
    REQUIRE(h->getParameter("vec", v) && v.size() == 3);
    REQUIRE_FALSE(h->getParameter("missing", i));
    REQUIRE_FALSE(h->getParameter("name", i));
    auto g = std::make_shared<StdImplementation>();
    g->setParameter("x", 1);
    REQUIRE(h->setGroup("G", g));
    g->setParameter("x", 2);   // setGroup stored a copy
    REQUIRE(h->getGroup("G").lock() && h->getGroup("G").lock()->getParameter("x", i) && i == 1);
    REQUIRE(h->getGroup("nope").lock() == nullptr);
    auto c = h->clone();
    c->setParameter("answer", 7);
    REQUIRE(h->getParameter("answer", i) && i == 42);
    std::chrono::nanoseconds ns;
    h->setParameter("dt", 0.01);
    REQUIRE(h->getParameter("dt", ns) && ns == 10ms);

    VariablesHandler vh;
    REQUIRE(vh.addVariable("dummy1", 10));
    REQUIRE(vh.addVariable(robotVelocity, 29));
    REQUIRE(vh.addVariable("dummy2", 15));
    REQUIRE_FALSE(vh.addVariable("dummy2", 3));
    REQUIRE(vh.getNumberOfVariables() == 54);
    REQUIRE(vh.getVariable(robotVelocity).offset == 10 && vh.getVariable(robotVelocity).size == 29);
    REQUIRE_FALSE(vh.getVariable("missing").isValid());
    auto vhh = std::make_shared<StdImplementation>();
    vhh->setParameter("variables_name", std::vector<std::string>{"a", "b"});
    vhh->setParameter("variables_size", std::vector<int>{3, 4});
    VariablesHandler vh2;
    REQUIRE(vh2.initialize(vhh) && vh2.getNumberOfVariables() == 7 && vh2.getVariable("b").offset == 3);

    ConstantWeightProvider w;
    REQUIRE_FALSE(w.isOutputValid());
    auto wh = std::make_shared<StdImplementation>();
    REQUIRE_FALSE(w.initialize(wh));
    wh->setParameter("weight", std::vector<double>{1, 2});
    REQUIRE(w.initialize(wh) && w.isOutputValid() && w.getOutput().size() == 2);
}

static void testTaskInitializationAndAddTaskRules(const char* modelPath)
{
    log()->setLevel(Logger::Level::off);   // the error paths below are expected
    auto kinDyn = loadKinDyn(modelPath);
    REQUIRE(kinDyn != nullptr);
    const std::size_t n = kinDyn->getNrOfDegreesOfFreedom();
    REQUIRE(n == 23);
    REQUIRE(kinDyn->getFrameIndex("l_sole") >= 0 && kinDyn->getFrameIndex("not_a_frame") == -1);

    auto se3 = std::make_shared<SE3Task>();
    auto h = std::make_shared<StdImplementation>();
    REQUIRE_FALSE(se3->initialize(h));                       // no kinDyn yet
    REQUIRE(se3->setKinDyn(kinDyn));
    REQUIRE_FALSE(se3->initialize(h));                       // robot_velocity_variable_name missing
    h->setParameter("robot_velocity_variable_name", robotVelocity);
    REQUIRE_FALSE(se3->initialize(h));                       // frame_name missing
    h->setParameter("frame_name", "not_a_frame");
    REQUIRE_FALSE(se3->initialize(h));
    h->setParameter("frame_name", "l_sole");
    REQUIRE_FALSE(se3->initialize(h));                       // gains missing
    h->setParameter("kp_linear", 7.0);
    h->setParameter("kp_angular", std::vector<double>{5.0, 5.0, 5.0});
    REQUIRE(se3->initialize(h));
    REQUIRE(se3->size() == 6 && se3->type() == LinearTask::Type::equality);
    h->setParameter("mask", std::vector<bool>{true, false, true});
    auto se3m = std::make_shared<SE3Task>();
    REQUIRE(se3m->setKinDyn(kinDyn) && se3m->initialize(h) && se3m->size() == 5);
    REQUIRE_FALSE(se3->update());                            // set-point never set (SE3Task.cpp:245-249)
    REQUIRE_FALSE(se3->isValid());

    auto jt = std::make_shared<JointTrackingTask>();
    auto jh = std::make_shared<StdImplementation>();
    jh->setParameter("robot_velocity_variable_name", robotVelocity);
    jh->setParameter("kp", std::vector<double>(n - 1, 1.0));
    REQUIRE(jt->setKinDyn(kinDyn));
    REQUIRE_FALSE(jt->initialize(jh));                       // wrong kp size (JointTrackingTask.cpp:100-110)
    jh->setParameter("kp", std::vector<double>(n, 1.0));
    REQUIRE(jt->initialize(jh) && jt->size() == n);
    REQUIRE_FALSE(jt->setSetPoint(VectorXd(n + 1, 0.0)));    // wrong set-point size
    REQUIRE(jt->setSetPoint(VectorXd(n, 0.0)));

    auto jl = std::make_shared<JointLimitsTask>();
    auto lh = std::make_shared<StdImplementation>();
    lh->setParameter("robot_velocity_variable_name", robotVelocity);
    lh->setParameter("klim", std::vector<double>(n, 1.5));
    lh->setParameter("sampling_time", 0.01);
    REQUIRE(jl->setKinDyn(kinDyn));
    REQUIRE_FALSE(jl->initialize(lh));                       // klim outside [0, 1]
    lh->setParameter("klim", std::vector<double>(n, 0.5));
    REQUIRE_FALSE(jl->initialize(lh));                       // use_model_limits is mandatory
    lh->setParameter("use_model_limits", true);
    REQUIRE(jl->initialize(lh) && jl->size() == 2 * n && jl->type() == LinearTask::Type::inequality);

    QPInverseKinematics ik;
    VectorXd w6(6, 1.0);
    REQUIRE_FALSE(ik.addTask(se3, "se3", 0));                // initialize() first (QPInverseKinematics.cpp:173-177)
    auto ikh = std::make_shared<StdImplementation>();
    REQUIRE_FALSE(ik.initialize(ikh));
    ikh->setParameter("robot_velocity_variable_name", robotVelocity);
    REQUIRE(ik.initialize(ikh));
    REQUIRE(ik.addTask(se3, "se3", 0));
    REQUIRE_FALSE(ik.addTask(se3, "se3", 0));                // duplicate name
    REQUIRE_FALSE(ik.addTask(se3m, "se3m", 2));              // priority must be 0 or 1
    REQUIRE_FALSE(ik.addTask(jl, "limits", 1, VectorXd(2 * n, 1.0)));   // inequality with priority 1
    REQUIRE_FALSE(ik.addTask(jt, "reg", 1));                 // priority 1 needs a weight provider
    REQUIRE_FALSE(ik.addTask(jt, "reg", 1, VectorXd(n - 1, 1.0)));      // wrong weight size
    REQUIRE(ik.addTask(jt, "reg", 1, VectorXd(n, 1.0)));
    REQUIRE(ik.addTask(jl, "limits", 0));
    REQUIRE_FALSE(ik.setTaskWeight("se3", w6));              // only priority-1 tasks have a weight (:268-276)
    REQUIRE_FALSE(ik.setTaskWeight("nope", w6));
    REQUIRE_FALSE(ik.setTaskWeight("reg", w6));              // wrong size
    REQUIRE(ik.setTaskWeight("reg", VectorXd(n, 10.0)));
    REQUIRE(ik.getTaskWeightProvider("reg").lock() && ik.getTaskWeightProvider("reg").lock()->getOutput()[0] == 10.0);
    REQUIRE(ik.getTaskNames().size() == 3);
    REQUIRE(ik.getTask("se3").lock() == se3 && ik.getTask("nope").lock() == nullptr);
    REQUIRE_FALSE(ik.advance());                             // finalize() first (:421-425)
    REQUIRE_FALSE(ik.isOutputValid());
    REQUIRE(ik.toString().find("se3") != std::string::npos);
    VariablesHandler bad;
    bad.addVariable("other", 29);
    REQUIRE_FALSE(ik.finalize(bad));                         // the robot velocity variable is not in the handler

    // DistanceTask / GravityTask initialisation rules (DistanceTask.cpp:95-175, GravityTask.cpp:76-147)
    {
        DistanceTask dist;
        auto dh = std::make_shared<StdImplementation>();
        dh->setParameter("robot_velocity_variable_name", robotVelocity);
        REQUIRE(dist.setKinDyn(kinDyn));
        REQUIRE_FALSE(dist.initialize(dh));                  // kp missing
        dh->setParameter("kp", 2.0);
        REQUIRE_FALSE(dist.initialize(dh));                  // target_frame_name missing
        dh->setParameter("target_frame_name", "l_hand_palm"), dh->setParameter("reference_frame_name", "not_a_frame");
        REQUIRE_FALSE(dist.initialize(dh));                  // unknown reference frame
        dh->setParameter("reference_frame_name", "");
        REQUIRE(dist.initialize(dh) && dist.size() == 1 && dist.type() == IKLinearTask::Type::equality);
        GravityTask grav;
        auto gh = std::make_shared<StdImplementation>();
        gh->setParameter("robot_velocity_variable_name", robotVelocity), gh->setParameter("kp", 2.0);
        REQUIRE(grav.setKinDyn(kinDyn));
        REQUIRE_FALSE(grav.initialize(gh));                  // target_frame_name missing
        gh->setParameter("target_frame_name", "chest");
        REQUIRE(grav.initialize(gh) && grav.size() == 2);
        REQUIRE(IKLinearTaskFactory::createInstance("DistanceTask") != nullptr && IKLinearTaskFactory::createInstance("GravityTask") != nullptr);
    }

    // integrator error conventions
    ForwardEuler<FloatingBaseSystemVelocityKinematics> integrator;
    REQUIRE_FALSE(integrator.integrate(0s, 10ms));           // no dynamical system
    REQUIRE_FALSE(integrator.setIntegrationStep(0ns));
    REQUIRE(integrator.setIntegrationStep(10ms));
    auto system = std::make_shared<FloatingBaseSystemVelocityKinematics>();
    REQUIRE(integrator.setDynamicalSystem(system));
    REQUIRE_FALSE(integrator.setDynamicalSystem(system));    // only once (Integrator.h:94-99)
    REQUIRE_FALSE(integrator.integrate(10ms, 0s));           // final time before the initial one
    REQUIRE(system->setState({{0, 0, 0}, SO3d::Identity(), VectorXd(n, 0.0)}));
    REQUIRE_FALSE(integrator.integrate(0s, 10ms));           // no control input yet
    REQUIRE_FALSE(system->setControlInput({{0, 0, 0, 0, 0, 0}, VectorXd(n + 1, 0.0)}));   // wrong size (.cpp:52-56)
    log()->setLevel(Logger::Level::warn);
}

// 'First order smoother' (tests/FirstOrderSmoother.cpp:22-100: step response against 1 - exp(-3 t / settling_time), 1e-2) and
// 'Multistate weight provider' (tests/MultiStateWeightProvider.cpp:22-118: walking <-> stance within 1e-2 after 50 steps)
static void testWeightProviders()
{
    using BipedalLocomotion::ContinuousDynamicalSystem::FirstOrderSmoother;
    using BipedalLocomotion::ContinuousDynamicalSystem::MultiStateWeightProvider;
    {
        auto params = std::make_shared<StdImplementation>();
        FirstOrderSmoother smoother;
        log()->setLevel(Logger::Level::off);
        REQUIRE_FALSE(smoother.initialize(params));              // settling_time missing
        REQUIRE_FALSE(smoother.reset(VectorXd(2, 0.0)));         // not initialised
        log()->setLevel(Logger::Level::warn);
        params->setParameter("settling_time", 0.1), params->setParameter("sampling_time", std::chrono::nanoseconds(100us));
        REQUIRE(smoother.initialize(params) && smoother.reset(VectorXd(2, 0.0)));
        log()->setLevel(Logger::Level::off);
        REQUIRE_FALSE(smoother.advance());                       // no input yet
        log()->setLevel(Logger::Level::warn);
        REQUIRE(smoother.setInput(VectorXd(2, 1.0)));
        double worst = 0, settled = -1;
        for (unsigned int i = 0; i < 1000; i++) {
            const double t = 100e-6 * i, expected = 1.0 - std::exp(-30.0 * t);
            worst = std::max(worst, std::abs(smoother.getOutput()[0] - expected));
            if (settled < 0 && smoother.getOutput()[1] > 0.95) settled = t;
            REQUIRE(smoother.advance());
        }
        REQUIRE(worst < 1e-2 && std::abs(settled - 0.1) < 1e-2);   // 5 % band reached at the settling time
    }
    {
        const VectorXd walking{10, 100, 1000}, stance{50, 30, 20};
        auto params = std::make_shared<StdImplementation>();
        params->setParameter("settling_time", 0.1), params->setParameter("sampling_time", std::chrono::nanoseconds(10ms));
        params->setParameter("states", std::vector<std::string>{"WALKING", "STANCE"});
        auto pw = std::make_shared<StdImplementation>(), ps = std::make_shared<StdImplementation>();
        pw->setParameter("name", "walking"), pw->setParameter("weight", walking);
        ps->setParameter("name", "stance"), ps->setParameter("weight", stance);
        REQUIRE(params->setGroup("WALKING", pw) && params->setGroup("STANCE", ps));
        MultiStateWeightProvider provider;
        REQUIRE(provider.initialize(params));
        REQUIRE(provider.getOutput() == walking);
        REQUIRE(provider.setState("stance"));
        log()->setLevel(Logger::Level::off);
        REQUIRE_FALSE(provider.setState("foo"));
        log()->setLevel(Logger::Level::warn);
        auto close = [](const VectorXd& a, const VectorXd& b, double rel) {
            double d = 0, nb = 0;
            for (std::size_t i = 0; i < a.size(); i++) d += (a[i] - b[i]) * (a[i] - b[i]), nb += b[i] * b[i];
            return std::sqrt(d) <= rel * std::sqrt(nb);
        };
        for (unsigned int i = 0; i < 50; i++) REQUIRE(provider.advance());
        REQUIRE(provider.isOutputValid() && close(provider.getOutput(), stance, 1e-2));
        REQUIRE(provider.reset("stance") && provider.getOutput() == stance && provider.setState("walking"));
        for (unsigned int i = 0; i < 50; i++) REQUIRE(provider.advance());
        REQUIRE(close(provider.getOutput(), walking, 1e-2));
        // states of different sizes are rejected (MultiStateWeightProvider.cpp:77-96)
        auto ps2 = std::make_shared<StdImplementation>();
        ps2->setParameter("name", "stance"), ps2->setParameter("weight", VectorXd{1, 2});
        REQUIRE(params->setGroup("STANCE", ps2));
        MultiStateWeightProvider bad;
        log()->setLevel(Logger::Level::off);
        REQUIRE_FALSE(bad.initialize(params));
        log()->setLevel(Logger::Level::warn);
    }
}

// ------------------------------------------------------------------------------------------ GPU
static double maxAbsDiff(const MatrixXd& A, std::size_t r0, std::size_t c0, const MatrixXd& B, std::size_t br0)
{
    double m = 0;
    for (std::size_t r = 0; r < B.rows - br0 && r + r0 < A.rows; r++)
        for (std::size_t c = 0; c < B.cols; c++) m = std::max(m, std::abs(A(r0 + r, c0 + c) - B(br0 + r, c)));
    return m;
}

static VectorXd nominalJoints(const KinDynBatch& kd, std::mt19937& gen, double spread)
{
    std::uniform_real_distribution<double> u(-spread, spread);
    VectorXd s(kd.getNrOfDegreesOfFreedom());
    for (std::size_t j = 0; j < s.size(); j++) s[j] = 0.5 * (kd.jointLowerLimits()[j] + kd.jointUpperLimits()[j]) + u(gen);
    return s;
}

// SE3TaskTest.cpp:79-139 / CoMTaskTest.cpp:82-130: A's robot-velocity block is the Jacobian, the neighbouring column
// blocks (dummy1, dummy2) stay zero, b is the control law
static void testTaskMatrices(const char* modelPath)
{
    auto kinDyn = loadKinDyn(modelPath);
    std::mt19937 gen(3);
    REQUIRE(kinDyn->setRobotState({0.1, -0.2, 0.7}, SO3d::Exp({0.1, -0.2, 0.3}), nominalJoints(*kinDyn, gen, 0.3)));
    VariablesHandler vh;
    vh.addVariable("dummy1", 10), vh.addVariable(robotVelocity, kinDyn->getNrOfDegreesOfFreedom() + 6), vh.addVariable("dummy2", 15);
    auto h = std::make_shared<StdImplementation>();
    h->setParameter("robot_velocity_variable_name", robotVelocity), h->setParameter("frame_name", "l_sole");
    h->setParameter("kp_linear", 2.0), h->setParameter("kp_angular", 3.0);
    SE3Task task;
    REQUIRE(task.setKinDyn(kinDyn) && task.initialize(h) && task.setVariablesHandler(vh));
    SE3d desired({0.2, 0.1, 0.05}, SO3d::Exp({0.0, 0.1, -0.1}));
    SE3d::Tangent ff;
    ff.linear = {0.1, 0.2, 0.3}, ff.angular = {-0.1, 0.0, 0.2};
    REQUIRE(task.setSetPoint(desired, ff));
    REQUIRE(task.update() && task.isValid());
    const MatrixXd& A = task.getA();
    REQUIRE(A.rows == 6 && A.cols == 10 + 29 + 15);
    MatrixXd J;
    SE3d pose;
    std::array<double, 9> R;
    REQUIRE(kinDyn->getFrameFreeFloatingJacobian(kinDyn->getFrameIndex("l_sole"), J));
    REQUIRE(kinDyn->getWorldTransform(kinDyn->getFrameIndex("l_sole"), pose, R));
    REQUIRE(maxAbsDiff(A, 0, 10, J, 0) == 0.0);
    double outside = 0;
    for (std::size_t r = 0; r < 6; r++)
        for (std::size_t c = 0; c < A.cols; c++)
            if (c < 10 || c >= 39) outside = std::max(outside, std::abs(A(r, c)));
    REQUIRE(outside == 0.0);
    for (int c = 0; c < 3; c++) REQUIRE(std::abs(task.getB()[c] - (ff.linear[c] + 2.0 * (desired.p[c] - pose.p[c]))) < 1e-12);
    // base block of the Jacobian: [I, -S(p_F - p_B); 0, I]
    REQUIRE(J(0, 0) == 1 && J(1, 1) == 1 && J(3, 3) == 1 && J(3, 0) == 0 && std::abs(J(0, 4) - (pose.p[2] - 0.7)) < 1e-12);
    task.setTaskControllerMode(SE3Task::Mode::Disable);
    REQUIRE(task.update());
    for (int c = 0; c < 3; c++) REQUIRE(task.getB()[c] == ff.linear[c] && task.getB()[3 + c] == ff.angular[c]);   // feed-forward only

    CoMTask com;
    auto ch = std::make_shared<StdImplementation>();
    ch->setParameter("robot_velocity_variable_name", robotVelocity), ch->setParameter("kp_linear", 4.0);
    REQUIRE(com.setKinDyn(kinDyn) && com.initialize(ch) && com.setVariablesHandler(vh) && com.setSetPoint({0, 0, 0.5}));
    REQUIRE(com.update());
    MatrixXd Jc;
    Vector3d c0;
    REQUIRE(kinDyn->getCenterOfMassJacobian(Jc) && kinDyn->getCenterOfMassPosition(c0));
    REQUIRE(maxAbsDiff(com.getA(), 0, 10, Jc, 0) == 0.0);
    REQUIRE(std::abs(com.getB()[2] - 4.0 * (0.5 - c0[2])) < 1e-12);

    JointTrackingTask jt;
    auto jh = std::make_shared<StdImplementation>();
    jh->setParameter("robot_velocity_variable_name", robotVelocity), jh->setParameter("kp", std::vector<double>(23, 3.0));
    REQUIRE(jt.setKinDyn(kinDyn) && jt.initialize(jh) && jt.setVariablesHandler(vh) && jt.setSetPoint(VectorXd(23, 0.0)) && jt.update());
    VectorXd s;
    kinDyn->getJointPos(s);
    for (std::size_t j = 0; j < 23; j++) {   // JointTrackingTaskTest.cpp:84-114: b = -Kp s for a zero set-point, A = [0 I]
        REQUIRE(std::abs(jt.getB()[j] + 3.0 * s[j]) < 1e-13);
        REQUIRE(jt.getA()(j, 10 + 6 + j) == 1.0 && jt.getA()(j, 10 + j) == 0.0);
    }

    // exogenous feedback (SE3Task.cpp:196-214, :261-268, :338-366; CoMTask.cpp:116, :192-195, :262-273): b follows the measured
    // pose, A stays the model's Jacobian; a part whose key is off is ignored
    {
        auto eh = std::make_shared<StdImplementation>();
        eh->setParameter("robot_velocity_variable_name", robotVelocity), eh->setParameter("frame_name", "l_sole");
        eh->setParameter("kp_linear", 2.0), eh->setParameter("kp_angular", 3.0), eh->setParameter("use_position_exogenous_feedback", true);
        SE3Task exo;
        REQUIRE(exo.setKinDyn(kinDyn) && exo.initialize(eh) && exo.setVariablesHandler(vh) && exo.setSetPoint(desired, ff));
        REQUIRE(exo.feedbackLength() == 7 && task.feedbackLength() == 0);
        REQUIRE(exo.update());
        for (int c = 0; c < 3; c++) REQUIRE(std::abs(exo.getB()[c] - (ff.linear[c] + 2.0 * desired.p[c])) < 1e-12);   // no feedback yet: p = 0
        const Vector3d measured{0.21, 0.12, 0.03};
        REQUIRE(exo.setFeedback(SE3d(measured, SO3d::Exp({0.3, 0.2, 0.1}))) && exo.update());
        REQUIRE(maxAbsDiff(exo.getA(), 0, 10, J, 0) == 0.0);
        task.setTaskControllerMode(SE3Task::Mode::Enable);
        REQUIRE(task.update());
        for (int c = 0; c < 3; c++) {
            REQUIRE(std::abs(exo.getB()[c] - (ff.linear[c] + 2.0 * (desired.p[c] - measured[c]))) < 1e-12);
            REQUIRE(std::abs(exo.getB()[3 + c] - task.getB()[3 + c]) < 1e-12);   // orientation key off: the model's orientation
        }
        CoMTask exoCom;
        auto ech = std::make_shared<StdImplementation>();
        ech->setParameter("robot_velocity_variable_name", robotVelocity), ech->setParameter("kp_linear", 4.0), ech->setParameter("use_exogenous_feedback", true);
        REQUIRE(exoCom.setKinDyn(kinDyn) && exoCom.initialize(ech) && exoCom.setVariablesHandler(vh) && exoCom.setSetPoint({0, 0, 0.5}));
        REQUIRE(exoCom.setFeedback({0.01, -0.02, 0.45}) && exoCom.update());
        REQUIRE(std::abs(exoCom.getB()[2] - 4.0 * (0.5 - 0.45)) < 1e-12 && std::abs(exoCom.getB()[0] + 4.0 * 0.01) < 1e-12);
    }

    // DistanceTaskTest.cpp / GravityTaskTest.cpp: A and b against the formulas of DistanceTask.cpp:196-243, GravityTask.cpp:160-199
    DistanceTask dist;
    auto dh = std::make_shared<StdImplementation>();
    dh->setParameter("robot_velocity_variable_name", robotVelocity), dh->setParameter("kp", 2.5);
    dh->setParameter("target_frame_name", "l_hand_palm"), dh->setParameter("reference_frame_name", "r_sole");
    REQUIRE(dist.setKinDyn(kinDyn) && dist.initialize(dh) && dist.setVariablesHandler(vh));
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(dist.update());   // no set-point yet
    log()->setLevel(Logger::Level::warn);
    REQUIRE(dist.setDesiredDistance(-0.4) && dist.update());   // the sign is dropped (DistanceTask.cpp:240)
    SE3d hand, sole;
    MatrixXd Jh, Js;
    REQUIRE(kinDyn->getWorldTransform(kinDyn->getFrameIndex("l_hand_palm"), hand, R) && kinDyn->getWorldTransform(kinDyn->getFrameIndex("r_sole"), sole, R));
    REQUIRE(kinDyn->getFrameFreeFloatingJacobian(kinDyn->getFrameIndex("l_hand_palm"), Jh) && kinDyn->getFrameFreeFloatingJacobian(kinDyn->getFrameIndex("r_sole"), Js));
    const Vector3d dv{hand.p[0] - sole.p[0], hand.p[1] - sole.p[1], hand.p[2] - sole.p[2]};
    const double dn = std::sqrt(dv[0] * dv[0] + dv[1] * dv[1] + dv[2] * dv[2]);
    REQUIRE(std::abs(dist.getDistance() - dn) < 1e-14 && std::abs(dist.getB()[0] - 2.5 * (0.4 - dn)) < 1e-13);
    double worst = 0, baseCols = 0;
    for (std::size_t c = 0; c < 29; c++) {
        double a = 0;
        for (int r = 0; r < 3; r++) a += dv[r] * (Jh(r, c) - Js(r, c));
        worst = std::max(worst, std::abs(dist.getA()(0, 10 + c) - a / dn));
        if (c < 6) baseCols = std::max(baseCols, std::abs(dist.getA()(0, 10 + c)));
    }
    REQUIRE(worst < 1e-14 && baseCols < 1e-14);   // a relative distance does not depend on the base (getRelativeJacobian has no base block)
    GravityTask grav;
    auto gh = std::make_shared<StdImplementation>();
    gh->setParameter("robot_velocity_variable_name", robotVelocity), gh->setParameter("kp", 3.0), gh->setParameter("target_frame_name", "chest");
    REQUIRE(grav.setKinDyn(kinDyn) && grav.initialize(gh) && grav.setVariablesHandler(vh));
    REQUIRE(grav.setSetPoint({0.0, 0.6, 1.8}, {0.1, -0.2, 0.3}) && grav.update());
    SE3d chest;
    MatrixXd Jchest;
    REQUIRE(kinDyn->getWorldTransform(kinDyn->getFrameIndex("chest"), chest, R) && kinDyn->getFrameFreeFloatingJacobian(kinDyn->getFrameIndex("chest"), Jchest));
    REQUIRE(grav.getA().rows == 2 && maxAbsDiff(grav.getA(), 0, 10, Jchest, 3) == 0.0);
    {
        const double nz = std::sqrt(0.6 * 0.6 + 1.8 * 1.8), z[3] = {0.0, 0.6 / nz, 1.8 / nz}, f[3] = {0.1, -0.2, 0.3};
        double zw[3], fw[3];
        for (int r = 0; r < 3; r++) {
            zw[r] = R[3 * r] * z[0] + R[3 * r + 1] * z[1] + R[3 * r + 2] * z[2];
            fw[r] = R[3 * r] * f[0] + R[3 * r + 1] * f[1] + R[3 * r + 2] * f[2];
        }
        REQUIRE(std::abs(grav.getB()[0] - (fw[0] + 3.0 * zw[1])) < 1e-13 && std::abs(grav.getB()[1] - (fw[1] - 3.0 * zw[0])) < 1e-13);
    }
}

// 'With builder' (QPInverseKinematicsTest.cpp:642-798) on the ergoCub configuration: build() from a parameter handler,
// closed loop advance() -> setControlInput -> integrate(0, dT) until the set-points are reached
static void testClosedLoop(const char* modelPath, std::size_t N, bool withLimits)
{
    constexpr auto dT = 10ms;
    auto kinDyn = loadKinDyn(modelPath, N);
    const std::size_t n = kinDyn->getNrOfDegreesOfFreedom();
    std::mt19937 gen(42);
    std::normal_distribution<double> noise(0.0, 0.05);

    // desired configuration -> set-points; perturbed configuration -> initial state
    auto desiredKinDyn = loadKinDyn(modelPath, N);
    std::vector<VectorXd> sDes(N);
    for (std::size_t i = 0; i < N; i++) {
        sDes[i] = nominalJoints(*kinDyn, gen, 0.15);
        REQUIRE(desiredKinDyn->setRobotState({0.0, 0.0, 0.75}, SO3d::Exp({0.0, 0.05, 0.1}), sDes[i], i));
    }
    auto problem = QPInverseKinematics::build(ergoCubHandler(n, withLimits), kinDyn);
    REQUIRE(problem.isValid());
    auto& [variablesHandler, weights, ik] = problem;
    REQUIRE(variablesHandler.getNumberOfVariables() == n + 6 && weights.size() == 2);
    auto comTask = std::dynamic_pointer_cast<CoMTask>(ik->getTask("COM_TASK").lock());
    auto rFoot = std::dynamic_pointer_cast<SE3Task>(ik->getTask("RIGHT_FOOT_TASK").lock());
    auto lFoot = std::dynamic_pointer_cast<SE3Task>(ik->getTask("LEFT_FOOT_TASK").lock());
    auto torso = std::dynamic_pointer_cast<SO3Task>(ik->getTask("TORSO_TASK").lock());
    auto reg = std::dynamic_pointer_cast<JointTrackingTask>(ik->getTask("JOINT_REGULARIZATION_TASK").lock());
    REQUIRE(comTask && rFoot && lFoot && torso && reg);

    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(ik->advance());   // no set-point yet
    log()->setLevel(Logger::Level::warn);

    std::vector<Vector3d> comDes(N);
    std::vector<SE3d> rDes(N), lDes(N);
    for (std::size_t i = 0; i < N; i++) {
        std::array<double, 9> R;
        SE3d chest;
        REQUIRE(desiredKinDyn->getCenterOfMassPosition(comDes[i], i));
        REQUIRE(desiredKinDyn->getWorldTransform(kinDyn->getFrameIndex("r_sole"), rDes[i], R, i));
        REQUIRE(desiredKinDyn->getWorldTransform(kinDyn->getFrameIndex("l_sole"), lDes[i], R, i));
        REQUIRE(desiredKinDyn->getWorldTransform(kinDyn->getFrameIndex("chest"), chest, R, i));
        REQUIRE(comTask->setSetPoint(comDes[i], {0, 0, 0}, i));
        REQUIRE(rFoot->setSetPoint(rDes[i], SE3d::Tangent::Zero(), i) && lFoot->setSetPoint(lDes[i], SE3d::Tangent::Zero(), i));
        REQUIRE(torso->setSetPoint(chest.R, SO3Tangentd::Zero(), i));
        REQUIRE(reg->setSetPoint(sDes[i], i));
    }

    auto system = std::make_shared<FloatingBaseSystemVelocityKinematics>();
    system->setBatchSize(N);
    for (std::size_t i = 0; i < N; i++) {
        VectorXd s0 = sDes[i];
        for (auto& v : s0) v += noise(gen);
        REQUIRE(system->setState({{0.0, 0.0, 0.75}, SO3d::Exp({0.0, 0.05 + noise(gen), 0.1}), s0}, i));
    }
    ForwardEuler<FloatingBaseSystemVelocityKinematics> integrator;
    REQUIRE(integrator.setIntegrationStep(dT));
    REQUIRE(integrator.setDynamicalSystem(system));

    for (int iteration = 0; iteration < 120; iteration++) {
        REQUIRE(kinDyn->setRobotStateBatch(system->basePositions().data(), system->baseQuaternions().data(), system->jointPositions().data()));
        REQUIRE(ik->advance());
        REQUIRE(ik->isOutputValid());
        if (N == 1) {
            REQUIRE(system->setControlInput({ik->getOutput().baseVelocity.coeffs(), ik->getOutput().jointVelocity}));
        } else {
            REQUIRE(system->setControlInputBatch(static_cast<QPInverseKinematics*>(ik.get())->getBatchOutput().data()));
        }
        REQUIRE(integrator.integrate(0s, dT));
        if (withLimits) {
            const auto& nu = static_cast<QPInverseKinematics*>(ik.get())->getBatchOutput();
            double vmax = 0;
            for (std::size_t i = 0; i < N; i++)
                for (std::size_t j = 0; j < n; j++) vmax = std::max(vmax, std::abs(nu[i * (n + 6) + 6 + j]));
            REQUIRE(vmax <= 2.0 + 1e-9);
        }
    }
    REQUIRE(kinDyn->setRobotStateBatch(system->basePositions().data(), system->baseQuaternions().data(), system->jointPositions().data()));
    for (std::size_t i = 0; i < N; i += std::max<std::size_t>(1, N / 7)) {
        Vector3d com;
        SE3d pose;
        std::array<double, 9> R;
        REQUIRE(kinDyn->getCenterOfMassPosition(com, i));
        for (int c = 0; c < 3; c++) REQUIRE(std::abs(com[c] - comDes[i][c]) < 2e-3);
        REQUIRE(kinDyn->getWorldTransform(kinDyn->getFrameIndex("l_sole"), pose, R, i));
        for (int c = 0; c < 3; c++) REQUIRE(std::abs(pose.p[c] - lDes[i].p[c]) < 2e-3);
        const auto qa = pose.quat().coeffs(), qb = lDes[i].quat().coeffs();
        const double dot = std::abs(qa[0] * qb[0] + qa[1] * qb[1] + qa[2] * qb[2] + qa[3] * qb[3]);
        REQUIRE(dot > 1.0 - 1e-5);
    }
    // setTaskWeight takes effect on the next advance() (QPInverseKinematics.cpp:450-451)
    const auto before = static_cast<QPInverseKinematics*>(ik.get())->getBatchOutput();
    REQUIRE(ik->setTaskWeight("JOINT_REGULARIZATION_TASK", VectorXd(n, 50.0)));
    for (std::size_t i = 0; i < N; i++) REQUIRE(comTask->setSetPoint({comDes[i][0] + 0.02, comDes[i][1], comDes[i][2]}, {0, 0, 0}, i));
    REQUIRE(ik->advance());
    REQUIRE(ik->getRawSolution().size() == n + 6);
    REQUIRE(std::abs(ik->getOutput().baseVelocity.coeffs()[0]) + std::abs(ik->getOutput().jointVelocity[0]) > 1e-6);
    (void)before;
}

// IntegratorFloatingBaseSystemVelocityKinematics.cpp:25-88 with a non-identity initial rotation
// QPFixedBaseInverseKinematics (QPFixedBaseInverseKinematicsTest.cpp: the base link is held by "base_se3_task", a hand reaches
// a pose with the joints regularised): the base twist stays zero on every tick and the hand converges
// src/Math/tests/SplineTest.cpp:21-167 "Quintic spline" replayed: 100 knots on a random quintic in R^4 between 320 ms and 642 s;
// the spline returns the polynomial, its first and its second derivative (isApprox 1e-5), through evaluatePoint, advance and
// evaluateOrderedPoints.
template <template <class> class SplineClass, int degree> static void splineSection()
{
    constexpr std::size_t numberOfPoints = 100;
    constexpr std::chrono::nanoseconds initTime = 320ms, finalTime = 2s + 640s;
    constexpr std::chrono::nanoseconds dT = (finalTime - initTime) / (numberOfPoints - 1);
    constexpr double tolerance = 1e-5;
    std::mt19937 gen(5);
    std::uniform_real_distribution<double> uni(-1.0, 1.0);
    double c[6][4];
    for (int i = 0; i < 6; i++)
        for (double& v : c[i]) v = i <= degree ? uni(gen) : 0.0;   // "Cubic spline" (:185-330): knots on a random cubic
    auto poly = [&](double t, int order) {
        VectorXd out(4, 0.0);
        for (int k = 0; k < 4; k++)
            for (int i = order; i < 6; i++) {
                double f = 1.0;
                for (int j = 0; j < order; j++) f *= i - j;
                out[k] += f * c[i][k] * std::pow(t, i - order);
            }
        return out;
    };
    auto approx = [&](const VectorXd& a, const VectorXd& b) {   // Eigen's isApprox
        double d = 0, na = 0, nb = 0;
        for (std::size_t k = 0; k < a.size(); k++) d += (a[k] - b[k]) * (a[k] - b[k]), na += a[k] * a[k], nb += b[k] * b[k];
        return d <= tolerance * tolerance * std::min(na, nb);
    };
    std::vector<VectorXd> knots;
    std::vector<std::chrono::nanoseconds> time;
    for (std::size_t i = 0; i < numberOfPoints; i++) {
        time.push_back(dT * i + initTime);
        knots.push_back(poly(std::chrono::duration<double>(time.back()).count(), 0));
    }
    const double initT = std::chrono::duration<double>(initTime).count(), lastT = std::chrono::duration<double>(time.back()).count();
    SplineClass<VectorXd> spline;
    VectorXd position, velocity, acceleration;
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(spline.evaluatePoint(1s, position));                                        // no knots
    REQUIRE_FALSE(spline.setKnots({knots[0]}, {time[0]}));                                    // Spline.h:338-342
    REQUIRE_FALSE(spline.setKnots(knots, std::vector<std::chrono::nanoseconds>(time.begin(), time.end() - 1)));
    REQUIRE(spline.setKnots(knots, time));
    REQUIRE_FALSE(spline.evaluatePoint(1s, position));                                        // no boundary conditions
    REQUIRE_FALSE(spline.setInitialConditions(VectorXd(4, 0.0), VectorXd(3, 0.0)));
    REQUIRE_FALSE(spline.setAdvanceTimeStep(0ns));
    log()->setLevel(Logger::Level::warn);
    REQUIRE(spline.setInitialConditions({poly(initT, 1), poly(initT, 2)}));
    REQUIRE(spline.setFinalConditions({poly(lastT, 1), poly(lastT, 2)}));
    constexpr std::size_t pointsToCheckNumber = 1000;
    const std::chrono::nanoseconds dTCheckPoints = (time.back() - initTime) / pointsToCheckNumber;
    bool allClose = true;
    for (std::size_t i = 0; i < pointsToCheckNumber; i++) {
        const std::chrono::nanoseconds t = dTCheckPoints * i + initTime;
        const double td = std::chrono::duration<double>(t).count();
        allClose = allClose && spline.evaluatePoint(t, position, velocity, acceleration);
        allClose = allClose && approx(poly(td, 0), position) && approx(poly(td, 1), velocity) && (degree == 3 || approx(poly(td, 2), acceleration));
    }
    REQUIRE(allClose);
    // "Advance capabilities"
    REQUIRE_FALSE(spline.isOutputValid());
    REQUIRE(spline.setAdvanceTimeStep(dTCheckPoints));
    allClose = true;
    for (std::size_t i = 0; i < pointsToCheckNumber; i++) {
        const double td = std::chrono::duration<double>(dTCheckPoints * i + initTime).count();
        allClose = allClose && spline.advance() && spline.isOutputValid();
        const auto& traj = spline.getOutput();
        allClose = allClose && approx(poly(td, 0), traj.position) && approx(poly(td, 1), traj.velocity) && (degree == 3 || approx(poly(td, 2), traj.acceleration));
    }
    REQUIRE(allClose);
    // "Query from a vector of times"
    std::vector<std::chrono::nanoseconds> timeVector;
    for (std::size_t i = 0; i < pointsToCheckNumber; i++) timeVector.push_back(dTCheckPoints * i + initTime);
    std::vector<VectorXd> pv, vv, av;
    REQUIRE(spline.evaluateOrderedPoints(timeVector, pv, vv, av));
    allClose = true;
    for (std::size_t i = 0; i < pointsToCheckNumber; i++) {
        const double td = std::chrono::duration<double>(timeVector[i]).count();
        allClose = allClose && approx(poly(td, 0), pv[i]) && approx(poly(td, 1), vv[i]) && (degree == 3 || approx(poly(td, 2), av[i]));
    }
    REQUIRE(allClose);
    // outside the knots the end states hold (Spline.h:607-625); two equal times are refused (:444-448)
    REQUIRE(spline.evaluatePoint(0ns, position, velocity) && approx(knots.front(), position) && approx(poly(initT, 1), velocity));
    REQUIRE(spline.evaluatePoint(finalTime + 10s, position, velocity) && approx(knots.back(), position) && approx(poly(lastT, 1), velocity));
    SplineClass<Vector3d> twin;
    REQUIRE(twin.setKnots({Vector3d{0, 0, 0}, Vector3d{1, 0, 0}, Vector3d{2, 0, 0}}, {0ms, 500ms, 500ms}));
    REQUIRE(twin.setInitialConditions(Vector3d{0, 0, 0}, Vector3d{0, 0, 0}) && twin.setFinalConditions(Vector3d{0, 0, 0}, Vector3d{0, 0, 0}));
    Vector3d p3;
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(twin.evaluatePoint(200ms, p3));
    log()->setLevel(Logger::Level::warn);
}

// src/Planners/tests/SO3PlannerTest.cpp:21-167 replayed: the rotation integrates its own velocity (right-plus for the body-fixed
// planner, left-plus for the inertial one), the velocity its acceleration, both ends are met (tolerance 1e-3).
template <class Planner> static void so3PlannerSection(bool body, bool initialVelocity)
{
    using namespace BipedalLocomotion::Planners;
    std::mt19937 gen(21);
    std::normal_distribution<double> nrm(0.0, 1.0);
    auto randomRotation = [&]() { return SO3d(Quaterniond{nrm(gen), nrm(gen), nrm(gen), nrm(gen)}); };
    const SO3d R0 = randomRotation(), R1 = randomRotation();
    constexpr std::chrono::nanoseconds T = 1s, dT = 100us;
    constexpr double tolerance = 1e-3, h = 100e-6;
    auto near = [&](const SO3d& a, const SO3d& b) { const Vector3d e = (b.inverse() * a).Log(); return std::sqrt(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]) < tolerance; };
    auto nearV = [&](const Vector3d& a, const Vector3d& b) { return std::abs(a[0] - b[0]) < tolerance && std::abs(a[1] - b[1]) < tolerance && std::abs(a[2] - b[2]) < tolerance; };
    Planner planner;
    SO3Tangentd v0;
    if (initialVelocity) {
        const Vector3d d = (R1 * R0.inverse()).Log();
        v0.w = {2 * d[0], 2 * d[1], 2 * d[2]};
    }
    SO3d rotation;
    SO3Tangentd velocity, acceleration;
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(planner.setRotations(R0, R1, 0s));
    REQUIRE(planner.setRotations(R0, R1, T));
    REQUIRE_FALSE(planner.evaluatePoint(0s, rotation, velocity, acceleration));   // boundary conditions missing
    REQUIRE_FALSE(planner.advance());
    log()->setLevel(Logger::Level::warn);
    REQUIRE(planner.setInitialConditions(v0, SO3Tangentd::Zero()) && planner.setFinalConditions(SO3Tangentd::Zero(), SO3Tangentd::Zero()));
    REQUIRE(planner.evaluatePoint(0s, rotation, velocity, acceleration));
    REQUIRE(near(rotation, R0) && nearV(velocity.w, v0.w) && nearV(acceleration.w, {0, 0, 0}));
    bool ok = true;
    for (std::size_t i = 1; i < static_cast<std::size_t>(T / dT); i++) {
        const SO3d step = SO3d::Exp({velocity.w[0] * h, velocity.w[1] * h, velocity.w[2] * h});
        const SO3d predicted = body ? rotation * step : step * rotation;
        const Vector3d predictedVelocity{velocity.w[0] + acceleration.w[0] * h, velocity.w[1] + acceleration.w[1] * h, velocity.w[2] + acceleration.w[2] * h};
        ok = ok && planner.evaluatePoint(i * dT, rotation, velocity, acceleration);
        ok = ok && near(predicted, rotation) && nearV(predictedVelocity, velocity.w);
    }
    REQUIRE(ok);
    REQUIRE(planner.evaluatePoint(T, rotation, velocity, acceleration));
    REQUIRE(near(rotation, R1) && nearV(velocity.w, {0, 0, 0}) && nearV(acceleration.w, {0, 0, 0}));
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(planner.evaluatePoint(T + 1ms, rotation, velocity, acceleration));   // SO3Planner.tpp:178-184
    const Vector3d d = (R1 * R0.inverse()).Log();
    SO3Tangentd skew;
    skew.w = {d[1] * 0.3 - d[2] * 0.2, d[2] * 1.0 - d[0] * 0.3, d[0] * 0.2 - d[1] * 1.0};   // d x (1, 0.2, 0.3)-like: not along d
    Planner refused;
    REQUIRE(refused.setRotations(R0, R1, T) && refused.setInitialConditions(skew, SO3Tangentd::Zero()) && refused.setFinalConditions(SO3Tangentd::Zero(), SO3Tangentd::Zero()));
    REQUIRE_FALSE(refused.evaluatePoint(500ms, rotation, velocity, acceleration));
    log()->setLevel(Logger::Level::warn);
    REQUIRE(planner.setAdvanceTimeStep(250ms) && planner.isOutputValid());
    for (int i = 0; i < 6; i++) REQUIRE(planner.advance());
    REQUIRE(near(planner.getOutput().rotation, R1));
}

static void testSO3Planner();
// Planners::SwingFootPlanner, replaying src/Planners/tests/SwingFootPlannerTest.cpp (the reference's test only requires that
// advance() succeeds over the whole contact list, in its three scenarios) plus what that trajectory must satisfy: the contact
// pose while in contact, continuity, the apex height, and -- the property the device hand-over rests on -- that re-planning on
// every advance() gives the curve planned once at take-off (planOnDevice).
static Contacts::ContactList swingContactList(bool modified)
{
    Contacts::ContactList list;
    auto rz = [](double a) { return SO3d::Exp({0.0, 0.0, a}); };
    REQUIRE(list.addContact(SE3d({0.9, 0, 0}, rz(1.5708)), 0s, 300ms));
    REQUIRE(list.addContact(SE3d({0.8899, 0.1345, 0.0600}, rz(1.7208)), 600ms, 1s + 500ms));
    if (modified) REQUIRE(list.addContact(SE3d({0.5104, 0.415, 0.400}, rz(2.4)), 1s + 800ms, 2s + 700ms));
    else REQUIRE(list.addContact(SE3d({0.8104, 0.3915, 0.1800}, rz(2.0208)), 1s + 800ms, 2s + 700ms));
    REQUIRE(list.addContact(SE3d({0.6585, 0.6135, 0.3000}, rz(2.3208)), 3s, 3s + 900ms));
    REQUIRE(list.addContact(SE3d({0.3261, 0.8388, 0.4800}, rz(2.7708)), 4s, 6s));
    return list;
}

static void testSwingFootPlanner()
{
    using namespace BipedalLocomotion::Planners;
    constexpr std::chrono::nanoseconds dT = 10ms;
    for (const char* method : {"min_acceleration", "min_jerk"}) {
        auto handler = std::make_shared<StdImplementation>();
        handler->setParameter("sampling_time", dT), handler->setParameter("step_height", 0.1), handler->setParameter("foot_apex_time", 0.5);
        handler->setParameter("foot_landing_velocity", 0.0), handler->setParameter("foot_landing_acceleration", 0.0);
        handler->setParameter("interpolation_method", std::string(method));
        const auto list = swingContactList(false), modifiedList = swingContactList(true);
        for (int scenario = 0; scenario < 3; scenario++) {   // constant list, modified during contact, modified during swing
            SwingFootPlanner planner;
            REQUIRE(planner.initialize(handler));
            REQUIRE(planner.setContactList(list));
            SwingFootPlannerState prev;
            double apex = -1.0;
            SwingFootPlanner::DeviceSwing swing;
            REQUIRE(planner.planOnDevice(list[0], swing));
            std::unique_ptr<Math::Spline<Vector3d>> once;
            if (swing.minJerk) once = std::make_unique<Math::QuinticSpline<Vector3d>>();
            else once = std::make_unique<Math::CubicSpline<Vector3d>>();
            REQUIRE(once->setInitialConditions(swing.initialVelocity, swing.initialAcceleration) && once->setFinalConditions(swing.finalVelocity, swing.finalAcceleration));
            REQUIRE(once->setKnots(swing.positions, swing.times));
            SO3PlannerInertial onceR;
            REQUIRE(onceR.setInitialConditions(SO3Tangentd::Zero(), SO3Tangentd::Zero()) && onceR.setFinalConditions(SO3Tangentd::Zero(), SO3Tangentd::Zero()));
            REQUIRE(onceR.setRotations(list[0].pose.R, swing.landingRotation, swing.duration));
            for (std::chrono::nanoseconds time = list.firstContact()->activationTime; time < list.lastContact()->deactivationTime; time += dT) {
                if (scenario == 1 && time == 1s) REQUIRE(planner.setContactList(modifiedList));
                if (scenario == 2 && time == 1s + 600ms) REQUIRE(planner.setContactList(modifiedList));
                REQUIRE(planner.advance());
                REQUIRE(planner.isOutputValid());
                const auto& out = planner.getOutput();
                REQUIRE(out.time == time);
                const auto& use = (scenario > 0 && time >= (scenario == 1 ? 1s : 1s + 600ms)) ? modifiedList : list;
                const auto active = use.getActiveContact(time);
                REQUIRE(out.isInContact == (active != use.cend()));
                if (out.isInContact) {
                    for (int c = 0; c < 3; c++) REQUIRE(out.transform.p[c] == active->pose.p[c] && out.mixedVelocity.linear[c] == 0.0);
                } else if (time > list.firstContact()->activationTime) {
                    for (int c = 0; c < 3; c++) {   // one sample apart: the foot moves by what its speed allows (trapezoid + curvature)
                        const double moved = std::abs(out.transform.p[c] - prev.transform.p[c]);
                        const double speed = std::max(std::abs(out.mixedVelocity.linear[c]), std::abs(prev.mixedVelocity.linear[c]));
                        REQUIRE(moved <= speed * 0.01 * 1.05 + 1e-4);
                    }
                }
                if (time >= 300ms && time < 600ms) {   // the first swing: the curve re-planned every period is the one planned at take-off
                    Vector3d p, v;
                    REQUIRE(once->evaluatePoint(time, p, v));
                    for (int c = 0; c < 3; c++) REQUIRE(std::abs(out.transform.p[c] - p[c]) < 1e-9 && std::abs(out.mixedVelocity.linear[c] - v[c]) < 1e-7);
                    apex = std::max(apex, out.transform.p[2]);
                    SO3d R;
                    SO3Tangentd w, dw;
                    REQUIRE(onceR.evaluatePoint(time - swing.times[0], R, w, dw));
                    const Vector3d dr = (R.inverse() * out.transform.R).Log();
                    for (int c = 0; c < 3; c++) REQUIRE(std::abs(dr[c]) < 1e-9 && std::abs(w.w[c] - out.mixedVelocity.angular[c]) < 1e-7);
                }
                prev = out;
            }
            REQUIRE(std::abs(apex - 0.16) < 1e-3);   // step_height above the higher contact (z = 0.06), reached at half swing
            log()->setLevel(Logger::Level::off);
            REQUIRE_FALSE(planner.advance());        // past the last contact (:503-509)
            log()->setLevel(Logger::Level::warn);
        }
    }
    log()->setLevel(Logger::Level::off);
    SwingFootPlanner planner;
    auto bad = std::make_shared<StdImplementation>();
    bad->setParameter("sampling_time", dT), bad->setParameter("step_height", 0.1), bad->setParameter("foot_apex_time", 1.5);
    REQUIRE_FALSE(planner.initialize(bad));
    REQUIRE_FALSE(planner.advance());                                      // empty contact list
    Contacts::ContactList list;
    REQUIRE_FALSE(list.addContact(SE3d(), 2s, 1s));                         // activation after deactivation
    REQUIRE(list.addContact(SE3d(), 0s, 1s));
    REQUIRE_FALSE(list.addContact(SE3d(), 500ms, 2s));                      // overlap
    log()->setLevel(Logger::Level::warn);
}

static void testSO3Planner()
{
    using namespace BipedalLocomotion::Planners;
    so3PlannerSection<SO3PlannerBody>(true, false);
    so3PlannerSection<SO3PlannerInertial>(false, false);
    so3PlannerSection<SO3PlannerInertial>(false, true);
}

static void testSplines()
{
    splineSection<BipedalLocomotion::Math::QuinticSpline, 5>();
    splineSection<BipedalLocomotion::Math::CubicSpline, 3>();
}

// IKLinearTask::setSetPointSpline: the hand of N fixed-base robots follows one spline per robot, evaluated on the device at
// the solver's time; the solution equals the reference's way (evaluate on the host, setSetPoint, advance) at every time.
static void testSetPointSpline(const char* modelPath, std::size_t N)
{
    using BipedalLocomotion::Math::QuinticSpline;
    auto make = [&](std::shared_ptr<KinDynBatch>& kinDyn, std::shared_ptr<SE3Task>& hand, QPFixedBaseInverseKinematics& ik) {
        kinDyn = loadKinDyn(modelPath, N);
        const std::size_t n = kinDyn->getNrOfDegreesOfFreedom();
        auto ikHandler = std::make_shared<StdImplementation>();
        ikHandler->setParameter("robot_velocity_variable_name", robotVelocity);
        REQUIRE(ik.setKinDyn(kinDyn) && ik.initialize(ikHandler));
        hand = std::make_shared<SE3Task>();
        auto hh = std::make_shared<StdImplementation>();
        hh->setParameter("robot_velocity_variable_name", robotVelocity), hh->setParameter("frame_name", "l_hand_palm");
        hh->setParameter("kp_linear", 5.0), hh->setParameter("kp_angular", 5.0);
        REQUIRE(hand->setKinDyn(kinDyn) && hand->initialize(hh) && ik.addTask(hand, "hand", 0));
        auto reg = std::make_shared<JointTrackingTask>();
        auto rh = std::make_shared<StdImplementation>();
        rh->setParameter("robot_velocity_variable_name", robotVelocity), rh->setParameter("kp", std::vector<double>(n, 1.0));
        REQUIRE(reg->setKinDyn(kinDyn) && reg->initialize(rh) && ik.addTask(reg, "reg", 1, VectorXd(n, 1.0)));
        VariablesHandler vh;
        vh.addVariable(robotVelocity, n + 6);
        REQUIRE(ik.finalize(vh));
        std::mt19937 gen(3);
        for (std::size_t i = 0; i < N; i++) {
            const VectorXd s = nominalJoints(*kinDyn, gen, 0.2);
            REQUIRE(kinDyn->setRobotState({0.0, 0.0, 0.7}, SO3d::Identity(), s, i) && reg->setSetPoint(s, i));
        }
    };
    std::shared_ptr<KinDynBatch> kdDev, kdHost;
    std::shared_ptr<SE3Task> handDev, handHost;
    QPFixedBaseInverseKinematics ikDev, ikHost;
    make(kdDev, handDev, ikDev), make(kdHost, handHost, ikHost);
    std::mt19937 gen(17);
    std::uniform_real_distribution<double> uni(-0.01, 0.01);
    std::vector<QuinticSpline<Vector3d>> splines(N);
    std::vector<SE3d> pose(N);
    const std::vector<std::chrono::nanoseconds> time{0ms, 300ms, 450ms, 1000ms};
    for (std::size_t i = 0; i < N; i++) {
        std::array<double, 9> R;
        REQUIRE(kdDev->getWorldTransform(kdDev->getFrameIndex("l_hand_palm"), pose[i], R, i));
        REQUIRE(handDev->setSetPoint(pose[i], SE3d::Tangent::Zero(), i));
        std::vector<Vector3d> knots(4, pose[i].translation());
        for (std::size_t k = 1; k < 4; k++)
            for (int c = 0; c < 3; c++) knots[k][c] += uni(gen) * k;
        REQUIRE(splines[i].setKnots(knots, time));
        REQUIRE(splines[i].setInitialConditions(Vector3d{uni(gen), uni(gen), uni(gen)}, Vector3d{0, 0, 0}));
        REQUIRE(splines[i].setFinalConditions(Vector3d{0, 0, 0}, Vector3d{0, 0, 0}));
    }
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(handDev->setSetPointSpline(std::vector<QuinticSpline<Vector3d>>(N + 1)));   // one per robot
    REQUIRE_FALSE(std::make_shared<SO3Task>()->setSetPointSpline(splines));                   // no position set-point
    log()->setLevel(Logger::Level::warn);
    REQUIRE(handDev->setSetPointSpline(splines));
    double worst = 0.0;
    for (const std::chrono::nanoseconds t : {-100ms, 0ms, 120ms, 300ms, 777ms, 1000ms, 1400ms}) {
        REQUIRE(ikDev.setTime(t) && ikDev.advance());
        for (std::size_t i = 0; i < N; i++) {
            Vector3d p, v;
            REQUIRE(splines[i].evaluatePoint(t, p, v));
            SE3d::Tangent ff = SE3d::Tangent::Zero();
            ff.linear = v;
            REQUIRE(handHost->setSetPoint(SE3d(p, pose[i].R), ff, i));
        }
        REQUIRE(ikHost.advance());
        for (std::size_t k = 0; k < ikDev.getBatchOutput().size(); k++) worst = std::max(worst, std::abs(ikDev.getBatchOutput()[k] - ikHost.getBatchOutput()[k]));
    }
    REQUIRE(worst < 1e-8);
    handDev->clearSetPointSpline();
    REQUIRE(ikDev.advance());
    // setSetPointMinimumJerk: from the stored pose to a turned and shifted one over 0.4 s; the device generator against
    // SO3PlannerInertial (rotation) and the two-knot minimum-jerk profile (position) evaluated on the host
    {
        using namespace BipedalLocomotion::Planners;
        std::vector<double> target(N * 13, 0.0);
        std::vector<SO3PlannerInertial> planners(N);
        std::vector<Vector3d> shift(N);
        for (std::size_t i = 0; i < N; i++) {
            const SO3d R1 = SO3d::Exp({0.1 * (i + 1), -0.05, 0.2}) * pose[i].R;
            shift[i] = {0.01 * (i + 1), -0.02, 0.015};
            const auto q = R1.quat().coeffs();
            for (int c = 0; c < 3; c++) target[i * 13 + c] = pose[i].translation()[c] + shift[i][c];
            for (int c = 0; c < 4; c++) target[i * 13 + 3 + c] = q[c];
            REQUIRE(planners[i].setRotations(pose[i].R, R1, 400ms));
            REQUIRE(planners[i].setInitialConditions(SO3Tangentd::Zero(), SO3Tangentd::Zero()) && planners[i].setFinalConditions(SO3Tangentd::Zero(), SO3Tangentd::Zero()));
            REQUIRE(handHost->setSetPoint(pose[i], SE3d::Tangent::Zero(), i));
        }
        log()->setLevel(Logger::Level::off);
        REQUIRE_FALSE(std::make_shared<JointTrackingTask>()->setSetPointMinimumJerk(target, 400ms));
        REQUIRE_FALSE(handDev->setSetPointMinimumJerk(std::vector<double>(5, 0.0), 400ms));
        log()->setLevel(Logger::Level::warn);
        REQUIRE(handDev->setSetPointMinimumJerk(target, 400ms));
        REQUIRE(ikDev.setTime(3s) && ikDev.advance() && ikHost.advance());   // tau = 0
        double worstMj = 0.0;
        for (const std::chrono::nanoseconds dtau : {100ms, 250ms, 400ms, 700ms}) {
            REQUIRE(ikDev.setTime(3s + dtau) && ikDev.advance());
            const double tau = std::min(1.0, std::chrono::duration<double>(dtau).count() / 0.4);
            const double sv = tau * tau * tau * (10.0 - 15.0 * tau + 6.0 * tau * tau), sd = tau < 1.0 ? 30.0 * tau * tau * (1 - tau) * (1 - tau) / 0.4 : 0.0;
            for (std::size_t i = 0; i < N; i++) {
                SO3PlannerState st;
                REQUIRE(planners[i].evaluatePoint(std::min(dtau, std::chrono::nanoseconds(400ms)), st));
                SE3d::Tangent ff = SE3d::Tangent::Zero();
                Vector3d p = pose[i].translation();
                for (int c = 0; c < 3; c++) p[c] += sv * shift[i][c], ff.linear[c] = sd * shift[i][c];
                ff.angular = st.velocity.w;
                REQUIRE(handHost->setSetPoint(SE3d(p, st.rotation), ff, i));
            }
            REQUIRE(ikHost.advance());
            for (std::size_t k = 0; k < ikDev.getBatchOutput().size(); k++) worstMj = std::max(worstMj, std::abs(ikDev.getBatchOutput()[k] - ikHost.getBatchOutput()[k]));
        }
        REQUIRE(worstMj < 1e-8);
        REQUIRE(handDev->setSetPointMinimumJerk({}, 0ns) && ikDev.advance());
    }
    // A swing phase handed to the device at take-off (SwingFootPlanner::planOnDevice: one 3D spline for the origin + the landing
    // rotation for the minimum-jerk generator) against the reference's loop: planner.advance() every period on the host, its
    // output given to setSetPoint.  One planner per robot, same contact timing, different landing spots.
    {
        using namespace BipedalLocomotion::Planners;
        constexpr std::chrono::nanoseconds dT = 10ms;
        auto ph = std::make_shared<StdImplementation>();
        ph->setParameter("sampling_time", dT), ph->setParameter("step_height", 0.03), ph->setParameter("foot_apex_time", 0.4);
        ph->setParameter("interpolation_method", std::string("min_jerk"));
        std::vector<SwingFootPlanner> planners(N);
        std::vector<QuinticSpline<Vector3d>> origin(N);
        std::vector<double> landing(N * 13, 0.0);
        std::chrono::nanoseconds duration{0};
        for (std::size_t i = 0; i < N; i++) {
            Contacts::ContactList list;
            const SE3d land({pose[i].p[0] + 0.02 * (i + 1), pose[i].p[1] - 0.015, pose[i].p[2] + 0.01}, SO3d::Exp({0.03, 0.05, 0.1 * (i + 1)}) * pose[i].R);
            // (every component of the rotation error non-zero: SO3Planner finds the scalar between a boundary velocity and the error
            // from the first component whose ratio is finite, SO3Planner.tpp:81-89 -- with an error component that is rounding noise
            // that ratio is noise too, in the reference as in the host class that mirrors it; the device generator plans once)
            REQUIRE(list.addContact(pose[i], 0s, 300ms) && list.addContact(land, 600ms, 2s));
            REQUIRE(planners[i].initialize(ph) && planners[i].setContactList(list));
            SwingFootPlanner::DeviceSwing swing;
            REQUIRE(planners[i].planOnDevice(list[0], swing) && swing.minJerk);
            REQUIRE(origin[i].setInitialConditions(swing.initialVelocity, swing.initialAcceleration) && origin[i].setFinalConditions(swing.finalVelocity, swing.finalAcceleration));
            REQUIRE(origin[i].setKnots(swing.positions, swing.times));
            const auto q = swing.landingRotation.quat().coeffs();
            for (int c = 0; c < 3; c++) landing[i * 13 + c] = land.p[c];
            for (int c = 0; c < 4; c++) landing[i * 13 + 3 + c] = q[c];
            duration = swing.duration;
            REQUIRE(handDev->setSetPoint(pose[i], SE3d::Tangent::Zero(), i));
        }
        REQUIRE(duration == 310ms);                                            // one period before the contact ends (createSE3Traj's previousTime)
        REQUIRE(handDev->setSetPointSpline(origin) && handDev->setSetPointMinimumJerk(landing, duration));
        REQUIRE(ikDev.setTime(290ms) && ikDev.advance());                      // the generators start here: tau = 0
        double worstSwing = 0.0;
        std::chrono::nanoseconds hostTime = 0ns;
        for (const std::chrono::nanoseconds t : {300ms, 350ms, 420ms, 520ms, 590ms, 600ms, 800ms}) {
            for (; hostTime <= t; hostTime += dT)
                for (std::size_t i = 0; i < N; i++) REQUIRE(planners[i].advance());   // the planner is sampled every period, as in the reference's loop
            for (std::size_t i = 0; i < N; i++) {
                const auto& out = planners[i].getOutput();
                REQUIRE(out.time == t);
                REQUIRE(handHost->setSetPoint(out.transform, out.mixedVelocity, i));
            }
            REQUIRE(ikHost.advance());
            REQUIRE(ikDev.setTime(t) && ikDev.advance());
            double wt = 0.0;
            for (std::size_t k = 0; k < ikDev.getBatchOutput().size(); k++) wt = std::max(wt, std::abs(ikDev.getBatchOutput()[k] - ikHost.getBatchOutput()[k]));
            if (std::getenv("BLF_TEST_VERBOSE")) {
                std::printf("swing t=%lld ms: max |dnu| %.3e\n", static_cast<long long>(t.count() / 1000000), wt);
                for (int task = 0; task < 3; task++) {
                    if (blfik_setpoint_len(ikDev.handle(), task) != 13) continue;
                    std::vector<double> sp(N * 13);
                    if (blfik_evaluate_setpoint(ikDev.handle(), task, std::chrono::duration<double>(t).count(), sp.data(), BLFIK_MEM_HOST) != BLFIK_OK) continue;
                    const auto& out = planners[0].getOutput();
                    const auto q = out.transform.R.quat().coeffs();
                    std::printf("  task %d device sp: p %.6f %.6f %.6f q %.6f %.6f %.6f %.6f v %.4f %.4f %.4f w %.4f %.4f %.4f\n", task, sp[0], sp[1], sp[2], sp[3], sp[4], sp[5], sp[6], sp[7], sp[8], sp[9], sp[10], sp[11], sp[12]);
                    std::printf("         host planner: p %.6f %.6f %.6f q %.6f %.6f %.6f %.6f v %.4f %.4f %.4f w %.4f %.4f %.4f\n", out.transform.p[0], out.transform.p[1], out.transform.p[2], q[0], q[1], q[2], q[3],
                                out.mixedVelocity.linear[0], out.mixedVelocity.linear[1], out.mixedVelocity.linear[2], out.mixedVelocity.angular[0], out.mixedVelocity.angular[1], out.mixedVelocity.angular[2]);
                }
            }
            worstSwing = std::max(worstSwing, wt);
        }
        REQUIRE(worstSwing < 1e-7);
        handDev->clearSetPointSpline();
        REQUIRE(handDev->setSetPointMinimumJerk({}, 0ns) && ikDev.advance());
    }
}

// A JointTrackingTask whose joint positions follow a Math::QuinticSpline<VectorXd> / CubicSpline<VectorXd>: evaluated on the device
// inside the tick (setSetPointSpline) against the reference's way -- the spline evaluated on the host, its point handed to
// setSetPoint(position, velocity) every period.
template <class SplineT> static void testJointSetPointSpline(const char* modelPath, std::size_t N)
{
    struct Solver {
        std::shared_ptr<KinDynBatch> kinDyn;
        std::shared_ptr<JointTrackingTask> reg;
        QPFixedBaseInverseKinematics ik;
    };
    std::vector<VectorXd> s0(N);
    auto make = [&](Solver& S) {
        S.kinDyn = loadKinDyn(modelPath, N);
        const std::size_t n = S.kinDyn->getNrOfDegreesOfFreedom();
        auto ikHandler = std::make_shared<StdImplementation>();
        ikHandler->setParameter("robot_velocity_variable_name", robotVelocity);
        REQUIRE(S.ik.setKinDyn(S.kinDyn) && S.ik.initialize(ikHandler));
        S.reg = std::make_shared<JointTrackingTask>();
        auto rh = std::make_shared<StdImplementation>();
        rh->setParameter("robot_velocity_variable_name", robotVelocity), rh->setParameter("kp", std::vector<double>(n, 2.0));
        REQUIRE(S.reg->setKinDyn(S.kinDyn) && S.reg->initialize(rh) && S.ik.addTask(S.reg, "reg", 1, VectorXd(n, 1.0)));
        VariablesHandler vh;
        vh.addVariable(robotVelocity, n + 6);
        REQUIRE(S.ik.finalize(vh));
        std::mt19937 gen(3);
        for (std::size_t i = 0; i < N; i++) {
            s0[i] = nominalJoints(*S.kinDyn, gen, 0.2);
            REQUIRE(S.kinDyn->setRobotState({0.0, 0.0, 0.7}, SO3d::Identity(), s0[i], i) && S.reg->setSetPoint(s0[i], i));
        }
    };
    Solver dev, host;
    make(dev), make(host);
    const std::size_t n = dev.kinDyn->getNrOfDegreesOfFreedom();
    std::mt19937 gen(23);
    std::uniform_real_distribution<double> uni(-0.05, 0.05);
    std::vector<SplineT> splines(N);
    const std::vector<std::chrono::nanoseconds> time{0ms, 250ms, 600ms, 1000ms};
    for (std::size_t i = 0; i < N; i++) {
        std::vector<VectorXd> knots(4, s0[i]);
        for (std::size_t k = 1; k < 4; k++)
            for (std::size_t j = 0; j < n; j++) knots[k][j] += uni(gen) * k;
        VectorXd v0(n), zero(n, 0.0);
        for (auto& v : v0) v = uni(gen);
        REQUIRE(splines[i].setKnots(knots, time));
        REQUIRE(splines[i].setInitialConditions(v0, zero) && splines[i].setFinalConditions(zero, zero));
    }
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(dev.reg->setSetPointSpline(std::vector<SplineT>(N + 1)));                              // one per robot
    REQUIRE_FALSE(dev.reg->setSetPointSpline(std::vector<BipedalLocomotion::Math::QuinticSpline<Vector3d>>(N)));   // no knots
    {
        std::vector<BipedalLocomotion::Math::QuinticSpline<Vector3d>> wrong(N);
        for (auto& w : wrong) REQUIRE(w.setKnots({Vector3d{0, 0, 0}, Vector3d{1, 1, 1}}, {0ms, 100ms}));
        REQUIRE_FALSE(dev.reg->setSetPointSpline(wrong));                                                // three coordinates, n joints
    }
    REQUIRE_FALSE(dev.reg->setSetPointMinimumJerk(std::vector<double>(N * 2 * n, 0.0), 100ms));          // no such generator in joint space
    log()->setLevel(Logger::Level::warn);
    REQUIRE(dev.reg->setSetPointSpline(splines));
    double worst = 0.0;
    for (const std::chrono::nanoseconds t : {-100ms, 0ms, 120ms, 250ms, 777ms, 1000ms, 1400ms}) {
        REQUIRE(dev.ik.setTime(t) && dev.ik.advance());
        for (std::size_t i = 0; i < N; i++) {
            VectorXd p, v, a;
            REQUIRE(splines[i].evaluatePoint(t, p, v, a));
            REQUIRE(host.reg->setSetPoint(p, v, i));
        }
        REQUIRE(host.ik.advance());
        for (std::size_t k = 0; k < dev.ik.getBatchOutput().size(); k++) worst = std::max(worst, std::abs(dev.ik.getBatchOutput()[k] - host.ik.getBatchOutput()[k]));
    }
    REQUIRE(worst < 1e-8);
    dev.reg->clearSetPointSpline();
    REQUIRE(dev.ik.advance());
}

static void testFixedBase(const char* modelPath, std::size_t N)
{
    constexpr auto dT = 10ms;
    auto kinDyn = loadKinDyn(modelPath, N);
    const std::size_t n = kinDyn->getNrOfDegreesOfFreedom();
    REQUIRE(kinDyn->getFloatingBase() == "root_link");
    auto ikHandler = std::make_shared<StdImplementation>();
    ikHandler->setParameter("robot_velocity_variable_name", robotVelocity);
    QPFixedBaseInverseKinematics ik;
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(ik.initialize(ikHandler));   // setKinDyn first (QPFixedBaseInverseKinematics.cpp:96-100)
    log()->setLevel(Logger::Level::warn);
    REQUIRE(ik.setKinDyn(kinDyn) && ik.initialize(ikHandler));
    REQUIRE(ik.getTaskNames().size() == 1 && ik.getTask("base_se3_task").lock() != nullptr);

    auto hand = std::make_shared<SE3Task>();
    auto hh = std::make_shared<StdImplementation>();
    hh->setParameter("robot_velocity_variable_name", robotVelocity), hh->setParameter("frame_name", "l_hand_palm");
    hh->setParameter("kp_linear", 5.0), hh->setParameter("kp_angular", 5.0);
    REQUIRE(hand->setKinDyn(kinDyn) && hand->initialize(hh) && ik.addTask(hand, "hand", 0));
    auto reg = std::make_shared<JointTrackingTask>();
    auto rh = std::make_shared<StdImplementation>();
    rh->setParameter("robot_velocity_variable_name", robotVelocity), rh->setParameter("kp", std::vector<double>(n, 1.0));
    REQUIRE(reg->setKinDyn(kinDyn) && reg->initialize(rh) && ik.addTask(reg, "reg", 1, VectorXd(n, 1.0)));
    VariablesHandler vh;
    vh.addVariable(robotVelocity, n + 6);
    REQUIRE(ik.finalize(vh));

    std::mt19937 gen(11);
    auto desired = loadKinDyn(modelPath, N);
    auto system = std::make_shared<FloatingBaseSystemVelocityKinematics>();
    system->setBatchSize(N);
    std::vector<SE3d> handDes(N);
    const Vector3d p0{0.1, -0.2, 0.8};
    const SO3d R0 = SO3d::Exp({0.05, -0.1, 0.3});
    for (std::size_t i = 0; i < N; i++) {
        VectorXd sDes = nominalJoints(*kinDyn, gen, 0.2), s0 = sDes;
        std::normal_distribution<double> noise(0.0, 0.08);
        for (auto& v : s0) v += noise(gen);
        std::array<double, 9> R;
        REQUIRE(desired->setRobotState(p0, R0, sDes, i));
        REQUIRE(desired->getWorldTransform(kinDyn->getFrameIndex("l_hand_palm"), handDes[i], R, i));
        REQUIRE(hand->setSetPoint(handDes[i], SE3d::Tangent::Zero(), i) && reg->setSetPoint(sDes, i));
        REQUIRE(system->setState({p0, R0, s0}, i));
    }
    ForwardEuler<FloatingBaseSystemVelocityKinematics> integrator;
    REQUIRE(integrator.setIntegrationStep(dT) && integrator.setDynamicalSystem(system));
    for (int iteration = 0; iteration < 150; iteration++) {
        REQUIRE(kinDyn->setRobotStateBatch(system->basePositions().data(), system->baseQuaternions().data(), system->jointPositions().data()));
        REQUIRE(ik.advance() && ik.isOutputValid());
        const auto& nu = ik.getBatchOutput();
        double baseTwist = 0;
        for (std::size_t i = 0; i < N; i++)
            for (int c = 0; c < 6; c++) baseTwist = std::max(baseTwist, std::abs(nu[i * (n + 6) + c]));
        REQUIRE(baseTwist < 1e-9);
        REQUIRE(system->setControlInputBatch(nu.data()));
        REQUIRE(integrator.integrate(0s, dT));
    }
    REQUIRE(kinDyn->setRobotStateBatch(system->basePositions().data(), system->baseQuaternions().data(), system->jointPositions().data()));
    for (std::size_t i = 0; i < N; i++) {
        SE3d pose;
        std::array<double, 9> R;
        REQUIRE(kinDyn->getWorldTransform(kinDyn->getFrameIndex("l_hand_palm"), pose, R, i));
        for (int c = 0; c < 3; c++) REQUIRE(std::abs(pose.p[c] - handDes[i].p[c]) < 2e-3);
        for (int c = 0; c < 3; c++) REQUIRE(std::abs(system->basePositions()[3 * i + c] - p0[c]) < 1e-9);   // the base never moved
    }
}

// KinDynBatch's dynamic quantities (what the TSID tasks read from kinDyn): mass matrix, gravity and bias forces, bias accelerations.
// Checked by the physics that ties them together, on a batch of 3: M symmetric positive definite with the total mass on the base
// translation block; bias forces at zero velocity = gravity forces; the Coriolis part is quadratic in the velocity; frame velocity
// = J nu; and along a motion with constant nu (MIXED) the frame velocity changes by Jdot nu dt (finite difference through
// ForwardEuler on the device).
static void testKinDynDynamics(const char* modelPath)
{
    const std::size_t N = 3;
    auto kinDyn = loadKinDyn(modelPath, N);
    const std::size_t n = kinDyn->getNrOfDegreesOfFreedom(), nv = n + 6;
    std::mt19937 gen(31);
    std::uniform_real_distribution<double> uni(-1.0, 1.0);
    std::vector<VectorXd> s(N), ds(N);
    std::vector<std::array<double, 6>> vb(N);
    std::vector<SO3d> R(N);
    std::vector<Vector3d> p(N);
    for (std::size_t i = 0; i < N; i++) {
        s[i] = nominalJoints(*kinDyn, gen, 0.3);
        ds[i].resize(n);
        for (auto& v : ds[i]) v = uni(gen);
        for (auto& v : vb[i]) v = uni(gen);
        R[i] = SO3d::Exp({uni(gen), uni(gen), uni(gen)}), p[i] = {uni(gen), uni(gen), 0.7};
        REQUIRE(kinDyn->setRobotState(p[i], R[i], s[i], i) && kinDyn->setRobotVelocity(vb[i], ds[i], i));
    }
    kinDyn->setGravityAcceleration(9.81);
    const int hand = kinDyn->getFrameIndex("l_hand_palm");
    REQUIRE(hand >= 0);
    double totalMass = 0.0;
    for (std::size_t i = 0; i < N; i++) {
        System::MatrixXd M, J;
        VectorXd g, h;
        REQUIRE(kinDyn->getFreeFloatingMassMatrix(M, i) && kinDyn->generalizedGravityForces(g, i) && kinDyn->generalizedBiasForces(h, i));
        REQUIRE(M.rows == nv && M.cols == nv && g.size() == nv && h.size() == nv);
        if (i == 0) totalMass = M(0, 0);
        REQUIRE(totalMass > 1.0 && std::abs(M(1, 1) - totalMass) < 1e-9 && std::abs(M(0, 1)) < 1e-9 && std::abs(M(0, 0) - totalMass) < 1e-9);
        REQUIRE(std::abs(g[2] - 9.81 * totalMass) < 1e-8 && std::abs(g[0]) < 1e-10);
        double asym = 0.0, mind = 1e300;
        for (std::size_t a = 0; a < nv; a++) {
            mind = std::min(mind, M(a, a));
            for (std::size_t b = 0; b < nv; b++) asym = std::max(asym, std::abs(M(a, b) - M(b, a)));
        }
        REQUIRE(asym < 1e-10 && mind > 0.0);
        std::array<double, 6> v, a;
        REQUIRE(kinDyn->getFrameVel(hand, v, i) && kinDyn->getFrameBiasAcc(hand, a, i) && kinDyn->getFrameFreeFloatingJacobian(hand, J, i));
        for (int r = 0; r < 6; r++) {
            double acc = 0.0;
            for (std::size_t c = 0; c < nv; c++) acc += J(r, c) * (c < 6 ? vb[i][c] : ds[i][c - 6]);
            REQUIRE(std::abs(acc - v[r]) < 1e-11);
        }
    }
    // zero velocity: bias = gravity; twice the velocity: four times the Coriolis part
    std::vector<VectorXd> h1(N), g1(N);
    for (std::size_t i = 0; i < N; i++) REQUIRE(kinDyn->generalizedBiasForces(h1[i], i) && kinDyn->generalizedGravityForces(g1[i], i));
    for (std::size_t i = 0; i < N; i++) {
        std::array<double, 6> v2 = vb[i];
        VectorXd d2 = ds[i];
        for (auto& v : v2) v *= 2.0;
        for (auto& v : d2) v *= 2.0;
        REQUIRE(kinDyn->setRobotVelocity(v2, d2, i));
    }
    for (std::size_t i = 0; i < N; i++) {
        VectorXd h2;
        REQUIRE(kinDyn->generalizedBiasForces(h2, i));
        for (std::size_t c = 0; c < nv; c++) REQUIRE(std::abs((h2[c] - g1[i][c]) - 4.0 * (h1[i][c] - g1[i][c])) < 1e-8 * (1.0 + std::abs(h1[i][c])));
        REQUIRE(kinDyn->setRobotVelocity({0, 0, 0, 0, 0, 0}, VectorXd(n, 0.0), i));
    }
    for (std::size_t i = 0; i < N; i++) {
        VectorXd h0;
        Vector3d cb;
        REQUIRE(kinDyn->generalizedBiasForces(h0, i) && kinDyn->getCenterOfMassBiasAcc(cb, i));
        for (std::size_t c = 0; c < nv; c++) REQUIRE(std::abs(h0[c] - g1[i][c]) < 1e-9 * (1.0 + std::abs(g1[i][c])));
        REQUIRE(std::abs(cb[0]) + std::abs(cb[1]) + std::abs(cb[2]) == 0.0);
    }
    // Jdot nu against the change of J nu along a motion with nu held (central difference, ForwardEuler on the device)
    const double d = 1e-5;
    for (std::size_t i = 0; i < N; i++) REQUIRE(kinDyn->setRobotVelocity(vb[i], ds[i], i));
    std::vector<std::array<double, 6>> bias(N), vp(N), vm(N);
    for (std::size_t i = 0; i < N; i++) REQUIRE(kinDyn->getFrameBiasAcc(hand, bias[i], i));
    for (int sign : {1, -1}) {
        auto system = std::make_shared<FloatingBaseSystemVelocityKinematics>();
        system->setBatchSize(N);
        for (std::size_t i = 0; i < N; i++) {
            std::array<double, 6> tw = vb[i];
            VectorXd sd = ds[i];
            for (auto& v : tw) v *= sign;
            for (auto& v : sd) v *= sign;
            REQUIRE(system->setState({p[i], R[i], s[i]}, static_cast<std::ptrdiff_t>(i)));
            REQUIRE(system->setControlInput({tw, sd}, static_cast<std::ptrdiff_t>(i)));
        }
        ForwardEuler<FloatingBaseSystemVelocityKinematics> integrator;
        REQUIRE(integrator.setIntegrationStep(10us) && integrator.setDynamicalSystem(system) && integrator.integrate(0s, 10us));
        REQUIRE(kinDyn->setRobotStateBatch(system->basePositions().data(), system->baseQuaternions().data(), system->jointPositions().data()));
        for (std::size_t i = 0; i < N; i++) REQUIRE(kinDyn->getFrameVel(hand, sign > 0 ? vp[i] : vm[i], i));
    }
    for (std::size_t i = 0; i < N; i++)
        for (int r = 0; r < 6; r++) REQUIRE(std::abs((vp[i][r] - vm[i][r]) / (2 * d) - bias[i][r]) < 2e-5 * (1.0 + std::abs(bias[i][r])));
}

static void testIntegrator()
{
    constexpr auto dT = 100us;
    const std::size_t n = 20;
    std::mt19937 gen(7);
    std::uniform_real_distribution<double> u(-1, 1);
    auto system = std::make_shared<FloatingBaseSystemVelocityKinematics>();
    std::array<double, 6> twist;
    for (auto& t : twist) t = u(gen);
    VectorXd ds(n), s0(n);
    for (std::size_t j = 0; j < n; j++) ds[j] = u(gen), s0[j] = u(gen);
    const Vector3d p0{0.1, 0.2, 0.3};
    const SO3d R0 = SO3d::Exp({0.4, -0.3, 0.2});
    REQUIRE(system->setState({p0, R0, s0}));
    REQUIRE(system->setControlInput({twist, ds}));
    ForwardEuler<FloatingBaseSystemVelocityKinematics> integrator;
    REQUIRE(integrator.setIntegrationStep(dT) && integrator.setDynamicalSystem(system));
    constexpr double T = 0.5;
    REQUIRE(integrator.integrate(0s, 500ms));   // 4999 full steps + the last one, in two launches
    const auto& [p, R, s] = integrator.getSolution();
    for (int c = 0; c < 3; c++) REQUIRE(std::abs(p[c] - (p0[c] + T * twist[c])) < 1e-9);
    for (std::size_t j = 0; j < n; j++) REQUIRE(std::abs(s[j] - (s0[j] + T * ds[j])) < 1e-9);
    // closed form R = Exp(w T) R0 (tangent + group = left-plus)
    const SO3d Rexp = SO3d::Exp({twist[3] * T, twist[4] * T, twist[5] * T});
    const auto a = Rexp.rotation(), b = R0.rotation(), got = R.rotation();
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double e = 0;
            for (int k = 0; k < 3; k++) e += a[3 * i + k] * b[3 * k + j];
            REQUIRE(std::abs(got[3 * i + j] - e) < 1e-9);
        }
    log()->setLevel(Logger::Level::off);
    REQUIRE_FALSE(integrator.integrate(0s, 50us));   // shorter than dT: rejected, not underflowed
    log()->setLevel(Logger::Level::warn);
}


// src/ContinuousDynamicalSystem/tests/IntegratorLinearSystem.cpp replayed: a mass-spring-damper driven by a step, against its
// closed-form response, ForwardEuler at the reference's 1e-3 and RK4 at 1e-8 (the reference accepts 1e-3 for both).
static void testLinearSystemIntegrators()
{
    constexpr auto dT = 100us;
    constexpr double tolEuler = 1e-3, tolRK4 = 1e-8, simulationTime = 2.0;
    const std::vector<double> A{0, 1, -2, -2}, B{0, 2};
    auto closedForm = [](double t) {
        return std::array<double, 2>{1 - std::exp(-t) * (std::cos(t) + std::sin(t)), 2 * std::exp(-t) * std::sin(t)};
    };
    auto run = [&](auto& integrator, auto system, double tol) {
        REQUIRE(system->setSystemMatrices(A, 2, 2, B, 2, 1));
        REQUIRE(system->setState({VectorXd{0.0, 0.0}}));
        REQUIRE(system->setControlInput({VectorXd{1.0}}));
        REQUIRE(integrator.setIntegrationStep(dT));
        REQUIRE(integrator.setDynamicalSystem(system));
        const int steps = static_cast<int>(simulationTime / 1e-4);
        double worst = 0;
        for (int i = 0; i < steps; i++) {
            const auto& [x] = integrator.getSolution();
            const auto e = closedForm(1e-4 * i);
            worst = std::max({worst, std::abs(x[0] - e[0]), std::abs(x[1] - e[1])});
            REQUIRE(integrator.integrate(0s, dT));
        }
        REQUIRE(worst < tol);
    };
    ForwardEuler<LinearTimeInvariantSystem> euler;
    run(euler, std::make_shared<LinearTimeInvariantSystem>(), tolEuler);
    RK4<LinearTimeInvariantSystem> rk4;
    run(rk4, std::make_shared<LinearTimeInvariantSystem>(), tolRK4);
    // the reference's error rules
    log()->setLevel(Logger::Level::off);
    LinearTimeInvariantSystem bad;
    REQUIRE_FALSE(bad.setSystemMatrices(A, 2, 2, B, 1, 2));                           // A and B: different rows
    REQUIRE_FALSE(bad.setSystemMatrices({0, 1, 2, 3, 4, 5}, 2, 3, {0, 2}, 2, 1));     // A not square
    LinearTimeInvariantSystem::StateDerivative dx;
    REQUIRE_FALSE(bad.dynamics(0s, dx));                                              // matrices never set
    REQUIRE(bad.setSystemMatrices(A, 2, 2, B, 2, 1));
    REQUIRE(bad.setState({VectorXd{0.0, 0.0, 0.0}}));
    REQUIRE(bad.setControlInput({VectorXd{1.0}}));
    REQUIRE_FALSE(bad.dynamics(0s, dx));                                              // state of the wrong size
    REQUIRE(bad.setState({VectorXd{0.0, 0.0}}));
    REQUIRE(bad.setControlInput({VectorXd{1.0, 2.0}}));
    REQUIRE_FALSE(bad.dynamics(0s, dx));                                              // input of the wrong size
    RK4<LinearTimeInvariantSystem> twice;
    REQUIRE(twice.setDynamicalSystem(std::make_shared<LinearTimeInvariantSystem>()));
    REQUIRE_FALSE(twice.setDynamicalSystem(std::make_shared<LinearTimeInvariantSystem>()));
    REQUIRE_FALSE(twice.setIntegrationStep(0ns));
    log()->setLevel(Logger::Level::warn);
}

// FloatingBaseSystemAccelerationKinematics under ForwardEuler, on the device: constant acceleration has the closed forms
// v = v0 + a T and (explicit Euler, derivative at the old state) p_K = p0 + K dt v0 + dt^2 a K (K - 1) / 2; the rotation is checked
// at zero angular acceleration, where R = Exp(w T) R0.  Batch of 3 robots with different inputs.
static void testAccelerationIntegrator()
{
    constexpr auto dT = 1ms;
    const std::size_t n = 7, N = 3;
    const int K = 250;
    std::mt19937 gen(11);
    std::uniform_real_distribution<double> u(-1, 1);
    auto system = std::make_shared<FloatingBaseSystemAccelerationKinematics>();
    system->setBatchSize(N);
    std::vector<FloatingBaseSystemAccelerationKinematics::State> x0(N);
    std::vector<FloatingBaseSystemAccelerationKinematics::Input> in(N);
    for (std::size_t i = 0; i < N; i++) {
        auto& [p, R, s, dp, w, ds] = x0[i];
        for (int c = 0; c < 3; c++) p[c] = u(gen), dp[c] = u(gen), w.w[c] = u(gen);
        R = SO3d::Exp({u(gen), u(gen), u(gen)});
        s.resize(n), ds.resize(n);
        for (std::size_t j = 0; j < n; j++) s[j] = u(gen), ds[j] = u(gen);
        auto& [dtw, dds] = in[i];
        for (int c = 0; c < 6; c++) dtw[c] = c < 3 ? u(gen) : 0.0;
        dds.resize(n);
        for (auto& v : dds) v = u(gen);
        REQUIRE(system->setState(x0[i], static_cast<std::ptrdiff_t>(i)));
        REQUIRE(system->setControlInput(in[i], static_cast<std::ptrdiff_t>(i)));
    }
    FloatingBaseSystemAccelerationKinematics::StateDerivative d;
    REQUIRE(system->dynamics(0s, d, 1));
    for (int c = 0; c < 3; c++) REQUIRE(std::get<0>(d)[c] == std::get<3>(x0[1])[c] && std::get<3>(d)[c] == std::get<0>(in[1])[c]);
    ForwardEuler<FloatingBaseSystemAccelerationKinematics> integrator;
    REQUIRE(integrator.setIntegrationStep(dT) && integrator.setDynamicalSystem(system));
    REQUIRE(integrator.integrate(0s, 250ms));
    const double dt = 1e-3, T = K * dt, quad = dt * dt * K * (K - 1) / 2.0;
    for (std::size_t i = 0; i < N; i++) {
        const auto& [p, R, s, dp, w, ds] = integrator.getSolution(i);
        const auto& [p0, R0, s0, dp0, w0, ds0] = x0[i];
        const auto& [dtw, dds] = in[i];
        for (int c = 0; c < 3; c++) {
            REQUIRE(std::abs(dp[c] - (dp0[c] + T * dtw[c])) < 1e-12);
            REQUIRE(std::abs(p[c] - (p0[c] + T * dp0[c] + quad * dtw[c])) < 1e-12);
            REQUIRE(std::abs(w.w[c] - w0.w[c]) < 1e-15);
        }
        for (std::size_t j = 0; j < n; j++) {
            REQUIRE(std::abs(ds[j] - (ds0[j] + T * dds[j])) < 1e-12);
            REQUIRE(std::abs(s[j] - (s0[j] + T * ds0[j] + quad * dds[j])) < 1e-12);
        }
        const SO3d Rexp = SO3d::Exp({w0.w[0] * T, w0.w[1] * T, w0.w[2] * T});
        const auto a = Rexp.rotation(), b = R0.rotation(), got = R.rotation();
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) {
                double e = 0;
                for (int k = 0; k < 3; k++) e += a[3 * r + k] * b[3 * k + c];
                REQUIRE(std::abs(got[3 * r + c] - e) < 1e-10);
            }
    }
}

// The batch path (N >= 64): packed page-locked records as the native storage of state and set-points, one blfik_tick_packed per
// shard.  Its results must not depend on how the batch is cut: one shard, three shards (all on device 0 here: the sharding
// logic is the same whichever device a block lands on), compact records; and each robot must get what it gets when it is solved
// in a small batch through the per-array calls.  Also: the regularisation set-point given to all robots at once travels as
// one batch-shared record, a per-robot call afterwards brings the records back, and advanceAndIntegrate without the host
// copy-back follows the same trajectory as with it.
static void testBatchPath(const char* modelPath)
{
    constexpr std::size_t N = 150;
    constexpr auto dT = 10ms;
    std::mt19937 gen(7);
    std::normal_distribution<double> noise(0.0, 0.04);
    struct Case { std::vector<int> devices; bool compact; std::size_t n; };
    const std::vector<Case> cases{{{}, false, N}, {{0, 0, 0}, false, N}, {{0, 0}, true, N}, {{}, false, 40}};
    std::vector<std::vector<double>> results;
    std::vector<VectorXd> sDes(N), s0(N);
    std::vector<double> tilt(N);
    {
        auto kd = loadKinDyn(modelPath, 1);
        for (std::size_t i = 0; i < N; i++) {
            sDes[i] = nominalJoints(*kd, gen, 0.15);
            s0[i] = sDes[i];
            for (auto& v : s0[i]) v += noise(gen);
            tilt[i] = 0.05 + noise(gen);
        }
    }
    for (const Case& c : cases) {
        auto kinDyn = loadKinDyn(modelPath, c.n);
        auto desired = loadKinDyn(modelPath, c.n);
        const std::size_t n = kinDyn->getNrOfDegreesOfFreedom();
        for (std::size_t i = 0; i < c.n; i++) REQUIRE(desired->setRobotState({0.0, 0.0, 0.75}, SO3d::Exp({0.0, 0.05, 0.1}), sDes[i], i));
        auto problem = QPInverseKinematics::build(ergoCubHandler(n, true, c.devices, c.compact), kinDyn);
        REQUIRE(problem.isValid());
        auto* ik = static_cast<QPInverseKinematics*>(problem.ik.get());
        auto comTask = std::dynamic_pointer_cast<CoMTask>(ik->getTask("COM_TASK").lock());
        auto rFoot = std::dynamic_pointer_cast<SE3Task>(ik->getTask("RIGHT_FOOT_TASK").lock());
        auto lFoot = std::dynamic_pointer_cast<SE3Task>(ik->getTask("LEFT_FOOT_TASK").lock());
        auto torso = std::dynamic_pointer_cast<SO3Task>(ik->getTask("TORSO_TASK").lock());
        auto reg = std::dynamic_pointer_cast<JointTrackingTask>(ik->getTask("JOINT_REGULARIZATION_TASK").lock());
        for (std::size_t i = 0; i < c.n; i++) {
            std::array<double, 9> R;
            Vector3d com;
            SE3d r, l, chest;
            REQUIRE(desired->getCenterOfMassPosition(com, i));
            REQUIRE(desired->getWorldTransform(kinDyn->getFrameIndex("r_sole"), r, R, i));
            REQUIRE(desired->getWorldTransform(kinDyn->getFrameIndex("l_sole"), l, R, i));
            REQUIRE(desired->getWorldTransform(kinDyn->getFrameIndex("chest"), chest, R, i));
            REQUIRE(comTask->setSetPoint(com, {0, 0, 0}, i));
            REQUIRE(rFoot->setSetPoint(r, SE3d::Tangent::Zero(), i) && lFoot->setSetPoint(l, SE3d::Tangent::Zero(), i));
            REQUIRE(torso->setSetPoint(chest.R, SO3Tangentd::Zero(), i));
            REQUIRE(kinDyn->setRobotState({0.0, 0.0, 0.75}, SO3d::Exp({0.0, tilt[i], 0.1}), s0[i], i));
        }
        REQUIRE(reg->setSetPoint(sDes[0]));                      // every robot at once: one batch-shared record
        REQUIRE(reg->setPointIsUniform());
        REQUIRE(ik->advance() && ik->isOutputValid());
        std::vector<double> first(ik->batchOutputData(), ik->batchOutputData() + c.n * (n + 6));
        for (std::size_t i = 0; i < c.n; i++) REQUIRE(reg->setSetPoint(sDes[i], i));   // and now one per robot: back into the records
        REQUIRE_FALSE(reg->setPointIsUniform());
        REQUIRE(ik->advance());
        for (std::size_t i = 0; i < c.n; i++) REQUIRE(ik->isOutputValid(i));
        REQUIRE(ik->getBatchStatus().size() == c.n);
        std::vector<double> out(ik->getBatchOutput());
        double moved = 0;
        for (std::size_t k = 0; k < out.size(); k++) moved = std::max(moved, std::abs(out[k] - first[k]));
        REQUIRE(moved > 1e-4);                                   // the per-robot postural targets differ from robot 0's
        // closed loop on the device, with and without reading the state back every period
        for (int it = 0; it < 5; it++) REQUIRE(ik->advanceAndIntegrate(dT, it % 2 == 0 ? false : true));
        REQUIRE(ik->advanceAndIntegrate(dT, false) && ik->syncState());
        const auto& sEnd = kinDyn->jointPositions();
        out.insert(out.end(), sEnd.begin(), sEnd.end());
        results.push_back(std::move(out));
    }
    const std::size_t nv = 29, n = 23;
    REQUIRE(results[0].size() == N * (nv + n));
    for (std::size_t k = 0; k < results[0].size(); k++) REQUIRE(results[1][k] == results[0][k]);            // cut differently: bit-identical
    double dCompact = 0, dSmall = 0;
    for (std::size_t k = 0; k < results[0].size(); k++) dCompact = std::max(dCompact, std::abs(results[2][k] - results[0][k]));
    REQUIRE(dCompact == 0.0);                                                                               // the feed-forward terms were zero
    for (std::size_t i = 0; i < 40; i++) {
        for (std::size_t c = 0; c < nv; c++) dSmall = std::max(dSmall, std::abs(results[3][i * nv + c] - results[0][i * nv + c]));
        for (std::size_t c = 0; c < n; c++) dSmall = std::max(dSmall, std::abs(results[3][40 * nv + i * n + c] - results[0][N * nv + i * n + c]));
    }
    REQUIRE(dSmall < 1e-7);   // the small batch takes the per-array calls (and, at 40 robots, the same kernels): same optimum, same rollout
}

int main(int argc, char** argv)
{
    if (argc < 3) {
        std::fprintf(stderr, "usage: %s cpu|gpu <ergocub23 model file>\n", argv[0]);
        return 2;
    }
    const bool gpu = std::strcmp(argv[1], "gpu") == 0;
    testParametersAndVariables();
    testTaskInitializationAndAddTaskRules(argv[2]);
    testWeightProviders();
    testSplines();
    testSO3Planner();
    testSwingFootPlanner();
    testLinearSystemIntegrators();
    if (gpu) {
        testTaskMatrices(argv[2]);
        testClosedLoop(argv[2], 1, false);
        testClosedLoop(argv[2], 1, true);
        testClosedLoop(argv[2], 48, true);
        testClosedLoop(argv[2], 200, true);   // the packed-record batch path
        testBatchPath(argv[2]);
        testFixedBase(argv[2], 1);
        testFixedBase(argv[2], 5);
        testSetPointSpline(argv[2], 3);
        testJointSetPointSpline<BipedalLocomotion::Math::QuinticSpline<VectorXd>>(argv[2], 3);
        testJointSetPointSpline<BipedalLocomotion::Math::CubicSpline<VectorXd>>(argv[2], 70);   // the packed-record path
        testIntegrator();
        testAccelerationIntegrator();
        testKinDynDynamics(argv[2]);
    }
    std::printf("%s: %d checks, %d failures\n", gpu ? "gpu" : "cpu", g_checks, g_failures);
    return g_failures == 0 ? 0 : 1;
}
