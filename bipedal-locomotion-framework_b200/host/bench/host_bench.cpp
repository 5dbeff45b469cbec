/**
 * host_bench.cpp -- the C3 control tick timed through the reference-named C++ call, QPInverseKinematics::advance().
 *
 *   host_bench <model file> <inputs file> <steps> <warmup> <devices, e.g. 0 or 0,1,2,3> <compact 0|1> <warm start 0|1> [<nu output file>]
 *
 * The inputs file is written by bench.py (--via host): int64 N, int64 n, then p[N][3] q[N][4] s[N][n] and the set-points of
 * the CoM (6), right foot (13), left foot (13), chest (7) and postural (2n) tasks, all float64, robot-major.  The caller's part
 * of a control period -- kinDyn->setRobotState and task->setSetPoint per robot, the calls of
 * QPInverseKinematicsTest.cpp:459-484 -- is done once before the timed loop and timed on its own (`fill_ms`): those calls write
 * straight into the solver's page-locked records, so advance() has nothing to gather.  The timed loop calls advance() `steps`
 * times; every call sends the records to the device(s), runs the tick and has nu + status back on the host when it returns.
 * One JSON line on stdout.
 */
#include <BipedalLocomotion/IK.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>

using namespace BipedalLocomotion;
using namespace BipedalLocomotion::IK;
using namespace BipedalLocomotion::ParametersHandler;

static std::shared_ptr<IParametersHandler> c3Handler(std::size_t n, const std::vector<int>& devices, bool compact)
{
    // gains and weights of utilities/balancing-position-control/config/robots/ergoCubGazeboV1/blf_balancing_position_control/ik.ini
    auto h = std::make_shared<StdImplementation>();
    h->setParameter("tasks", std::vector<std::string>{"COM_TASK", "RIGHT_FOOT_TASK", "LEFT_FOOT_TASK", "JOINT_LIMITS_TASK", "JOINT_VELOCITY_LIMITS_TASK",
                                                       "TORSO_TASK", "JOINT_REGULARIZATION_TASK"});
    auto ik = std::make_shared<StdImplementation>();
    ik->setParameter("robot_velocity_variable_name", std::string("robot_velocity"));
    ik->setParameter("cuda_devices", devices);
    ik->setParameter("compact_records", compact);
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
    auto lim = std::make_shared<StdImplementation>();
    lim->setParameter("type", "JointLimitsTask"), lim->setParameter("klim", std::vector<double>(n, 0.5));
    lim->setParameter("sampling_time", 0.01), lim->setParameter("use_model_limits", true), lim->setParameter("priority", 0);
    h->setGroup("JOINT_LIMITS_TASK", lim);
    auto vel = std::make_shared<StdImplementation>();
    vel->setParameter("type", "JointVelocityLimitsTask"), vel->setParameter("upper_limits", std::vector<double>(n, 2.0));
    vel->setParameter("lower_limits", std::vector<double>(n, -2.0)), vel->setParameter("priority", 0);
    h->setGroup("JOINT_VELOCITY_LIMITS_TASK", vel);
    auto torso = std::make_shared<StdImplementation>();
    torso->setParameter("type", "SO3Task"), torso->setParameter("frame_name", "chest"), torso->setParameter("kp_angular", 5.0);
    torso->setParameter("priority", 1), torso->setParameter("weight", std::vector<double>{5.0, 5.0, 5.0});
    h->setGroup("TORSO_TASK", torso);
    std::vector<double> w(n, 1.0);
    for (std::size_t j = 3; j < 11 && j < n; j++) w[j] = 2.0;   // ik.ini:62-66
    auto reg = std::make_shared<StdImplementation>();
    reg->setParameter("type", "JointTrackingTask"), reg->setParameter("kp", std::vector<double>(n, 5.0));
    reg->setParameter("priority", 1), reg->setParameter("weight", w);
    h->setGroup("JOINT_REGULARIZATION_TASK", reg);
    return h;
}

static double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char** argv)
{
    if (argc < 8) {
        std::fprintf(stderr, "usage: %s <model> <inputs> <steps> <warmup> <devices> <compact 0|1> <warm start 0|1> [<nu out>]\n", argv[0]);
        return 2;
    }
    const int steps = std::atoi(argv[3]), warmup = std::atoi(argv[4]);
    std::vector<int> devices;
    {
        std::stringstream ss(argv[5]);
        std::string tok;
        while (std::getline(ss, tok, ',')) devices.push_back(std::atoi(tok.c_str()));
    }
    const bool compact = std::atoi(argv[6]) != 0, warm = std::atoi(argv[7]) != 0;
    std::ifstream in(argv[2], std::ios::binary);
    int64_t N64 = 0, n64 = 0;
    in.read(reinterpret_cast<char*>(&N64), 8), in.read(reinterpret_cast<char*>(&n64), 8);
    const std::size_t N = static_cast<std::size_t>(N64), n = static_cast<std::size_t>(n64);
    auto readArr = [&](std::size_t cols) {
        std::vector<double> v(N * cols);
        in.read(reinterpret_cast<char*>(v.data()), static_cast<std::streamsize>(v.size() * sizeof(double)));
        return v;
    };
    const auto p = readArr(3), q = readArr(4), s = readArr(n), spCom = readArr(6), spR = readArr(13), spL = readArr(13), spChest = readArr(7), spReg = readArr(2 * n);
    if (!in) {
        std::fprintf(stderr, "host_bench: short inputs file\n");
        return 2;
    }
    log()->setLevel(Logger::Level::error);
    auto kinDyn = std::make_shared<KinDynBatch>();
    if (!kinDyn->loadRobotModelFromFile(argv[1]) || !kinDyn->setBatchSize(N, devices.empty() ? 0 : devices[0])) return 3;
    auto problem = QPInverseKinematics::build(c3Handler(n, devices, compact), kinDyn);
    if (!problem.isValid()) return 4;
    auto* ik = static_cast<QPInverseKinematics*>(problem.ik.get());
    ik->setWarmStart(warm);   // false: every advance() solves from scratch, like the tick bench.py times
    auto comTask = std::dynamic_pointer_cast<CoMTask>(ik->getTask("COM_TASK").lock());
    auto rFoot = std::dynamic_pointer_cast<SE3Task>(ik->getTask("RIGHT_FOOT_TASK").lock());
    auto lFoot = std::dynamic_pointer_cast<SE3Task>(ik->getTask("LEFT_FOOT_TASK").lock());
    auto torso = std::dynamic_pointer_cast<SO3Task>(ik->getTask("TORSO_TASK").lock());
    auto reg = std::dynamic_pointer_cast<JointTrackingTask>(ik->getTask("JOINT_REGULARIZATION_TASK").lock());
    if (!comTask || !rFoot || !lFoot || !torso || !reg) return 5;

    // the caller's part of a period, through the reference's calls
    auto fill = [&]() {
        bool ok = true;
        VectorXd sj(n);
        bool uniformReg = true;
        for (std::size_t i = 1; i < N && uniformReg; i++)
            for (std::size_t c = 0; c < 2 * n; c++)
                if (spReg[i * 2 * n + c] != spReg[c]) {
                    uniformReg = false;
                    break;
                }
        for (std::size_t i = 0; i < N; i++) {
            sj.assign(s.begin() + i * n, s.begin() + (i + 1) * n);
            ok &= kinDyn->setRobotState({p[3 * i], p[3 * i + 1], p[3 * i + 2]}, SO3d(Quaterniond{q[4 * i], q[4 * i + 1], q[4 * i + 2], q[4 * i + 3]}), sj, i);
            const double* c = &spCom[6 * i];
            ok &= comTask->setSetPoint({c[0], c[1], c[2]}, {c[3], c[4], c[5]}, i);
            for (int side = 0; side < 2; side++) {
                const double* f = side ? &spL[13 * i] : &spR[13 * i];
                SE3d::Tangent tw;
                for (int k = 0; k < 3; k++) tw.linear[k] = f[7 + k], tw.angular[k] = f[10 + k];
                ok &= (side ? lFoot : rFoot)->setSetPoint(SE3d({f[0], f[1], f[2]}, SO3d(Quaterniond{f[3], f[4], f[5], f[6]})), tw, i);
            }
            const double* t = &spChest[7 * i];
            SO3Tangentd tw;
            for (int k = 0; k < 3; k++) tw.w[k] = t[4 + k];
            ok &= torso->setSetPoint(SO3d(Quaterniond{t[0], t[1], t[2], t[3]}), tw, i);
            if (!uniformReg) {
                VectorXd sd(spReg.begin() + i * 2 * n, spReg.begin() + i * 2 * n + n), vd(spReg.begin() + i * 2 * n + n, spReg.begin() + (i + 1) * 2 * n);
                ok &= reg->setSetPoint(sd, vd, i);
            }
        }
        if (uniformReg) {   // the postural target is one vector for the whole batch: given once, for every robot
            VectorXd sd(spReg.begin(), spReg.begin() + n), vd(spReg.begin() + n, spReg.begin() + 2 * n);
            ok &= reg->setSetPoint(sd, vd);
        }
        return ok;
    };
    if (!fill()) return 6;
    for (int k = 0; k < warmup; k++)
        if (!ik->advance()) return 7;
    const double f0 = now_ms();
    if (!fill()) return 6;
    const double fill_ms = now_ms() - f0;
    const double t0 = now_ms();
    for (int k = 0; k < steps; k++)
        if (!ik->advance()) return 7;
    const double ms = now_ms() - t0;
    int64_t unsolved = 0;
    const int32_t* st = ik->batchStatusData();
    for (std::size_t i = 0; i < N; i++) unsolved += st[i] != 0;
    if (argc > 8) {
        std::ofstream out(argv[8], std::ios::binary);
        out.write(reinterpret_cast<const char*>(ik->batchOutputData()), static_cast<std::streamsize>(N * (n + 6) * sizeof(double)));
        out.write(reinterpret_cast<const char*>(st), static_cast<std::streamsize>(N * sizeof(int32_t)));
    }
    std::printf("{\"call\": \"QPInverseKinematics::advance()\", \"instances\": %zu, \"devices\": %zu, \"steps\": %d, \"ms_per_step\": %.6f, "
                "\"ticks_per_s\": %.1f, \"fill_ms\": %.3f, \"unsolved\": %lld, \"compact_records\": %s, \"warm_start\": %s}\n",
                N, devices.size(), steps, ms / steps, N * steps / (ms * 1e-3), fill_ms, static_cast<long long>(unsolved), compact ? "true" : "false", warm ? "true" : "false");
    return 0;
}
