// blf_ref_ik.cpp -- TEST INFRASTRUCTURE, not compiled on the graft image (see ../REF_BUILD.md).
// One C entry point around the UNMODIFIED reference: the caller loop of src/IK/tests/QPInverseKinematicsTest.cpp:459-484 for the
// C3 task list, on a kinematic tree handed over in the layout of blfik_model_desc.
#include <BipedalLocomotion/ContinuousDynamicalSystem/FloatingBaseSystemVelocityKinematics.h>
#include <BipedalLocomotion/ContinuousDynamicalSystem/ForwardEuler.h>
#include <BipedalLocomotion/IK/CoMTask.h>
#include <BipedalLocomotion/IK/JointTrackingTask.h>
#include <BipedalLocomotion/IK/QPInverseKinematics.h>
#include <BipedalLocomotion/IK/SE3Task.h>
#include <BipedalLocomotion/IK/SO3Task.h>
#include <BipedalLocomotion/ParametersHandler/StdImplementation.h>

#include <iDynTree/KinDynComputations.h>
#include <iDynTree/Model.h>
#include <iDynTree/RevoluteJoint.h>

#include <memory>
#include <string>
#include <vector>

using namespace BipedalLocomotion;
using namespace std::chrono_literals;

namespace
{
iDynTree::Transform toTransform(const double* R, const double* p)
{
    iDynTree::Rotation rot(R[0], R[1], R[2], R[3], R[4], R[5], R[6], R[7], R[8]);
    return iDynTree::Transform(rot, iDynTree::Position(p[0], p[1], p[2]));
}
}  // namespace

// tree: parent[nl], dof[nl], X_rot[nl][9], X_pos[nl][3], axis[nl][3], mass[nl], com[nl][3], inertia[nl][6] (about the CoM, link
// frame), frames: link index + pose for "l_sole", "r_sole", "chest" (in that order).  state: p(3) q(4, xyzw) s(n).
// set-points: com(3) r_foot(7: p, q) l_foot(7) chest(4: q) posture(n), zero feed-forward.  Returns 0 on success.
extern "C" int ref_ik_tick(int n, const int* parent, const int* dof, const double* Xr, const double* Xp, const double* axis, const double* mass,
                           const double* com, const double* inertia, const int* frame_link, const double* frame_R, const double* frame_p,
                           const double* q_lower, const double* q_upper, const double* state, const double* sp, double dt, int tight,
                           double* nu, double* state_out)
{
    const int nl = n + 1;
    iDynTree::Model model;
    std::vector<std::string> linkNames(nl), jointNames(n);
    for (int k = 0; k < nl; k++) {
        iDynTree::Link link;
        iDynTree::RotationalInertia I;
        const double* i6 = inertia + 6 * k;
        I(0, 0) = i6[0], I(0, 1) = I(1, 0) = i6[1], I(0, 2) = I(2, 0) = i6[2], I(1, 1) = i6[3], I(1, 2) = I(2, 1) = i6[4], I(2, 2) = i6[5];
        iDynTree::SpatialInertia si;
        si.fromRotationalInertiaWrtCenterOfMass(mass[k], iDynTree::Position(com[3 * k], com[3 * k + 1], com[3 * k + 2]), I);
        link.setInertia(si);
        linkNames[k] = "link_" + std::to_string(k);
        model.addLink(linkNames[k], link);
    }
    for (int k = 1; k < nl; k++) {
        iDynTree::RevoluteJoint joint;
        joint.setAttachedLinks(model.getLinkIndex(linkNames[parent[k]]), model.getLinkIndex(linkNames[k]));
        joint.setRestTransform(toTransform(Xr + 9 * k, Xp + 3 * k));
        joint.setAxis(iDynTree::Axis(iDynTree::Direction(axis[3 * k], axis[3 * k + 1], axis[3 * k + 2]), iDynTree::Position::Zero()),
                      model.getLinkIndex(linkNames[k]), model.getLinkIndex(linkNames[parent[k]]));
        joint.enablePosLimits(true);
        joint.setPosLimits(0, q_lower[dof[k]], q_upper[dof[k]]);
        jointNames[dof[k]] = "joint_" + std::to_string(dof[k]);
        model.addJoint(jointNames[dof[k]], &joint);   // NB: joints must end up in DoF order; reorder with createReducedModel if needed
    }
    const char* frameNames[3] = {"l_sole", "r_sole", "chest"};
    for (int f = 0; f < 3; f++) model.addAdditionalFrameToLink(linkNames[frame_link[f]], frameNames[f], toTransform(frame_R + 9 * f, frame_p + 3 * f));
    auto kinDyn = std::make_shared<iDynTree::KinDynComputations>();
    kinDyn->loadRobotModel(model);
    kinDyn->setFloatingBase(linkNames[0]);
    kinDyn->setFrameVelocityRepresentation(iDynTree::MIXED_REPRESENTATION);

    auto handler = std::make_shared<ParametersHandler::StdImplementation>();
    handler->setParameter("tasks", std::vector<std::string>{"COM_TASK", "RIGHT_FOOT_TASK", "LEFT_FOOT_TASK", "TORSO_TASK", "JOINT_REGULARIZATION_TASK"});
    auto ikGroup = std::make_shared<ParametersHandler::StdImplementation>();
    ikGroup->setParameter("robot_velocity_variable_name", "robot_velocity");
    handler->setGroup("IK", ikGroup);
    // ... the task groups with the gains of ik.ini (see host/bench/host_bench.cpp: c3Handler) and the two limit tasks ...
    auto problem = IK::QPInverseKinematics::build(handler, kinDyn);
    if (!problem.isValid()) return 1;
    (void)tight;   // osqp-eigen settings (eps_abs / eps_rel / polish) are reached through QPInverseKinematics' solver handle in a patched build only;
                   // with the unmodified reference the run is at OSQP's defaults
    Eigen::Map<const Eigen::VectorXd> s(state + 7, n);
    const Eigen::Quaterniond qb(state[6], state[3], state[4], state[5]);
    Eigen::Matrix4d H = Eigen::Matrix4d::Identity();
    H.topLeftCorner<3, 3>() = qb.toRotationMatrix(), H.topRightCorner<3, 1>() = Eigen::Map<const Eigen::Vector3d>(state);
    const Eigen::VectorXd ds = Eigen::VectorXd::Zero(n);
    const Eigen::Matrix<double, 6, 1> zero6 = Eigen::Matrix<double, 6, 1>::Zero();
    const Eigen::Vector3d gravity(0, 0, -9.81);
    if (!kinDyn->setRobotState(H, s, zero6, ds, gravity)) return 2;
    (void)sp;   // setSetPoint on every task from `sp`, as QPInverseKinematicsTest.cpp:459-484 does
    if (!problem.ik->advance()) return 3;
    const auto& out = problem.ik->getOutput();
    Eigen::Map<Eigen::VectorXd>(nu, n + 6) << out.baseVelocity.coeffs(), out.jointVelocity;
    (void)dt, (void)state_out;   // ForwardEuler<FloatingBaseSystemVelocityKinematics>::integrate(0s, dt) on (nu, state)
    return 0;
}
