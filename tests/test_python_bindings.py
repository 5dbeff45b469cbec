"""The Python surface (blf_b200.bindings), exercised the way the reference's own binding tests exercise theirs
(bindings/python/IK/tests/test_QP_inverse_kinematics.py, bindings/python/ContinuousDynamicalSystem/tests/
test_floating_base_system_kinematics.py, bindings/python/System/tests, bindings/python/ParametersHandler/tests).
The CPU part covers the host logic (handlers, key / size / priority rules); everything that solves runs on the GPU."""
import logging
from datetime import timedelta

import numpy as np
import pytest

import blf_b200  # noqa: F401
import blf_b200.bindings as blf
from blf_b200 import models as M
from blf_b200 import workloads as W
from blf_b200.bindings import manif

logging.getLogger("blf").setLevel(logging.CRITICAL)   # the error paths below are expected


def get_kindyn(model=None, batch_size=1):
    model = model or M.icub26()
    return model.joint_names, blf.KinDynComputations(model, batch_size)


# ------------------------------------------------------------------------------------------------ CPU: host logic
def test_parameters_handler():
    h = blf.parameters_handler.StdParametersHandler()
    h.set_parameter_bool(name="b", value=True)
    h.set_parameter_int(name="i", value=3)
    h.set_parameter_float(name="f", value=1.5)
    h.set_parameter_string(name="s", value="blf")
    h.set_parameter_vector_float(name="vf", value=[1.0, 2.0])
    h.set_parameter_vector_bool(name="vb", value=[True, False, True])
    h.set_parameter_datetime(name="dt", value=timedelta(milliseconds=10))
    assert h.get_parameter_bool("b") is True and h.get_parameter_int("i") == 3 and h.get_parameter_float("f") == 1.5
    assert h.get_parameter_float("i") == 3.0                 # an int converts to a float, as in the C++ handler
    assert h.get_parameter_string("s") == "blf" and list(h.get_parameter_vector_float("vf")) == [1.0, 2.0]
    assert h.get_parameter_datetime("dt") == timedelta(milliseconds=10)
    with pytest.raises(ValueError):
        h.get_parameter_int("b")                             # a bool is not an int
    with pytest.raises(ValueError):
        h.get_parameter_string("missing")
    g = blf.parameters_handler.StdParametersHandler()
    g.set_parameter_int("x", 1)
    assert h.set_group("G", g)
    g.set_parameter_int("x", 2)                              # the group was cloned (StdImplementation::setGroup)
    assert h.get_group("G").get_parameter_int("x") == 1
    with pytest.raises(ValueError):
        h.get_group("nope")
    assert not h.is_empty()
    h.clear()
    assert h.is_empty()


def test_variables_handler():
    v = blf.system.VariablesHandler()
    assert v.add_variable("a", 10) and v.add_variable("robotVelocity", 32) and not v.add_variable("a", 3)
    var = v.get_variable("robotVelocity")
    assert var.is_valid() and var.offset == 10 and var.size == 32 and v.get_number_of_variables() == 42
    assert not v.get_variable("missing").is_valid()
    h = blf.parameters_handler.StdParametersHandler()
    h.set_parameter_vector_string("variables_name", ["x", "y"])
    h.set_parameter_vector_int("variables_size", [2, 5])
    v2 = blf.system.VariablesHandler()
    assert v2.initialize(h) and v2.get_variable("y").offset == 2
    h.set_parameter_vector_int("variables_size", [2])
    assert not blf.system.VariablesHandler().initialize(h)


def _handler(**kw):
    h = blf.parameters_handler.StdParametersHandler()
    h.set_parameter_string(name="robot_velocity_variable_name", value="robotVelocity")
    for k, v in kw.items():
        if isinstance(v, bool):
            h.set_parameter_bool(k, v)
        elif isinstance(v, float):
            h.set_parameter_float(k, v)
        elif isinstance(v, str):
            h.set_parameter_string(k, v)
        elif isinstance(v, timedelta):
            h.set_parameter_datetime(k, v)
        elif len(v) and isinstance(v[0], bool):
            h.set_parameter_vector_bool(k, v)
        else:
            h.set_parameter_vector_float(k, v)
    return h


def test_task_initialisation_rules():
    """the per-task tests of test_QP_inverse_kinematics.py:82-283 plus the failure branches of the C++ initialize()s"""
    joints, kindyn = get_kindyn()
    n = kindyn.getNrOfDegreesOfFreedom()
    var = blf.system.VariablesHandler()
    assert var.add_variable("robotVelocity", n + 6)
    bad_var = blf.system.VariablesHandler()
    assert bad_var.add_variable("robotVelocity", n + 5)

    com = blf.ik.CoMTask()
    assert not com.initialize(param_handler=_handler(kp_linear=10.0))          # set_kin_dyn first
    assert com.set_kin_dyn(kindyn)
    assert not com.initialize(param_handler=_handler())                        # kp_linear missing
    assert com.initialize(param_handler=_handler(kp_linear=10.0))
    assert com.set_variables_handler(variables_handler=var) and not com.set_variables_handler(variables_handler=bad_var)
    assert com.set_set_point(position=np.array([1., 1., 1.5]), velocity=np.array([0., 0.5, 0.5]))
    assert com.size() == 3 and com.type() == blf.ik.IKLinearTask.Type.equality

    se3 = blf.ik.SE3Task()
    assert se3.set_kin_dyn(kindyn)
    assert not se3.initialize(param_handler=_handler(frame_name="not_a_frame", kp_linear=10.0, kp_angular=10.0))
    assert not se3.initialize(param_handler=_handler(frame_name="r_sole", kp_linear=10.0))
    assert se3.initialize(param_handler=_handler(frame_name="r_sole", kp_linear=10.0, kp_angular=[1.0, 2.0, 3.0], mask=[True, False, True]))
    assert se3.size() == 5
    assert se3.set_set_point(I_H_F=manif.SE3(position=[1.0, -2.0, 3.3], quaternion=[0, 0, -1, 0]), mixed_velocity=manif.SE3Tangent([0.0] * 6))

    exo = blf.ik.SE3Task()
    assert exo.set_kin_dyn(kindyn) and exo.initialize(param_handler=_handler(frame_name="r_sole", kp_linear=10.0, kp_angular=10.0,
                                                                             use_position_exogenous_feedback=True))
    assert exo._spec("t", 0, None).record_len(n) == 20 and se3._spec("t", 0, None).record_len(n) == 13
    assert exo.set_feedback(I_H_F=manif.SE3(position=[0.1, 0.2, 0.3], quaternion=[0, 0, 1, 0]))
    assert np.allclose(exo._fb[0], [0.1, 0.2, 0.3, 0, 0, 0, 1])       # the orientation key is off: that part is ignored
    assert se3.set_feedback(I_H_F=manif.SE3.Identity()) and se3._fb is None
    exo_com = blf.ik.CoMTask()
    assert exo_com.set_kin_dyn(kindyn) and exo_com.initialize(param_handler=_handler(kp_linear=10.0, use_exogenous_feedback=True))
    assert exo_com.set_feedback([0.0, 0.1, 0.5]) and np.allclose(exo_com._fb[0], [0.0, 0.1, 0.5]) and com.set_feedback([0, 0, 0]) and com._fb is None

    so3 = blf.ik.SO3Task()
    assert so3.set_kin_dyn(kindyn) and so3.initialize(param_handler=_handler(frame_name="chest", kp_angular=5.0))
    assert so3.set_set_point(I_R_F=manif.SO3(quaternion=[0, 0, -1, 0]), angular_velocity=manif.SO3Tangent([0.0] * 3))

    jt = blf.ik.JointTrackingTask()
    assert jt.set_kin_dyn(kindyn)
    assert not jt.initialize(param_handler=_handler(kp=[5.0] * (n - 1)))       # size-checked (JointTrackingTask.cpp:100-110)
    assert jt.initialize(param_handler=_handler(kp=[5.0] * n))
    assert jt.set_set_point(joint_position=np.zeros(n)) and jt.set_set_point(joint_position=np.zeros(n), joint_velocity=np.ones(n))
    assert not jt.set_set_point(joint_position=np.zeros(n - 1))

    jl = blf.ik.JointLimitsTask()
    assert jl.set_kin_dyn(kindyn)
    base = dict(klim=[0.5] * n, upper_limits=np.full(n, 0.01), lower_limits=np.full(n, -0.01), sampling_time=timedelta(milliseconds=10))
    assert not jl.initialize(param_handler=_handler(**base))                   # use_model_limits is mandatory
    assert not jl.initialize(param_handler=_handler(use_model_limits=False, **{**base, "klim": [1.5] * n}))
    assert not jl.initialize(param_handler=_handler(use_model_limits=False, **{**base, "sampling_time": timedelta(0)}))
    assert jl.initialize(param_handler=_handler(use_model_limits=False, **base))
    assert jl.type() == blf.ik.IKLinearTask.Type.inequality and jl.size() == 2 * n

    jv = blf.ik.JointVelocityLimitsTask()
    assert jv.set_kin_dyn(kindyn)
    assert not jv.initialize(param_handler=_handler(upper_limits=np.ones(n)))
    assert jv.initialize(param_handler=_handler(upper_limits=np.ones(n), lower_limits=-np.ones(n)))

    dist = blf.ik.DistanceTask()
    assert dist.set_kin_dyn(kindyn)
    assert not dist.initialize(param_handler=_handler(target_frame_name="chest", reference_frame_name="nope", kp=5.0))
    assert dist.initialize(param_handler=_handler(target_frame_name="chest", reference_frame_name="root_link", kp=5.0))
    assert dist.set_desired_distance(desired_distance=-0.3) and dist.set_set_point(desired_distance=0.2)
    assert dist._sp[0, 0] == 0.2

    grav = blf.ik.GravityTask()
    assert grav.set_kin_dyn(kindyn) and grav.initialize(param_handler=_handler(target_frame_name="chest", kp=5.0))
    assert grav.set_set_point(desired_gravity_direction=[0.0, 3.0, 4.0], feedforward_velocity=[0.1, 0.2, 0.3])
    assert np.allclose(grav._sp[0], [0.0, 0.6, 0.8, 0.1, 0.2, 0.3])

    am = blf.ik.AngularMomentumTask()
    assert am.set_kin_dyn(kindyn) and am.initialize(param_handler=_handler())
    assert am.set_variables_handler(variables_handler=var) and am.set_set_point([0., 0., 0.]) and am.size() == 3

    state = blf.ik.IntegrationBasedIKState()
    state.joint_velocity = [0.1] * n
    assert state.joint_velocity == pytest.approx([0.1] * n)


def test_add_task_rules_and_custom_task():
    """QPInverseKinematics::addTask (QPInverseKinematics.cpp:167-249) and the custom-task test (:12-56)"""
    joints, kindyn = get_kindyn()
    n = kindyn.getNrOfDegreesOfFreedom()
    qp_ik = blf.ik.QPInverseKinematics()
    jt = blf.ik.JointTrackingTask()
    assert jt.set_kin_dyn(kindyn) and jt.initialize(param_handler=_handler(kp=[5.0] * n))
    assert not qp_ik.add_task(task=jt, task_name="reg", priority=1, weight=[1.0] * n)     # initialize() first
    assert not qp_ik.initialize(handler=blf.parameters_handler.StdParametersHandler())
    assert qp_ik.initialize(handler=_handler())
    assert not qp_ik.add_task(task=jt, task_name="reg", priority=1)                        # weight is mandatory
    assert not qp_ik.add_task(task=jt, task_name="reg", priority=1, weight=[1.0] * (n - 1))
    assert not qp_ik.add_task(task=jt, task_name="reg", priority=2, weight=[1.0] * n)
    assert qp_ik.add_task(task=jt, task_name="reg", priority=1, weight=[1.0] * n)
    assert not qp_ik.add_task(task=jt, task_name="reg", priority=1, weight=[1.0] * n)      # duplicate name
    jv = blf.ik.JointVelocityLimitsTask()
    assert jv.set_kin_dyn(kindyn) and jv.initialize(param_handler=_handler(upper_limits=np.ones(n), lower_limits=-np.ones(n)))
    assert not qp_ik.add_task(task=jv, task_name="lim", priority=1, weight=[1.0] * 2 * n)  # inequality as a cost
    assert qp_ik.add_task(task=jv, task_name="lim", priority=0)
    assert qp_ik.get_task_names() == ["reg", "lim"] and qp_ik.get_task("reg") is jt and qp_ik.get_task("x") is None
    assert not qp_ik.set_task_weight("lim", weight=[1.0] * 2 * n) and not qp_ik.set_task_weight("reg", weight=[1.0])
    assert qp_ik.set_task_weight("reg", weight=[2.0] * n) and qp_ik.get_task_weight_provider("reg").get_output()[0] == 2.0
    assert not qp_ik.advance() and not qp_ik.is_output_valid()                            # finalize() first

    class CustomTask(blf.ik.IKLinearTask):
        def set_variables_handler(self, variables_handler):
            return variables_handler.get_variable("robotVelocity").is_valid()

        def size(self):
            return 10

        def is_valid(self):
            return True

    qp2 = blf.ik.QPInverseKinematics()
    assert qp2.initialize(handler=_handler())
    var = blf.system.VariablesHandler()
    assert var.add_variable("robotVelocity", 38)
    assert qp2.add_task(CustomTask(), "custom_task", 0) and qp2.finalize(var)
    assert not qp2.advance()      # a task with its own A and b is accepted by the interface but not solved by the batched kernels


def test_manif_stand_in():
    rng = np.random.default_rng(0)
    w = rng.normal(size=3)
    R = manif.SO3Tangent(w).exp()
    assert np.allclose(R.log().coeffs(), w) and np.allclose(R.rotation() @ R.rotation().T, np.eye(3))
    R0 = manif.SO3Tangent(rng.normal(size=3)).exp()
    left = manif.SO3Tangent(w) + R0                            # left-plus: Exp(w) R0
    assert np.allclose(left.rotation(), R.rotation() @ R0.rotation())
    assert np.allclose((left - R0).coeffs(), (R0.inverse() * R * R0).log().coeffs())


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_qp_inverse_kinematics_icub_vector():
    """test_QP_inverse_kinematics.py:293-425 line by line: iCub with 32 joints, CoM + SE3(r_sole) hard, JointTracking soft;
    advance() succeeds, and with new desiderata the solution changes."""
    model = M.icub32()
    joints_list, kindyn = get_kindyn(model)
    n = kindyn.getNrOfDegreesOfFreedom()
    assert n == 32
    qp_ik = blf.ik.QPInverseKinematics()
    assert qp_ik.initialize(handler=_handler())
    joint_values = [0.4147046680648114, 0.020109925966605414, -0.008920202688283316, -0.9417931898020863, -0.5243497613177115,
                    -0.021969353979888327, 0.41490863370740766, 0.019729971064546152, -0.00875663754461344, -0.942238731363784,
                    -0.5245880299799013, -0.021615580177534933, 0.10368807143683698, -1.7513061346176159e-06, -3.762644846952276e-05,
                    -0.1599473872698949, 0.4341604471687158, 0.18350557928473807, 0.5390720970077302, 6.0766272665334824e-05,
                    -0.0010549820549572227, -6.457409077219657e-05, 0.0002255313408076292, 2.3204341497780863e-05, 5.920527253092593e-07,
                    -0.1591984430641039, 0.4340643739146118, 0.18311274358002663, 0.5398797698508592, 5.367856260643534e-05,
                    -0.0010455524092736824, 6.414366731688693e-05]
    s = np.array(joint_values)
    world_gravity = [0.0, 0.0, blf.math.StandardAccelerationOfGravitation]
    assert kindyn.setRobotState(np.eye(4), s, [0.0] * 6, [0.0] * n, world_gravity)

    com_task = blf.ik.CoMTask()
    assert com_task.set_kin_dyn(kindyn) and com_task.initialize(param_handler=_handler(kp_linear=10.0))
    assert com_task.set_set_point(position=np.array([0., 0., 0.53]), velocity=np.array([0., 0., 0.]))
    assert qp_ik.add_task(task=com_task, task_name="Com_task", priority=0)

    se3_task = blf.ik.SE3Task()
    assert se3_task.set_kin_dyn(kindyn) and se3_task.initialize(param_handler=_handler(frame_name="r_sole", kp_linear=10.0, kp_angular=10.0))
    assert se3_task.set_set_point(I_H_F=manif.SE3(position=[-0.02, 0.1, 0.0], quaternion=[0, 0, 0, 1]), mixed_velocity=manif.SE3Tangent([0.0] * 6))
    assert qp_ik.add_task(task=se3_task, task_name="se3_task", priority=0)

    joint_tracking_task = blf.ik.JointTrackingTask()
    assert joint_tracking_task.set_kin_dyn(kindyn) and joint_tracking_task.initialize(param_handler=_handler(kp=[5.0] * n))
    rng = np.random.default_rng(0)
    assert joint_tracking_task.set_set_point(joint_position=s + rng.uniform(-5e-3, 5e-3, n))
    assert qp_ik.add_task(task=joint_tracking_task, task_name="joint_tracking_task", priority=1, weight=[1.0] * n)

    task_names = qp_ik.get_task_names()
    assert len(task_names) == 3 and 'Com_task' in task_names and 'joint_tracking_task' in task_names and 'se3_task' in task_names
    qp_ik_var_handler = blf.system.VariablesHandler()
    assert qp_ik_var_handler.add_variable("robotVelocity", 38) is True
    assert qp_ik.finalize(handler=qp_ik_var_handler)
    assert qp_ik.advance()
    state = qp_ik.get_output()
    assert qp_ik.is_output_valid()
    first_joint, first_base = state.joint_velocity.copy(), state.base_velocity.coeffs().copy()

    # the hard constraints hold for the returned velocity: J_com nu = v*, J_sole nu = twist*
    nu = qp_ik.get_raw_solution()
    Jc, c = kindyn.getCenterOfMassJacobian(), kindyn.getCenterOfMassPosition()
    assert np.abs(Jc @ nu - 10.0 * (np.array([0., 0., 0.53]) - c)).max() < 1e-8

    assert com_task.set_set_point(position=np.array([0., 0., 0.5]), velocity=np.array([0., 0., 0.]))
    assert se3_task.set_set_point(I_H_F=manif.SE3(position=[-0.02, 0.2, 0.0], quaternion=[0, 0, 0, 1]), mixed_velocity=manif.SE3Tangent([0.0] * 6))
    assert joint_tracking_task.set_set_point(joint_position=s + rng.uniform(-5e-3, 5e-3, n))
    assert qp_ik.advance()
    updated_state = qp_ik.get_output()
    assert qp_ik.is_output_valid()
    assert first_joint != pytest.approx(updated_state.joint_velocity)
    assert first_base != pytest.approx(updated_state.base_velocity.coeffs())


@pytest.mark.gpu
def test_floating_base_system_kinematics():
    """test_floating_base_system_kinematics.py:9-48 (closed form, tolerance 1e-3) -- with a non-identity R0 so that the
    left-plus convention is pinned, and a second run with a batch of states"""
    tolerance = 1e-3
    dt = 0.0001
    simulation_time = 0.05
    np.random.seed(42)
    twist = np.random.uniform(-2, 2, (6))
    joint_velocity = np.random.uniform(-2, 2, (23))
    rotation0 = manif.SO3Tangent([0.3, -0.2, 0.5]).exp()
    position0 = np.array([0.0, 0.0, 0.0])
    joint_position0 = np.zeros(23)
    system = blf.continuous_dynamical_system.FloatingBaseSystemVelocityKinematics()
    system.state = position0, rotation0, joint_position0
    assert system.set_control_input((twist, joint_velocity))
    assert not system.set_control_input((twist, joint_velocity[:-1]))
    integrator = blf.continuous_dynamical_system.FloatingBaseSystemVelocityKinematicsForwardEulerIntegrator()
    assert not integrator.integrate(timedelta(seconds=0), timedelta(seconds=dt))     # no system yet
    assert integrator.set_dynamical_system(system) and not integrator.set_dynamical_system(system)
    integrator.integration_step = timedelta(seconds=dt)
    assert not integrator.integrate(timedelta(seconds=dt), timedelta(seconds=0))
    for i in range(0, int(simulation_time / dt), 25):
        base_position, base_rotation, joint_position = integrator.get_solution()
        t = i * dt
        assert base_position == pytest.approx(position0 + t * twist[0:3], abs=tolerance)
        exact = manif.SO3Tangent(twist[3:6] * t) + rotation0
        assert base_rotation.rotation() == pytest.approx(exact.rotation(), abs=tolerance)
        assert joint_position == pytest.approx(joint_position0 + t * joint_velocity, abs=tolerance)
        assert integrator.integrate(timedelta(seconds=0), timedelta(seconds=25 * dt))


def test_linear_system():
    """bindings/python/ContinuousDynamicalSystem/tests/test_linear_system.py replayed (ForwardEuler, tolerance 1e-3), and the same
    run with the RK4 integrator at 1e-8; then the reference's error rules."""
    def close_form_solution(t):
        return np.array([1 - np.exp(-t) * (np.cos(t) + np.sin(t)), 2 * np.exp(-t) * np.sin(t)])

    A = np.array([[0, 1], [-2, -2]])
    B = np.array([0, 2])
    u = np.array([1])
    x0 = np.array([0, 0])
    dt = 0.0001
    simulation_time = 1
    cds = blf.continuous_dynamical_system
    for cls, tolerance in ((cds.LinearTimeInvariantSystemForwardEulerIntegrator, 1e-3), (cds.LinearTimeInvariantSystemRK4Integrator, 1e-8)):
        system = cds.LinearTimeInvariantSystem()
        assert system.set_system_matrices(A, B)
        assert system.set_state((x0,))
        assert system.set_control_input((u,))
        integrator = cls()
        assert integrator.set_dynamical_system(system)
        assert integrator.set_integration_step(timedelta(seconds=dt))
        for i in range(0, int(simulation_time / dt)):
            solution, = integrator.get_solution()
            assert solution == pytest.approx(close_form_solution(i * dt), abs=tolerance)
            assert integrator.integrate(timedelta(seconds=0), timedelta(seconds=dt))
    bad = cds.LinearTimeInvariantSystem()
    assert bad.dynamics() is None                                             # matrices never set
    assert not bad.set_system_matrices(A, np.zeros((3, 1)))                    # A and B: different rows
    assert not bad.set_system_matrices(np.zeros((2, 3)), np.zeros((2, 1)))     # A not square
    assert bad.set_system_matrices(A, B) and bad.set_state((np.zeros(3),)) and bad.set_control_input((u,))
    assert bad.dynamics() is None                                             # state of the wrong size
    integrator = cds.LinearTimeInvariantSystemRK4Integrator()
    assert not integrator.integrate(timedelta(0), timedelta(seconds=dt))       # no system
    assert integrator.set_dynamical_system(bad) and not integrator.set_dynamical_system(bad)
    assert not integrator.set_integration_step(timedelta(0))
    assert integrator.set_integration_step(timedelta(seconds=dt))
    assert not integrator.integrate(timedelta(0), timedelta(seconds=dt))       # the dynamics cannot be evaluated
    assert not integrator.integrate(timedelta(0), timedelta(seconds=dt / 2))   # shorter than the step


@pytest.mark.gpu
def test_floating_base_system_acceleration_kinematics():
    """FloatingBaseSystemAccelerationKinematics under ForwardEuler on the device: constant accelerations against the closed forms
    v = v0 + a T and (explicit Euler, derivative at the old state) p_K = p0 + K dt v0 + dt^2 a K (K - 1) / 2; with no angular
    acceleration the rotation is Exp(omega T) R0 (left-plus).  One state, then a batch."""
    cds = blf.continuous_dynamical_system
    rng = np.random.default_rng(5)
    dt, K = 0.001, 200
    for N in (1, 5):
        shape = (N,) if N > 1 else ()
        p0, s0 = rng.uniform(-1, 1, shape + (3,)), rng.uniform(-1, 1, shape + (23,))
        dp0, w0, ds0 = rng.uniform(-1, 1, shape + (3,)), rng.uniform(-1, 1, shape + (3,)), rng.uniform(-1, 1, shape + (23,))
        phi0 = rng.uniform(-1, 1, shape + (3,))
        R0 = manif.SO3Tangent(phi0).exp() if N == 1 else W.so3_exp(phi0)
        dtw = np.concatenate([rng.uniform(-1, 1, shape + (3,)), np.zeros(shape + (3,))], axis=-1)
        dds = rng.uniform(-1, 1, shape + (23,))
        system = cds.FloatingBaseSystemAccelerationKinematics()
        system.state = p0, R0, s0, dp0, w0, ds0
        assert system.set_control_input((dtw, dds))
        assert not system.set_control_input((dtw, dds[..., :-1]))
        integrator = cds.FloatingBaseSystemAccelerationKinematicsForwardEulerIntegrator()
        assert integrator.set_dynamical_system(system)
        integrator.integration_step = timedelta(seconds=dt)
        assert integrator.integrate(timedelta(0), timedelta(seconds=K * dt))
        p, R, s, dp, w, ds = integrator.get_solution()
        T_, quad = K * dt, dt * dt * K * (K - 1) / 2
        assert dp == pytest.approx(dp0 + T_ * dtw[..., :3], abs=1e-12) and ds == pytest.approx(ds0 + T_ * dds, abs=1e-12)
        assert p == pytest.approx(p0 + T_ * dp0 + quad * dtw[..., :3], abs=1e-12)
        assert s == pytest.approx(s0 + T_ * ds0 + quad * dds, abs=1e-12)
        assert w == pytest.approx(w0, abs=1e-15)
        if N == 1:
            exact = manif.SO3Tangent(w0 * T_) + R0
            assert R.rotation() == pytest.approx(exact.rotation(), abs=1e-10)
        else:
            exact = W.quat_mul(W.so3_exp(w0 * T_), R0)
            assert np.abs(W.quat_to_rot(R) - W.quat_to_rot(exact)).max() < 1e-10


@pytest.mark.gpu
def test_kindyn_dynamic_quantities():
    """KinDynComputations' dynamic queries under iDynTree's names (what the TSID tasks read): the mass matrix is symmetric with the
    total mass on the base translation block, bias forces at zero velocity are the gravity forces, the frame velocity is J nu, and
    the Coriolis part is quadratic in the velocity."""
    model = M.ergocub23()
    n, nv = model.n, model.n + 6
    _, kindyn = get_kindyn(model, 2)
    rng = np.random.default_rng(8)
    H = np.eye(4)
    H[:3, 3] = [0.1, -0.2, 0.7]
    s, v, sd = model.q_nominal + rng.uniform(-0.3, 0.3, n), rng.uniform(-1, 1, 6), rng.uniform(-1, 1, n)
    g = np.array([0.0, 0.0, -9.81])
    assert kindyn.setRobotState(H, s, v, sd, g)
    Mm, G, h = kindyn.getFreeFloatingMassMatrix(), kindyn.generalizedGravityForces(), kindyn.generalizedBiasForces()
    assert Mm.shape == (nv, nv) and np.abs(Mm - Mm.T).max() < 1e-10 and np.abs(Mm[:3, :3] - model.mass.sum() * np.eye(3)).max() < 1e-9
    assert abs(G[2] - 9.81 * model.mass.sum()) < 1e-8
    J = kindyn.getFrameFreeFloatingJacobian("l_sole")
    assert kindyn.getFrameVel("l_sole") == pytest.approx(J @ np.concatenate([v, sd]), abs=1e-11)
    a1 = kindyn.getFrameBiasAcc("l_sole")
    assert kindyn.setRobotState(H, s, 2 * v, 2 * sd, g)
    assert kindyn.generalizedBiasForces() - G == pytest.approx(4 * (h - G), abs=1e-8)
    assert kindyn.getFrameBiasAcc("l_sole") == pytest.approx(4 * a1, abs=1e-9)
    assert kindyn.setRobotState(H, s, 0 * v, 0 * sd, g)
    assert kindyn.generalizedBiasForces() == pytest.approx(G, abs=1e-9) and np.abs(kindyn.getCenterOfMassBiasAcc()).max() == 0.0


@pytest.mark.gpu
def test_build_closed_loop_batch_and_fixed_base():
    """QPInverseKinematics.build from a parameter handler (QPInverseKinematicsTest.cpp:642-798) on a batch of 16 ergoCubs,
    closed with the Python integrator; then QPFixedBaseInverseKinematics: zero base twist on every tick."""
    model = M.ergocub23()
    N = 16
    _, kindyn = get_kindyn(model, N)
    n = model.n
    h = blf.parameters_handler.StdParametersHandler()
    h.set_parameter_vector_string("tasks", ["COM_TASK", "LEFT_FOOT_TASK", "RIGHT_FOOT_TASK", "TORSO_TASK", "REG_TASK", "LIMITS_TASK"])
    ik_group = _handler()
    h.set_group("IK", ik_group)

    def group(type_, priority, **kw):
        g = _handler(**kw)
        g.set_parameter_string("type", type_)
        g.set_parameter_int("priority", priority)
        return g
    h.set_group("COM_TASK", group("CoMTask", 0, kp_linear=2.0))
    h.set_group("LEFT_FOOT_TASK", group("SE3Task", 0, frame_name="l_sole", kp_linear=7.0, kp_angular=5.0))
    h.set_group("RIGHT_FOOT_TASK", group("SE3Task", 0, frame_name="r_sole", kp_linear=7.0, kp_angular=5.0))
    h.set_group("TORSO_TASK", group("SO3Task", 1, frame_name="chest", kp_angular=5.0, weight=[5.0, 5.0, 5.0]))
    h.set_group("REG_TASK", group("JointTrackingTask", 1, kp=[5.0] * n, weight=[1.0] * n))
    h.set_group("LIMITS_TASK", group("JointVelocityLimitsTask", 0, upper_limits=[2.0] * n, lower_limits=[-2.0] * n))
    variables, weights, ik = blf.ik.QPInverseKinematics.build(param_handler=h, kin_dyn=kindyn)
    assert ik is not None and set(weights) == {"TORSO_TASK", "REG_TASK"} and variables.get_number_of_variables() == n + 6

    rng = np.random.default_rng(5)
    from blf_b200 import workloads as W
    s_des = np.clip(0.5 * (model.q_lower + model.q_upper)[None] + rng.uniform(-0.15, 0.15, (N, n)), model.q_lower, model.q_upper)
    p0 = np.tile([0.0, 0.0, 0.75], (N, 1))
    q0 = W.so3_exp(np.tile([0.0, 0.05, 0.1], (N, 1)))
    desired = blf.KinDynComputations(model, N)
    desired.set_robot_state_batch(p0, q0, s_des)
    k = desired.batch_kinematics(["l_sole", "r_sole", "chest"], jac=False, com=True)
    quat = W.rot_to_quat(k["frame_rot"].reshape(-1, 3, 3)).reshape(N, 3, 4)
    zeros6 = np.zeros((N, 6))
    assert ik.get_task("COM_TASK").set_set_point(position=k["com"])
    assert ik.get_task("LEFT_FOOT_TASK").set_set_point(np.hstack([k["frame_pos"][:, 0], quat[:, 0], zeros6]))
    assert ik.get_task("RIGHT_FOOT_TASK").set_set_point(np.hstack([k["frame_pos"][:, 1], quat[:, 1], zeros6]))
    assert not ik.advance()                                    # TORSO / REG set-points still missing
    assert ik.get_task("TORSO_TASK").set_set_point(np.hstack([quat[:, 2], np.zeros((N, 3))]))
    assert ik.get_task("REG_TASK").set_set_point(joint_position=s_des)

    system = blf.continuous_dynamical_system.FloatingBaseSystemVelocityKinematics()
    system.state = p0.copy(), q0.copy(), s_des + rng.normal(0, 0.05, (N, n))
    integrator = blf.continuous_dynamical_system.FloatingBaseSystemVelocityKinematicsForwardEulerIntegrator()
    assert integrator.set_dynamical_system(system)
    integrator.integration_step = timedelta(milliseconds=10)
    for _ in range(120):
        p, q, s = integrator.get_solution()
        assert kindyn.set_robot_state_batch(p, q, s)
        assert ik.advance() and ik.is_output_valid()
        nu = ik.get_batch_output()
        assert np.abs(nu[:, 6:]).max() <= 2.0 + 1e-9
        assert system.set_control_input((nu[:, :6], nu[:, 6:]))
        assert integrator.integrate(timedelta(0), timedelta(milliseconds=10))
    p, q, s = integrator.get_solution()
    kindyn.set_robot_state_batch(p, q, s)
    kk = kindyn.batch_kinematics(["l_sole"], jac=False, com=True)
    assert np.abs(kk["com"] - k["com"]).max() < 2e-3 and np.abs(kk["frame_pos"][:, 0] - k["frame_pos"][:, 0]).max() < 2e-3

    # fixed base
    fixed = blf.ik.QPFixedBaseInverseKinematics()
    assert not fixed.initialize(handler=_handler())            # set_kin_dyn first
    assert fixed.set_kin_dyn(kindyn) and fixed.initialize(handler=_handler())
    assert fixed.get_task_names() == ["base_se3_task"]
    hand = blf.ik.SE3Task()
    assert hand.set_kin_dyn(kindyn) and hand.initialize(param_handler=_handler(frame_name="l_hand_palm", kp_linear=5.0, kp_angular=5.0))
    reg = blf.ik.JointTrackingTask()
    assert reg.set_kin_dyn(kindyn) and reg.initialize(param_handler=_handler(kp=[1.0] * n))
    assert fixed.add_task(hand, "hand", 0) and fixed.add_task(reg, "reg", 1, weight=[1.0] * n)
    var = blf.system.VariablesHandler()
    assert var.add_variable("robotVelocity", n + 6) and fixed.finalize(var)
    kh = desired.batch_kinematics(["l_hand_palm"], jac=False, com=False)
    qh = W.rot_to_quat(kh["frame_rot"].reshape(-1, 3, 3))
    assert hand.set_set_point(np.hstack([kh["frame_pos"][:, 0], qh, zeros6])) and reg.set_set_point(joint_position=s_des)
    system2 = blf.continuous_dynamical_system.FloatingBaseSystemVelocityKinematics()
    system2.state = p0.copy(), q0.copy(), s_des + rng.normal(0, 0.05, (N, n))
    integrator2 = blf.continuous_dynamical_system.FloatingBaseSystemVelocityKinematicsForwardEulerIntegrator()
    assert integrator2.set_dynamical_system(system2)
    integrator2.integration_step = timedelta(milliseconds=10)
    for _ in range(100):
        p, q, s = integrator2.get_solution()
        kindyn.set_robot_state_batch(p, q, s)
        assert fixed.advance()
        nu = fixed.get_batch_output()
        assert np.abs(nu[:, :6]).max() < 1e-9
        system2.set_control_input((nu[:, :6], nu[:, 6:]))
        assert integrator2.integrate(timedelta(0), timedelta(milliseconds=10))
    p, q, s = integrator2.get_solution()
    kindyn.set_robot_state_batch(p, q, s)
    assert np.abs(kindyn.batch_kinematics(["l_hand_palm"], jac=False, com=False)["frame_pos"] - kh["frame_pos"]).max() < 2e-3
    assert np.abs(p - p0).max() < 1e-9


@pytest.mark.gpu
def test_exogenous_feedback_through_the_bindings():
    """set_feedback (bindings/python/IK/src/SE3Task.cpp, CoMTask.cpp): with the measured pose equal to the model's the solution
    equals the plain tasks'; a shifted CoM measurement changes it by the controller law."""
    model = M.ergocub23()
    n = model.n

    def solve(exo, com_fb=None):
        _, kindyn = get_kindyn(model, 2)
        H = np.eye(4)
        H[:3, 3] = [0.0, 0.0, 0.75]
        assert kindyn.setRobotState(H, model.q_nominal)
        ik = blf.ik.QPInverseKinematics()
        assert ik.initialize(handler=_handler())
        foot = blf.ik.SE3Task()
        assert foot.set_kin_dyn(kindyn) and foot.initialize(param_handler=_handler(frame_name="l_sole", kp_linear=5.0, kp_angular=5.0,
                                                                                  use_position_exogenous_feedback=exo, use_orientation_exogenous_feedback=exo))
        com = blf.ik.CoMTask()
        assert com.set_kin_dyn(kindyn) and com.initialize(param_handler=_handler(kp_linear=3.0, use_exogenous_feedback=exo))
        reg = blf.ik.JointTrackingTask()
        assert reg.set_kin_dyn(kindyn) and reg.initialize(param_handler=_handler(kp=[1.0] * n))
        pose, c = kindyn.getWorldTransform("l_sole"), kindyn.getCenterOfMassPosition()
        from blf_b200 import workloads as W
        q = W.rot_to_quat(pose[None, :3, :3])[0]
        assert foot.set_set_point(manif.SE3(position=pose[:3, 3] + [0.01, 0.0, 0.0], quaternion=q))
        assert com.set_set_point(position=c + [0.0, 0.01, 0.0]) and reg.set_set_point(joint_position=model.q_nominal)
        if exo:
            assert foot.set_feedback(I_H_F=manif.SE3(position=pose[:3, 3], quaternion=q))
            assert com.set_feedback(c if com_fb is None else c + com_fb)
        assert ik.add_task(foot, "foot", 0) and ik.add_task(com, "com", 0) and ik.add_task(reg, "reg", 1, weight=[1.0] * n)
        var = blf.system.VariablesHandler()
        assert var.add_variable("robotVelocity", n + 6) and ik.finalize(var) and ik.advance()
        return ik.get_raw_solution().copy(), kindyn.getCenterOfMassJacobian()

    nu_plain, Jc = solve(False)
    nu_exo, _ = solve(True)
    assert np.abs(nu_plain - nu_exo).max() < 1e-9
    shift = np.array([0.0, 0.0, 0.02])
    nu_shift, _ = solve(True, shift)
    assert np.abs(Jc @ (nu_shift - nu_plain) + 3.0 * shift).max() < 1e-8      # J_com nu = v_ff + kp (p_d - p_measured)


# ------------------------------------------------------------------------------------------------ Math: splines
def _quintic(c, t, k=0):
    return sum(c[i] * np.prod(np.arange(i, i - k, -1)) * t ** (i - k) for i in range(k, 6))


def test_quintic_spline():
    """bindings/python/Math/tests/test_quintic_spline.py replayed: 100 knots on a random quintic between 0.32 s and 2.64 s, the
    spline returns the polynomial and its derivatives (atol 1e-5) -- plus the advance() / evaluate_ordered_points sections
    and the refusals of src/Math/tests/SplineTest.cpp."""
    rng = np.random.default_rng(11)
    number_of_points, init_time, final_time = 100, 0.32, 2.64
    dT = (final_time - init_time) / (number_of_points - 1)
    coefficients = [rng.random((4, 1)) for _ in range(6)]
    time = [init_time + dT * i for i in range(number_of_points)]
    knots = [_quintic(coefficients, t) for t in time]
    spline = blf.math.QuinticSpline()
    position, velocity, acceleration = np.zeros(4), np.zeros(4), np.zeros(4)
    assert not spline.evaluate_point(1.0, position)                                     # no knots yet
    assert not spline.set_knots(knots[:1], time[:1]) and not spline.set_knots(knots, time[:-1])
    assert spline.set_knots(knots, time)
    assert not spline.evaluate_point(1.0, position)                                     # no boundary conditions yet
    assert not spline.set_initial_conditions(np.zeros(4), np.zeros(3))
    assert spline.set_initial_conditions(_quintic(coefficients, init_time, 1), _quintic(coefficients, init_time, 2))
    assert spline.set_final_conditions(_quintic(coefficients, final_time, 1), _quintic(coefficients, final_time, 2))
    points_to_check_number = 10000
    dT_check_points = (final_time - init_time) / points_to_check_number
    for i in range(0, points_to_check_number, 3):
        t = dT_check_points * i + init_time
        assert spline.evaluate_point(t, position, velocity, acceleration)
        assert np.allclose(_quintic(coefficients, t).T[0], position, atol=1e-5)
        assert np.allclose(_quintic(coefficients, t, 1).T[0], velocity, atol=1e-5)
        assert np.allclose(_quintic(coefficients, t, 2).T[0], acceleration, atol=1e-5)
    # "Advance capabilities"
    assert not spline.is_output_valid()
    assert not spline.set_advance_time_step(timedelta(0)) and spline.set_advance_time_step(timedelta(seconds=dT_check_points * 50))
    for i in range(0, points_to_check_number, 50):
        t = dT_check_points * i + init_time
        assert spline.advance() and spline.is_output_valid()
        assert np.allclose(_quintic(coefficients, t).T[0], spline.get_output().position, atol=1e-5)
        assert np.allclose(_quintic(coefficients, t, 1).T[0], spline.get_output().velocity, atol=1e-5)
    # "Query from a vector of times", and timedelta arguments as in the reference's binding
    ts = [dT_check_points * i + init_time for i in range(0, points_to_check_number, 97)]
    ok, ps, vs, acs = spline.evaluate_ordered_points(ts)
    assert ok and all(np.allclose(_quintic(coefficients, t).T[0], p, atol=1e-5) for t, p in zip(ts, ps))
    assert all(np.allclose(_quintic(coefficients, t, 2).T[0], a, atol=1e-5) for t, a in zip(ts, acs))
    assert spline.evaluate_point(timedelta(seconds=1.0), position)
    assert np.allclose(_quintic(coefficients, 1.0).T[0], position, atol=1e-5)
    # two consecutive knots with the same time are refused when the coefficients are computed (Spline.h:444-448)
    twin = blf.math.QuinticSpline()
    assert twin.set_knots(knots[:3], [0.0, 0.5, 0.5])
    assert twin.set_initial_conditions(np.zeros(4), np.zeros(4)) and twin.set_final_conditions(np.zeros(4), np.zeros(4))
    assert not twin.evaluate_point(0.2, position)


@pytest.mark.gpu
def test_set_point_spline_through_the_bindings():
    """The batch extension set_set_point_spline: the CoM task of three robots follows one QuinticSpline object per robot on
    the device; at every queried time the solution equals the one obtained the reference's way (evaluate the spline on the
    host, set_set_point, advance)."""
    model = M.ergocub23()
    n, N = model.n, 3

    def make():
        _, kindyn = get_kindyn(model, N)
        H = np.eye(4)
        H[:3, 3] = [0.0, 0.0, 0.75]
        assert kindyn.setRobotState(H, model.q_nominal)
        ik = blf.ik.QPInverseKinematics()
        assert ik.initialize(handler=_handler())
        com = blf.ik.CoMTask()
        assert com.set_kin_dyn(kindyn) and com.initialize(param_handler=_handler(kp_linear=3.0))
        reg = blf.ik.JointTrackingTask()
        assert reg.set_kin_dyn(kindyn) and reg.initialize(param_handler=_handler(kp=[1.0] * n))
        c = kindyn.getCenterOfMassPosition()
        assert com.set_set_point(position=c) and reg.set_set_point(joint_position=model.q_nominal)
        assert ik.add_task(com, "com", 0) and ik.add_task(reg, "reg", 1, weight=[1.0] * n)
        var = blf.system.VariablesHandler()
        assert var.add_variable("robotVelocity", n + 6) and ik.finalize(var)
        return ik, com, c, reg

    rng = np.random.default_rng(4)
    ik_dev, com_dev, c, reg_dev = make()
    ik_host, com_host, _, reg_host = make()
    times = [0.0, 0.3, 0.5, 1.0]
    splines = []
    for i in range(N):
        sp = blf.math.QuinticSpline()
        assert sp.set_knots([c + rng.uniform(-0.01, 0.01, 3) * k for k in range(4)], times)
        assert sp.set_initial_conditions(rng.uniform(-0.01, 0.01, 3), np.zeros(3)) and sp.set_final_conditions(np.zeros(3), np.zeros(3))
        splines.append(sp)
    reg_only = blf.ik.SO3Task()
    assert not reg_only.set_set_point_spline(splines[0])                # no position set-point
    assert not blf.ik.JointTrackingTask().set_set_point_spline(splines[0])   # no kinDyn yet
    assert not reg_dev.set_set_point_spline(splines)                    # R^3 knots, n joints
    assert not com_dev.set_set_point_spline(splines[:2])                # one per robot
    assert com_dev.set_set_point_spline(splines)
    pos, vel = np.zeros(3), np.zeros(3)
    for t in (-0.1, 0.0, 0.12, 0.3, 0.77, 1.0, 1.4):
        assert ik_dev.set_time(t) and ik_dev.advance()
        for i in range(N):
            assert splines[i].evaluate_point(t, pos, vel)
            assert com_host.set_set_point(position=pos, velocity=vel, instance=i)
        assert ik_host.advance()
        assert np.abs(ik_dev.get_batch_output() - ik_host.get_batch_output()).max() < 1e-8
    assert com_dev.set_set_point_spline(None) and ik_dev.advance()
    assert com_host.set_set_point(position=c) and ik_host.advance()
    assert np.abs(ik_dev.get_batch_output() - ik_host.get_batch_output()).max() < 1e-8
    # the minimum-jerk generator through the task: from c to c + d over 0.5 s, starting at the solver's time of the next advance
    d = np.array([0.01, -0.02, 0.005])
    assert not reg_dev.set_set_point_minimum_jerk(np.zeros(2 * n), 0.5)   # no such generator in joint space
    assert com_dev.set_set_point_minimum_jerk(np.concatenate([c + d, np.zeros(3)]), timedelta(seconds=0.5))
    assert ik_dev.set_time(2.0) and ik_dev.advance()                     # tau = 0: still at c
    assert np.abs(ik_dev.get_batch_output() - ik_host.get_batch_output()).max() < 1e-8
    for tau in (0.3, 0.8, 1.0, 1.7):
        sv = 10 * min(tau, 1.0) ** 3 - 15 * min(tau, 1.0) ** 4 + 6 * min(tau, 1.0) ** 5
        sd = 30 * tau ** 2 * (1 - tau) ** 2 / 0.5 if tau < 1.0 else 0.0
        assert ik_dev.set_time(2.0 + 0.5 * tau) and ik_dev.advance()
        assert com_host.set_set_point(position=c + sv * d, velocity=sd * d) and ik_host.advance()
        assert np.abs(ik_dev.get_batch_output() - ik_host.get_batch_output()).max() < 1e-8
    assert com_dev.set_set_point_minimum_jerk(None, 0.0) and ik_dev.advance()
    assert com_host.set_set_point(position=c) and ik_host.advance()
    assert np.abs(ik_dev.get_batch_output() - ik_host.get_batch_output()).max() < 1e-8

    # a joint-space spline (QuinticSpline / CubicSpline over R^n) on the JointTrackingTask, one object per robot
    for cls in (blf.math.QuinticSpline, blf.math.CubicSpline):
        jsplines = []
        for i in range(N):
            sp = cls()
            assert sp.set_knots([model.q_nominal + rng.uniform(-0.02, 0.02, n) * k for k in range(4)], times)
            assert sp.set_initial_conditions(rng.uniform(-0.02, 0.02, n), np.zeros(n)) and sp.set_final_conditions(np.zeros(n), np.zeros(n))
            jsplines.append(sp)
        assert reg_dev.set_set_point_spline(jsplines)
        jp, jv = np.zeros(n), np.zeros(n)
        for t in (-0.1, 0.0, 0.12, 0.3, 0.77, 1.0, 1.4):
            assert ik_dev.set_time(t) and ik_dev.advance()
            for i in range(N):
                assert jsplines[i].evaluate_point(t, jp, jv)
                assert reg_host.set_set_point(joint_position=jp, joint_velocity=jv, instance=i)
            assert ik_host.advance()
            assert np.abs(ik_dev.get_batch_output() - ik_host.get_batch_output()).max() < 1e-8
        assert reg_dev.set_set_point_spline(None)
    assert reg_host.set_set_point(joint_position=model.q_nominal) and ik_host.advance() and ik_dev.advance()
    assert np.abs(ik_dev.get_batch_output() - ik_host.get_batch_output()).max() < 1e-8


# ------------------------------------------------------------------------------------------------ Planners: SO3Planner
@pytest.mark.parametrize("kind", ["body", "inertial", "inertial_v0"])
def test_so3_planner(kind):
    """src/Planners/tests/SO3PlannerTest.cpp:21-167 replayed: between two random rotations over 1 s, sampled every 100 us, the
    rotation integrates its own velocity (right-plus for the body-fixed planner, left-plus for the inertial one), the velocity
    its acceleration, and both ends are met at rest (tolerance 1e-3); third section: a non-zero initial velocity along the
    displacement."""
    rng = np.random.default_rng(8)
    R0, R1 = manif.SO3.Random(rng), manif.SO3.Random(rng)
    T, dT, tol = 1.0, 100e-6, 1e-3
    planner = blf.planners.SO3PlannerBody() if kind == "body" else blf.planners.SO3PlannerInertial()
    assert not planner.set_rotations(R0, R1, 0.0)
    assert planner.set_rotations(R0, R1, timedelta(seconds=T))
    assert planner.evaluate_point(0.0) is False                                  # boundary conditions missing
    v0 = manif.SO3Tangent.Zero()
    if kind == "inertial_v0":
        v0 = (R1 * R0.inverse()).log() * 2.0
    assert planner.set_initial_conditions(v0, manif.SO3Tangent.Zero())
    assert planner.set_final_conditions(manif.SO3Tangent.Zero(), manif.SO3Tangent.Zero())
    st = planner.evaluate_point(0.0)
    near = lambda A, B: np.linalg.norm((A - B).coeffs()) < tol
    assert near(st.rotation, R0) and np.allclose(st.velocity.coeffs(), v0.coeffs(), atol=tol) and np.allclose(st.acceleration.coeffs(), 0, atol=tol)
    for i in range(1, int(round(T / dT))):
        step = st.velocity * dT
        predicted = (st.rotation + step) if kind == "body" else (step + st.rotation)
        predicted_velocity = st.velocity + st.acceleration * dT
        st = planner.evaluate_point(i * dT)
        assert near(predicted, st.rotation)
        assert np.allclose(predicted_velocity.coeffs(), st.velocity.coeffs(), atol=tol)
    st = planner.evaluate_point(T)
    assert near(st.rotation, R1) and np.allclose(st.velocity.coeffs(), 0, atol=tol) and np.allclose(st.acceleration.coeffs(), 0, atol=tol)
    assert planner.evaluate_point(T + 0.1) is False and planner.evaluate_point(-0.1) is False   # :178-184
    # a boundary velocity that is not along the displacement is refused (:89-99)
    skew = blf.planners.SO3PlannerInertial()
    assert skew.set_rotations(R0, R1, T)
    d = (R1 * R0.inverse()).log().coeffs()
    assert skew.set_initial_conditions(manif.SO3Tangent(np.cross(d, [1.0, 0.3, -0.2])), manif.SO3Tangent.Zero())
    assert skew.set_final_conditions(manif.SO3Tangent.Zero(), manif.SO3Tangent.Zero())
    assert skew.evaluate_point(0.5) is False
    # advance()
    assert not planner.advance() and planner.set_advance_time_step(timedelta(seconds=0.25)) and planner.is_output_valid()
    for _ in range(6):
        assert planner.advance()
    assert near(planner.get_output().rotation, R1)
    pv = blf.planners.SO3PlannerInertial.project_tangent_vector(R0, R1, manif.SO3Tangent([0.3, -0.1, 0.2]))
    assert np.linalg.norm(np.cross(pv.coeffs(), d)) < 1e-12


@pytest.mark.gpu
def test_device_minimum_jerk_generator_is_the_inertial_so3_planner():
    """blfik_set_setpoint_minimum_jerk on an SO3 task == SO3PlannerInertial with zero boundary conditions, sample by sample
    (rotation and angular velocity), read back with blfik_evaluate_setpoint."""
    from blf_b200 import capi, tasks as T, workloads as W
    model = M.ergocub23()
    task = T.so3_task(model, "chest", "chest", [5.0, 5.0, 5.0], priority=1, weight=[1.0, 1.0, 1.0])
    reg = T.joint_tracking_task(model, "reg", [1.0] * model.n, priority=1, weight=[1.0] * model.n)
    N, dur, t0 = 5, 0.8, 2.5
    rng = np.random.default_rng(2)
    R0 = [manif.SO3.Random(rng) for _ in range(N)]
    R1 = [manif.SO3.Random(rng) for _ in range(N)]
    cm = capi.Model(model)
    prob = capi.Problem(cm, [task, reg], N)
    sp = np.zeros((N, 7))
    sp[:, :4] = [r.quat() for r in R0]
    prob.set_setpoint(0, sp)
    tg = sp.copy()
    tg[:, :4] = [r.quat() for r in R1]
    prob.set_time(t0)
    prob.set_setpoint_minimum_jerk(0, tg, dur)
    planners = []
    for a, b in zip(R0, R1):
        p = blf.planners.SO3PlannerInertial()
        assert p.set_rotations(a, b, dur)
        assert p.set_initial_conditions(manif.SO3Tangent.Zero(), manif.SO3Tangent.Zero())
        assert p.set_final_conditions(manif.SO3Tangent.Zero(), manif.SO3Tangent.Zero())
        planners.append(p)
    for tau in (0.0, 0.1, 0.37, 0.5, 0.93, 1.0):
        got = prob.evaluate_setpoint(0, t0 + tau * dur)
        for i, p in enumerate(planners):
            st = p.evaluate_point(tau * dur)
            q = st.rotation.quat()
            assert min(np.abs(got[i, :4] - q).max(), np.abs(got[i, :4] + q).max()) < 1e-12
            assert np.abs(got[i, 4:7] - st.velocity.coeffs()).max() < 1e-11
    prob.close()


def test_cubic_spline():
    """src/Math/tests/SplineTest.cpp "Cubic spline" (:185-330) through the Python class: knots on a random cubic, position and
    velocity reproduced; the acceleration is continuous across the knots."""
    rng = np.random.default_rng(13)
    coefficients = [rng.random((4, 1)) for _ in range(4)] + [np.zeros((4, 1))] * 2
    init_time, final_time = 0.32, 2.64
    time = [init_time + (final_time - init_time) / 99 * i for i in range(100)]
    spline = blf.math.CubicSpline()
    assert spline.set_knots([_quintic(coefficients, t) for t in time], time)
    assert spline.set_initial_conditions(_quintic(coefficients, init_time, 1), _quintic(coefficients, init_time, 2))
    assert spline.set_final_conditions(_quintic(coefficients, final_time, 1), _quintic(coefficients, final_time, 2))
    position, velocity, acceleration = np.zeros(4), np.zeros(4), np.zeros(4)
    for i in range(0, 10000, 7):
        t = init_time + (final_time - init_time) / 1e4 * i
        assert spline.evaluate_point(t, position, velocity, acceleration)
        assert np.allclose(_quintic(coefficients, t).T[0], position, atol=1e-5)
        assert np.allclose(_quintic(coefficients, t, 1).T[0], velocity, atol=1e-5)
        assert np.allclose(_quintic(coefficients, t, 2).T[0], acceleration, atol=1e-4)
