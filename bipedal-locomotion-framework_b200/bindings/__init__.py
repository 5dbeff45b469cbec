"""Python surface of the batched engine, laid out like the reference's ``bipedal_locomotion_framework.bindings``
(bindings/python/**): ``parameters_handler``, ``system``, ``ik``, ``continuous_dynamical_system``, ``math``, ``planners``.

    import blf_b200.bindings as blf
    handler = blf.parameters_handler.StdParametersHandler()
    task = blf.ik.SE3Task(); task.set_kin_dyn(kindyn); task.initialize(param_handler=handler)
    qp_ik = blf.ik.QPInverseKinematics(); ...; qp_ik.advance(); qp_ik.get_output().joint_velocity

Same class / method / keyword names and the same "return False and log" error convention as the reference bindings
(bindings/python/IK/src/*.cpp, bindings/python/System/src/VariablesHandler.cpp, bindings/python/ParametersHandler/src/
Module.cpp, bindings/python/ContinuousDynamicalSystem/src/*.cpp).  Differences, all additive: ``kindyn`` is a
:class:`KinDynComputations` of this package (a batch of N robots on one kinematic tree; iDynTree is not in this image),
``manif`` is the small stand-in in :mod:`.manif`, and every ``set_set_point`` / state setter also accepts arrays with a
leading batch dimension.  Everything on the IK tick (kinematics, task assembly, QP, integration) runs in the CUDA library
through :mod:`..capi`; there is no CPU path for it.  The one-trajectory planner classes (``math.QuinticSpline``,
``planners.SO3Planner*``) are host code as in the reference; for a batch their trajectories are evaluated on the device
(``IKLinearTask.set_set_point_spline``, ``capi.Problem.set_setpoint_minimum_jerk``).
"""
from . import manif, math, parameters_handler, system, continuous_dynamical_system, ik, planners  # noqa: F401
from .kindyn import KinDynComputations  # noqa: F401
