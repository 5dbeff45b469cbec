"""Python surface of the batched engine, laid out like the reference's ``bipedal_locomotion_framework.bindings``
(bindings/python/**): ``parameters_handler``, ``system``, ``ik``, ``continuous_dynamical_system``, ``math``.

    import blf_b200.bindings as blf
    handler = blf.parameters_handler.StdParametersHandler()
    task = blf.ik.SE3Task(); task.set_kin_dyn(kindyn); task.initialize(param_handler=handler)
    qp_ik = blf.ik.QPInverseKinematics(); ...; qp_ik.advance(); qp_ik.get_output().joint_velocity

Same class / method / keyword names and the same "return False and log" error convention as the reference bindings
(bindings/python/IK/src/*.cpp, bindings/python/System/src/VariablesHandler.cpp, bindings/python/ParametersHandler/src/
Module.cpp, bindings/python/ContinuousDynamicalSystem/src/*.cpp).  Differences, all additive: ``kindyn`` is a
:class:`KinDynComputations` of this package (a batch of N robots on one kinematic tree; iDynTree is not in this image),
``manif`` is the small stand-in in :mod:`.manif`, and every ``set_set_point`` / state setter also accepts arrays with a
leading batch dimension.  All arithmetic runs in the CUDA library through :mod:`..capi`; there is no CPU path.
"""
from . import manif, math, parameters_handler, system, continuous_dynamical_system, ik  # noqa: F401
from .kindyn import KinDynComputations  # noqa: F401
