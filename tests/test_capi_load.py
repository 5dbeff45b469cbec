"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/blfik.h declares, validates its
arguments, and FAILS LOUDLY without a CUDA device (no CPU fallback).  No compute call is made here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import blf_b200  # noqa: F401
from blf_b200 import capi, models as M, tasks as T, workloads as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    header = open(os.path.join(ROOT, "include", "blfik.h")).read()
    declared = set(re.findall(r"\b(blfik_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    for sym in declared:
        assert hasattr(lib, sym), f"{sym} declared in include/blfik.h but not exported"
    assert declared == set(capi.EXPORTS)


def test_default_settings_and_model_validation():
    lib = capi.load()
    s = capi.SolverSettings()
    lib.blfik_default_settings(C.byref(s))
    assert (s.rho, s.sigma, s.alpha, s.eps) == (1.0, 1e-2, 1.6, 1e-7) and s.admm_iters == 10
    km = M.ergocub23()
    model = capi.Model(km)          # host-only: works without a GPU
    assert model.handle
    bad = M.ergocub23()
    bad.parent = bad.parent.copy()
    bad.parent[3] = 7               # not topologically ordered
    with pytest.raises(capi.BlfikError) as e:
        capi.Model(bad)
    assert e.value.code == -1 and "topologically" in str(e.value)
    bad2 = M.ergocub23()
    bad2.dof = bad2.dof.copy()
    bad2.dof[2] = bad2.dof[1]
    with pytest.raises(capi.BlfikError):
        capi.Model(bad2)


def test_no_cpu_fallback():
    """Without a GPU, creating a problem must fail with a CUDA error -- never run on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    w = W.config(3)
    model = capi.Model(w.model)
    with pytest.raises(capi.BlfikError) as e:
        capi.Problem(model, w.tasks, 8)
    assert e.value.code == -2 and "no CPU path" in str(e.value)


def test_task_spec_rules():
    """addTask rules of QPInverseKinematics.cpp:167-249 enforced by the host-side validation."""
    m = M.ergocub23()
    ok = [T.com_task(m, "c", 1.0), T.joint_tracking_task(m, "j", np.ones(m.n), priority=1, weight=np.ones(m.n))]
    T.validate_tasks(m, ok)
    with pytest.raises(ValueError):
        T.validate_tasks(m, ok + [T.com_task(m, "c", 1.0)])                      # duplicate name
    with pytest.raises(ValueError):
        T.validate_tasks(m, [T.joint_tracking_task(m, "j", np.ones(m.n), priority=1)])   # no weight
    jl = T.joint_limits_task(m, "l", np.full(m.n, 0.5), 0.01)
    jl.priority = 1
    with pytest.raises(ValueError):
        T.validate_tasks(m, [jl])                                                 # inequality as a cost
    with pytest.raises(ValueError):
        T.joint_limits_task(m, "l", np.full(m.n, 1.5), 0.01)                      # klim outside [0,1]
    with pytest.raises(ValueError):
        T.joint_limits_task(m, "l", np.full(m.n, 0.5), 0.0)                       # sampling_time <= 0
    with pytest.raises(KeyError):
        T.se3_task(m, "s", "no_such_frame", 1.0, 1.0)


def test_shard_ranges_partition_and_inputs_are_shard_invariant():
    N = 1000
    for world in (1, 2, 3, 8):
        r = [W.shard_range(N, k, world) for k in range(world)]
        assert r[0][0] == 0 and r[-1][1] == N and all(r[k][1] == r[k + 1][0] for k in range(world - 1))
    w = W.config(3)
    full = W.synth_inputs(w, 64, 3)
    a, b = W.shard_range(64, 1, 2)
    part = full.slice(a, b)
    assert np.array_equal(part.s, full.s[32:]) and np.array_equal(part.setpoints[1], full.setpoints[1][32:])
