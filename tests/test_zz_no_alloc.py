"""blfik_kinematics_noalloc: the kinematics entry point for callers that are not allowed to allocate GPU memory
while running -- a hard-real-time control thread (cudaMalloc / cudaFree lock the driver and stall the device), or a
process whose device memory belongs to some other allocator.  It must give exactly what blfik_kinematics gives,
call after call, out of one fixed staging array that every call on the device shares and that is never cleared
between calls, and it must refuse, with a readable message, a request that does not fit that array.

The file name keeps this test last in collection order, so a run that stops at the first failure has already run
everything else; it does not depend on the other tests having run (or not)."""
import ctypes as C

import numpy as np
import pytest

import blf_b200  # noqa: F401
from blf_b200 import capi, models as M, tasks as T, workloads as W

pytestmark = pytest.mark.gpu

TOL_KIN = 1e-12     # the tolerance the kinematics parity tests use
SCRATCH_BYTES = 4 << 20


def _call(lib, fn, prob, frames, want_jac, want_com):
    """Run `fn` (either kinematics entry point) with host outputs; returns the dict of arrays."""
    N, nv, nf = prob.N, prob.nv, len(frames)
    fr = np.ascontiguousarray(frames, dtype=np.int32)
    out = dict(frame_pos=np.full((N, nf, 3), np.nan), frame_rot=np.full((N, nf, 3, 3), np.nan),
               frame_jac=np.full((N, nf, 6, nv), np.nan) if want_jac else None,
               com=np.full((N, 3), np.nan) if want_com else None,
               com_jac=np.full((N, 3, nv), np.nan) if (want_com and want_jac) else None)
    ptr = lambda a: None if a is None or a.size == 0 else a.ctypes.data
    rc = fn(prob.handle, ptr(fr), nf, ptr(out["frame_pos"]), ptr(out["frame_rot"]), ptr(out["frame_jac"]), ptr(out["com"]),
            ptr(out["com_jac"]), capi.MEM_HOST)
    return rc, out


def test_noalloc_kinematics_matches_allocating_entry_point_twice_in_a_row():
    lib = capi.load()
    m = M.ergocub23()
    N = 300
    rng = np.random.default_rng(21)
    tasks = [T.joint_tracking_task(m, "j", np.ones(m.n), priority=1, weight=np.ones(m.n))]
    model = capi.Model(m)
    prob = capi.Problem(model, tasks, N)
    frames_a = [m.frame_index("l_sole"), m.frame_index("r_sole")]
    frames_b = [m.frame_index("chest")]
    # two calls in a row on different states and with different output sets: the second call finds the first
    # call's data still in the staging array and must not be affected by it
    for call, (frames, jac, com) in enumerate([(frames_a, True, True), (frames_b, False, True), (frames_a, True, False)]):
        p = rng.uniform(-1, 1, (N, 3)); q = W.so3_exp(rng.uniform(-1.5, 1.5, (N, 3))); s = m.q_nominal + rng.uniform(-0.6, 0.6, (N, m.n))
        prob.set_state(p, q, s)
        rc, got = _call(lib, lib.blfik_kinematics_noalloc, prob, frames, jac, com)
        assert rc == 0, lib.blfik_last_error().decode()
        rc, ref = _call(lib, lib.blfik_kinematics, prob, frames, jac, com)
        assert rc == 0, lib.blfik_last_error().decode()
        for key, a in got.items():
            if a is None:
                continue
            assert np.isfinite(a).all(), f"call {call}: {key} was not fully written"
            assert np.abs(a - ref[key]).max() <= TOL_KIN * (1 + np.abs(ref[key]).max()), f"call {call}: {key}"
    prob.close()


def test_noalloc_kinematics_rejects_requests_beyond_the_staging_array():
    lib = capi.load()
    m = M.ergocub23()
    N = 4096                               # 4096 robots x 2 Jacobians = 11 MiB > 4 MiB
    tasks = [T.joint_tracking_task(m, "j", np.ones(m.n), priority=1, weight=np.ones(m.n))]
    model = capi.Model(m)
    prob = capi.Problem(model, tasks, N)
    prob.set_state(np.zeros((N, 3)), np.tile([0.0, 0, 0, 1], (N, 1)), np.tile(m.q_nominal, (N, 1)))
    frames = [m.frame_index("l_sole"), m.frame_index("r_sole")]
    rc, _ = _call(lib, lib.blfik_kinematics_noalloc, prob, frames, True, True)
    msg = lib.blfik_last_error().decode()
    assert rc == -3 and "bytes of staging" in msg and str(SCRATCH_BYTES) in msg
    # the same problem still works when the request is small enough (positions only: 4096 x 2 x 12 doubles < 4 MiB)
    rc, got = _call(lib, lib.blfik_kinematics_noalloc, prob, frames, False, False)
    assert rc == 0, lib.blfik_last_error().decode()
    rc, ref = _call(lib, lib.blfik_kinematics, prob, frames, False, False)
    assert rc == 0
    assert np.abs(got["frame_pos"] - ref["frame_pos"]).max() <= TOL_KIN and np.abs(got["frame_rot"] - ref["frame_rot"]).max() <= TOL_KIN
    prob.close()
