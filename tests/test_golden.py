"""The oracle against the committed fixtures (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

import blf_b200  # noqa: F401
from blf_b200 import workloads as W
from oracle.oracle import Oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("cid", [1, 3, 4])
def test_oracle_reproduces_golden(cid):
    g = np.load(os.path.join(GOLD, f"golden_c{cid}.npz"))
    w = W.config(cid)
    N = g["s"].shape[0]
    inp = W.synth_inputs(w, N, cid)
    # the workload generator itself is part of the fixture
    assert np.array_equal(inp.s, g["s"]) and np.array_equal(inp.base_quat, g["base_quat"])
    for t, sp in enumerate(inp.setpoints):
        if sp is not None:
            assert np.array_equal(sp, g[f"sp{t}"])
    o = Oracle(w.model, w.tasks)
    p, q, s = inp.base_pos.copy(), inp.base_quat.copy(), inp.s.copy()
    nu, status, _ = o.tick_batch(p, q, s, inp.setpoints, dt_integrate=w.dt)
    assert np.array_equal(status, g["status"])
    assert np.abs(nu - g["nu"]).max() < 1e-9
    assert np.abs(s - g["next_s"]).max() < 1e-11 and np.abs(p - g["next_pos"]).max() < 1e-11
    for i in range(N):
        R, P = o.fk(inp.base_pos[i], inp.base_quat[i], inp.s[i])
        for fi, f in enumerate(g["frames"]):
            assert np.abs(o.frame_jacobian(R, P, int(f)) - g["frame_jac"][i, fi]).max() < 1e-13
        c, J = o.com(R, P)
        assert np.abs(c - g["com"][i]).max() < 1e-13 and np.abs(J - g["com_jac"][i]).max() < 1e-13
