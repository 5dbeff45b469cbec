"""Oracle against the real reference (oracle/_ref/libblf_ref_ik.so, built by the recipe of oracle/REF_BUILD.md from the UNMODIFIED
tree under /root/reference).  The libraries the reference's IK path needs are not in this image, so the binary does not exist
here and the comparison is skipped with that reason -- the oracle stays "parity unpinned"."""
import os

import pytest

REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libblf_ref_ik.so")


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/libblf_ref_ik.so is absent: the reference's IK path needs Eigen, iDynTree, manif, "
                                                    "LieGroupControllers and OsqpEigen, none of which is in this image (oracle/REF_BUILD.md)")
def test_oracle_matches_the_reference_on_the_committed_fixtures():
    import ctypes as C
    import numpy as np
    import blf_b200  # noqa: F401
    from blf_b200 import workloads as W
    from oracle.oracle import Oracle
    lib = C.CDLL(REF)
    w = W.config(3)
    inp = W.synth_inputs(w, 16, 3)
    o = Oracle(w.model, w.tasks)
    nu, st, _ = o.tick_batch(inp.base_pos.copy(), inp.base_quat.copy(), inp.s.copy(), inp.setpoints)
    assert hasattr(lib, "ref_ik_tick")
    # the call itself mirrors oracle/ref_shim/blf_ref_ik.cpp; tolerances: nu 1e-5 (tight OSQP), state 1e-9 per tick
    pytest.skip("the shim's set-point and task-group plumbing is completed on the machine that can build it")
