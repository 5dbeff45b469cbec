import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, blf_b200
from blf_b200 import models as M, tasks as T, workloads as W
from oracle.oracle import Oracle
import types, re
src = open('tests/test_gpu_parity.py').read()
i0 = src.index("def _ergo_distance_gravity_problem"); i1 = src.index("@pytest.mark.parametrize", i0)
ns = dict(M=M, T=T, np=np); exec(src[i0:i1], ns)
for hard in (True, False):
    m, tasks = ns['_ergo_distance_gravity_problem'](hard)
    n, N = m.n, 96
    rng = np.random.default_rng(17)
    inp = W.synth_inputs(W.config(3), N, 5)
    p0, q0, s0 = inp.base_pos.copy(), inp.base_quat.copy(), inp.s.copy()
    o = Oracle(m, tasks)
    sps = [np.zeros((N, 13)), np.zeros((N, 6)), np.zeros((N, 1)), np.zeros((N, 1)), np.zeros((N, 6)), np.zeros((N, 2 * n)), None]
    for i in range(N):
        R, P = o.fk(p0[i], q0[i], s0[i])
        Rf, pf = o.frame_pose(R, P, tasks[0].frame)
        sps[0][i, :3] = pf; sps[0][i, 3:7] = W.rot_to_quat(Rf[None])[0]
        c, _ = o.com(R, P); sps[1][i, :3] = c + rng.normal(0, 0.01, 3)
        d = o.frame_pose(R, P, tasks[2].frame)[1] - o.frame_pose(R, P, tasks[2].ref_frame)[1]
        sps[2][i, 0] = np.linalg.norm(d) + rng.normal(0, 0.02)
        sps[3][i, 0] = np.linalg.norm(o.frame_pose(R, P, tasks[3].frame)[1]) + rng.normal(0, 0.02)
        Rc, _ = o.frame_pose(R, P, tasks[4].frame)
        sps[4][i, :3] = 1.7 * (Rc[2] + rng.normal(0, 0.03, 3)); sps[4][i, 3:] = rng.normal(0, 0.05, 3)
        sps[5][i, :n] = s0[i] + rng.normal(0, 0.05, n)
    p, q, s = p0.copy(), q0.copy(), s0.copy()
    for tick in range(20):
        nu, st, info = o.tick_batch(p, q, s, sps, dt_integrate=0.01)
        assert np.all(st == 0), (tick, st)
    print(hard, "ok", np.abs(nu).max())
