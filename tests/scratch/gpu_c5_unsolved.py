"""C5 closed loop: are the robots that exhaust the iteration budget at a tick infeasible (LP referee) or slow?"""
import sys; sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W, tasks as T
from oracle.oracle import Oracle
from helpers import referee
N = 8192
w = W.config(5)
inp = W.synth_inputs(w, N, 5)
rng = np.random.Generator(np.random.Philox(key=W.SEED + 50))
phases = rng.uniform(0.0, 2.0 * np.pi, size=(N, len(w.tasks)))
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp)
amp = 0.02
for t, task in enumerate(w.tasks):
    if task.type == T.TASK_SE3:
        prob.set_setpoint_trajectory(t, [amp] * 3 + [2 * amp] * 3, 2.0, np.ascontiguousarray(phases[:, t]))
o = Oracle(w.model, w.tasks)
for warm in (0, 5, 50, 300):
    if warm: prob.tick(warm, w.dt, integrate=True, warm_start=True)
    p, q, s = prob.get_state()
    tnow = prob.time
    sps = [prob.evaluate_setpoint(t, tnow) if task.record_len(w.model.n) else None for t, task in enumerate(w.tasks)]
    prob.tick(1, w.dt, integrate=True, warm_start=True)
    st, it = prob.get_status()
    bad = np.flatnonzero(st != 0)
    verdicts = []
    for i in bad[:12]:
        H, g, Cm, l, u = o.assemble(p[i], q[i], s[i], [None if sp is None else sp[i] for sp in sps])
        feas, _ = referee(H, g, Cm, l, u)
        verdicts.append((int(i), int(st[i]), int(it[i]), bool(feas)))
    print(f"after {prob.time / w.dt:.0f} ticks: unsolved {len(bad)}; (robot, status, iters, feasible): {verdicts}")
