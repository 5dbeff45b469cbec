"""C3 at 65,536: the sampled robots on which GPU and oracle disagree on the status -- objective and feasibility of both answers."""
import sys; sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
from oracle.oracle import Oracle
from helpers import referee
w = W.config(3); N = 65536
inp = W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.tick(1, w.dt)
nu = prob.get_output(); st, it = prob.get_status()
sub = np.unique(np.linspace(0, N - 1, 512).astype(int))
o = Oracle(w.model, w.tasks)
nur, str_, _ = o.tick_batch(inp.base_pos[sub].copy(), inp.base_quat[sub].copy(), inp.s[sub].copy(), [None if sp is None else sp[sub] for sp in inp.setpoints], nthreads=8)
ok = (st[sub] == 0) & (str_ == 0)
for k in np.flatnonzero(~ok):
    i = int(sub[k])
    H, g, Cm, l, u = o.assemble(inp.base_pos[i], inp.base_quat[i], inp.s[i], [None if sp is None else sp[i] for sp in inp.setpoints])
    feas, xe = referee(H, g, Cm, l, u)
    f = lambda x: 0.5 * x @ H @ x + g @ x
    viol = lambda x: max(np.maximum(Cm @ x - u, 0).max(), np.maximum(l - Cm @ x, 0).max())
    print(f"robot {i}: gpu status {st[i]} iters {it[i]} oracle status {str_[k]} feasible {feas}")
    if feas:
        print(f"   f(gpu) {f(nu[i]):.12f} viol {viol(nu[i]):.2e} | f(slsqp) {f(xe):.12f} viol {viol(xe):.2e} | f(oracle) {f(nur[k]):.12f} viol {viol(nur[k]):.2e} | |gpu-slsqp| {np.abs(nu[i]-xe).max():.2e} |gpu-oracle| {np.abs(nu[i]-nur[k]).max():.2e}")
