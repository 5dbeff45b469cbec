"""The infeasibility certificate of the ADMM fallback (BLFIK_EPS_INFEASIBLE): on C3 / survey inputs, which robots does it stop,
after how many mat-vecs, and does the LP referee agree that they have no feasible point?"""
import os, sys; sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
from oracle.oracle import Oracle
from helpers import referee
survey = len(sys.argv) > 1 and sys.argv[1] == "survey"
w = W.config(3); N = 65536
inp = W.synth_inputs_survey(w, N, 3) if survey else W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.kernel_timing(True); prob.tick(1, w.dt); prob.sync()
st, it = prob.get_status()
print(f"eps {os.environ.get('BLFIK_EPS_INFEASIBLE', '0')} survey {survey}: status {np.bincount(st, minlength=3).tolist()} kernels {np.round(prob.kernel_times(), 3).tolist()} iters of status-2 robots: {np.sort(it[st == 2])[:30].tolist()}")
o = Oracle(w.model, w.tasks)
wrong = 0
idx = np.flatnonzero(st == 2)[:200]
for i in idx:
    H, g, Cm, l, u = o.assemble(inp.base_pos[i], inp.base_quat[i], inp.s[i], [None if sp is None else sp[i] for sp in inp.setpoints])
    feas, _ = referee(H, g, Cm, l, u)
    wrong += bool(feas)
print(f"   LP referee on {len(idx)} robots declared infeasible: {wrong} have a feasible point")
one = np.flatnonzero(st == 1)[:60]
nf = 0
for i in one:
    H, g, Cm, l, u = o.assemble(inp.base_pos[i], inp.base_quat[i], inp.s[i], [None if sp is None else sp[i] for sp in inp.setpoints])
    feas, _ = referee(H, g, Cm, l, u)
    nf += not feas
print(f"   of {len(one)} robots left at status 1, {nf} are infeasible by the LP")
