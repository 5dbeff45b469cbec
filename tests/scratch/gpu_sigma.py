"""Mat-vec count, time and accuracy of the C3 tick against the proximal weight sigma."""
import os, sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
from oracle.oracle import Oracle
cid, N = 3, 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model)
o = Oracle(w.model, w.tasks)
sub = np.unique(np.linspace(0, N - 1, 512).astype(int))
nur, str_, _ = o.tick_batch(inp.base_pos[sub].copy(), inp.base_quat[sub].copy(), inp.s[sub].copy(), [None if sp is None else sp[sub] for sp in inp.setpoints], nthreads=16)
for sigma in (1e-2, 1e-3, 3e-4, 1e-4, 1e-5):
    prob = capi.Problem(model, w.tasks, N); prob.set_solver(sigma=sigma); prob.set_inputs(inp)
    stream = torch.cuda.ExternalStream(prob.stream)
    ts = []
    for rep in range(6):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    nu = prob.get_output(); st, it = prob.get_status()
    ok = (st[sub] == 0) & (str_ == 0)
    print(f"sigma {sigma:g}: {np.median(ts[1:]):.3f} ms matvecs {it.mean():.2f} unsolved {int((st != 0).sum())} (status 1: {int((st == 1).sum())}) err vs oracle {np.abs(nu[sub] - nur)[ok].max():.2e}")
    prob.close()
