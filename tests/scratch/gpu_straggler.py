"""How long does a robot that exhausts the iteration budget take on its own?  C3: the robots with >= 200 mat-vecs as a batch."""
import sys, copy; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3); N = 65536
inp = W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.tick(1, w.dt)
st, it = prob.get_status()
for lo_, hi_ in ((200, 10**9), (100, 200), (48, 100), (17, 48), (0, 7)):
    idx = np.flatnonzero((it >= lo_) & (it < hi_))[:64]
    if len(idx) == 0: continue
    inp2 = copy.copy(inp)
    inp2.base_pos, inp2.base_quat, inp2.s = inp.base_pos[idx], inp.base_quat[idx], inp.s[idx]
    inp2.setpoints = [None if sp is None else sp[idx] for sp in inp.setpoints]
    p2 = capi.Problem(capi.Model(w.model), w.tasks, len(idx)); p2.set_inputs(inp2); p2.kernel_timing(True)
    ks = []
    for rep in range(6):
        p2.set_state(inp2.base_pos, inp2.base_quat, inp2.s); p2.tick(1, w.dt); p2.sync(); ks.append(p2.kernel_times())
    st2, it2 = p2.get_status()
    k = np.median(np.array(ks[2:]), axis=0)
    print(f"{len(idx)} robots with {lo_} <= mat-vecs < {hi_}: assemble {k[0]*1e3:.0f} us solve {k[1]*1e3:.0f} us; iters {it2.min()}..{it2.max()} status {np.bincount(st2).tolist()}")
