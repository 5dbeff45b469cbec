"""Cost of one robot per difficulty class: fill the batch with copies of robots that needed exactly k mat-vecs."""
import os, sys, copy; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
cid, N = 3, 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model); prob = capi.Problem(model, w.tasks, N)
stream = torch.cuda.ExternalStream(prob.stream)
def timeit(inp):
    prob.set_inputs(inp)
    ts = []
    for rep in range(8):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    st, it = prob.get_status()
    return np.median(ts[2:]), it
t_real, it = timeit(inp)
print(f"real batch {t_real:.3f} ms")
for k in (6, 10, 12, 14, 16, 18, 24):
    src = np.flatnonzero(it == k)
    if len(src) < 8: continue
    idx = src[np.arange(N) % len(src)]
    rng = np.random.default_rng(0); rng.shuffle(idx)
    inp2 = copy.copy(inp)
    inp2.base_pos, inp2.base_quat, inp2.s = inp.base_pos[idx], inp.base_quat[idx], inp.s[idx]
    inp2.setpoints = [None if sp is None else sp[idx] for sp in inp.setpoints]
    t, it2 = timeit(inp2)
    print(f"all robots of class {k:2d} mat-vecs ({len(src):5d} distinct): {t:.3f} ms  ({t / N * 2368 * 1e3:.1f} us per robot-warp)  check iters {np.unique(it2)[:4]}")
