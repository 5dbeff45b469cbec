"""How much of the C3 kernel time is the straggler tail?  Time the real batch, then a batch where the instances that
needed more than the median work are replaced by copies of easy ones (same N)."""
import os, sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
cid, N = 3, 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model); prob = capi.Problem(model, w.tasks, N)
stream = torch.cuda.ExternalStream(prob.stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
def timeit(inp, label):
    prob.set_inputs(inp)
    ts = []
    for rep in range(10):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
        flush.fill_(rep); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    st, it = prob.get_status()
    print(f"{label}: median {np.median(ts[2:]):.3f} ms; iters mean {it.mean():.2f} hist {np.bincount(np.minimum(it, 40))[:41].tolist()} >40: {(it > 40).sum()} unsolved {(st != 0).sum()}")
    return st, it
st, it = timeit(inp, "real batch")
import copy
for thr in (0, 1, 2, 4, 8, 16):
    easy = np.flatnonzero(it <= thr)
    if len(easy) == 0: continue
    idx = np.arange(N)
    hard = np.flatnonzero(it > thr)
    idx[hard] = easy[np.arange(len(hard)) % len(easy)]
    inp2 = copy.copy(inp)
    inp2.base_pos, inp2.base_quat, inp2.s = inp.base_pos[idx], inp.base_quat[idx], inp.s[idx]
    inp2.setpoints = [None if sp is None else sp[idx] for sp in inp.setpoints]
    timeit(inp2, f"instances with iters > {thr} replaced ({len(hard)})")
