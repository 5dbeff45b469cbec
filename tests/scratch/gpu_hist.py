import os, sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
cid, N = int(sys.argv[1]) if len(sys.argv) > 1 else 3, 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model); prob = capi.Problem(model, w.tasks, N)
prob.set_inputs(inp)
stream = torch.cuda.ExternalStream(prob.stream)
ts = []
for rep in range(8):
    prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
st, it = prob.get_status()
h = np.bincount(np.minimum(it, 60))
print({k: os.environ[k] for k in os.environ if k.startswith("BLFIK_")}, f"median {np.median(ts[2:]):.3f} ms mean iters {it.mean():.2f} unsolved {(st != 0).sum()}")
print(" hist:", {i: int(c) for i, c in enumerate(h) if c})
