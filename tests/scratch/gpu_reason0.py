import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3); N = 65536
inp = W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.tick(1, w.dt)
st, r = prob.get_status()
code = r % 10; asit = (r // 10) % 100; a = r // 1000
print("reason histogram (0 solved in the first attempt, 1 too many active, 2 singular active-set matrix, 3 polish not converged, 4 refinements exhausted):", np.bincount(code, minlength=5).tolist())
for c in (2, 3):
    m = code == c
    print(f" reason {c}: at refinement {np.bincount(asit[m], minlength=13).tolist()}, active joints {np.bincount(a[m]).tolist()}")
