import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3); N = 65536
inp = W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.tick(1, w.dt)
st, it = prob.get_status()
print("status", np.bincount(st))
edges = [0, 4, 6, 8, 12, 16, 24, 32, 48, 64, 100, 150, 200, 1000]
h, _ = np.histogram(it, bins=edges)
for a, b, c in zip(edges[:-1], edges[1:], h): print(f"iters [{a},{b}): {c}")
big = np.flatnonzero(it >= 48)
print("robots with >= 48 mat-vecs:", big.tolist()[:40], "status", st[big].tolist()[:40], "iters", it[big].tolist()[:40])
