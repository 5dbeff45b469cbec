"""Can the slow robots of a cold C3 tick be told in advance from the joint box (the velocity bounds the limit tasks produce)?"""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3); N = 65536
inp = W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.tick(1, w.dt)
st, it = prob.get_status()
m = w.model
hi = np.minimum(0.5 * (m.q_upper - inp.s) / w.dt, 2.0)
lo = np.maximum(0.5 * (m.q_lower - inp.s) / w.dt, -2.0)
width = hi - lo
slow = it > 16
vslow = it >= 48
for name, score in (("joints with box narrower than 1 rad/s", (width < 1.0).sum(1)), ("narrower than 2", (width < 2.0).sum(1)), ("narrower than 3.9", (width < 3.9).sum(1)),
                    ("joints with a bound within 0.05 of zero", ((np.abs(lo) < 0.05) | (np.abs(hi) < 0.05)).sum(1)),
                    ("min width (negated)", -width.min(1)), ("sum of 1/width", (1.0 / np.maximum(width, 1e-3)).sum(1))):
    order = np.argsort(-score, kind="stable")
    for frac in (0.02, 0.05, 0.10, 0.2):
        top = order[: int(frac * N)]
        print(f"{name}: top {100*frac:.0f} % holds {100 * slow[top].sum() / slow.sum():.0f} % of the {slow.sum()} robots > 16 mat-vecs, {100 * vslow[top].sum() / vslow.sum():.0f} % of the {vslow.sum()} >= 48")
print("score histogram (joints with a bound within 0.05 of zero):", np.bincount(((np.abs(lo) < 0.05) | (np.abs(hi) < 0.05)).sum(1))[:8], "mean iters by score:", [round(it[(((np.abs(lo) < 0.05) | (np.abs(hi) < 0.05)).sum(1)) == k].mean(), 1) for k in range(5)])
