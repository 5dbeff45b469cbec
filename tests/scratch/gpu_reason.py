import os, sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
cid, N = 3, 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model); prob = capi.Problem(model, w.tasks, N)
prob.set_inputs(inp); prob.tick(1, w.dt); prob.sync()
st, raw = prob.get_status()
it, reason, last_as, last_a = raw & 0xfff, (raw >> 12) & 0xf, (raw >> 16) & 0xf, (raw >> 20) & 0x3f
hard = it >= 16
print("hard robots", hard.sum())
print("reason hist (0 none/exhausted, 1 cap, 2 singular, 3 noconv):", np.bincount(reason[hard], minlength=4))
print("last_as hist among hard:", np.bincount(last_as[hard]))
print("last active-set size among hard:", np.bincount(last_a[hard]))
print("active-set size among easy:", np.bincount(last_a[~hard]))
print("iters of reason-2:", np.bincount(np.minimum(it[hard & (reason == 2)], 60)))
print("iters of reason-0:", np.bincount(np.minimum(it[hard & (reason == 0)], 60)))
