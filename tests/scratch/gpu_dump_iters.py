"""Dump per-robot iteration counts and statuses of a cold C3 tick (for offline analysis of what predicts the slow robots)."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3); N = 65536
for seed in (3, 11):
    inp = W.synth_inputs(w, N, seed)
    prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.tick(1, w.dt)
    st, it = prob.get_status()
    np.savez_compressed(f'/root/repo/gpurun_out/iters_seed{seed}.npz', st=st.astype(np.int8), it=it.astype(np.int16))
    prob.close()
print("ok")
