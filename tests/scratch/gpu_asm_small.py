"""A small batch (argv[1] robots, default 2,048) through the two-launch path a few times: the assemble kernel at its latency floor."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
w = W.config(3)
inp = W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp); prob.kernel_timing(True)
for _ in range(5):
    prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.tick(1, w.dt); prob.sync()
print(N, np.round(prob.kernel_times(), 4).tolist())
