"""Per-kernel time of the two-launch tick against the batch size (C3 tasks, cold, warm L2)."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3)
model = capi.Model(w.model)
for N in (1, 256, 2048, 8192, 16384, 32768, 65536):
    inp = W.synth_inputs(w, N, 3)
    prob = capi.Problem(model, w.tasks, N); prob.set_inputs(inp); prob.kernel_timing(True)
    ta, tb = [], []
    for rep in range(8):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.tick(1, w.dt)
        a, b = prob.kernel_times(); ta.append(a); tb.append(b)
    print(f"N={N:6d}: assemble {np.median(ta[2:]) * 1e3:7.1f} us  solve {np.median(tb[2:]) * 1e3:7.1f} us  -> {N / (np.median(ta[2:]) + np.median(tb[2:])) / 1e3:6.2f} M ticks/s")
    prob.close()
