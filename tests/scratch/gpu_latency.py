"""Latency of one control tick for small batches through the C ABI (host buffers in, host buffers out)."""
import sys, time; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3)
for N in (1, 16, 256, 2048):
    inp = W.synth_inputs(w, N, 3)
    prob = capi.Problem(capi.Model(w.model), w.tasks, N)
    prob.set_inputs(inp)
    nu = np.empty((N, w.model.n + 6)); st = np.empty(N, dtype=np.int32)
    rec = prob.pack_inputs(inp)
    for mode in ("separate", "packed"):
        ts = []
        for rep in range(300):
            t0 = time.perf_counter()
            if mode == "separate":
                prob.set_state(inp.base_pos, inp.base_quat, inp.s)
                for t, sp in enumerate(inp.setpoints):
                    if sp is not None:
                        prob.set_setpoint(t, sp)
                prob.tick(1, w.dt, integrate=True, warm_start=True)
                prob.get_output(nu)
            else:
                prob.tick_packed(rec, nu, st, dt=w.dt, integrate=True, warm_start=True, nchunks=1)
            ts.append(time.perf_counter() - t0)
        ts = np.array(ts[50:]) * 1e6
        print(f"N={N:5d} {mode:9s}: median {np.median(ts):7.1f} us  p99 {np.percentile(ts, 99):7.1f} us per tick (host to host)")
    prob.close()
