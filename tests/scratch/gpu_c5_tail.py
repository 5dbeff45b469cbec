"""C5 closed loop (bench amplitudes): per-tick kernel times and the per-tick iteration distribution, N robots on one GPU."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W, tasks as T
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
w = W.config(5)
inp = W.synth_inputs(w, N, 5)
rng = np.random.Generator(np.random.Philox(key=W.SEED + 50))
phases = rng.uniform(0.0, 2.0 * np.pi, size=(N, len(w.tasks)))
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp)
amp = 0.02
for t, task in enumerate(w.tasks):
    if task.type == T.TASK_SE3:
        prob.set_setpoint_trajectory(t, [amp] * 3 + [2 * amp] * 3, 2.0, np.ascontiguousarray(phases[:, t]))
prob.kernel_timing(True)
prob.tick(50, w.dt, integrate=True, warm_start=True); prob.sync()
rows = []
for k in range(40):
    prob.tick(1, w.dt, integrate=True, warm_start=True); prob.sync()
    st, it = prob.get_status()
    rows.append((prob.kernel_times(), it.mean(), it.max(), (it > 16).sum(), (st != 0).sum()))
ks = np.array([r[0] for r in rows])
print(f"N={N}: per tick assemble {np.median(ks[:, 0]):.3f} ms solve median {np.median(ks[:, 1]):.3f} min {ks[:, 1].min():.3f} max {ks[:, 1].max():.3f} ms; "
      f"iters mean {np.mean([r[1] for r in rows]):.2f} max per tick median {np.median([r[2] for r in rows]):.0f}; robots > 16 iters per tick {np.mean([r[3] for r in rows]):.1f}; unsolved per tick {np.mean([r[4] for r in rows]):.1f}")
import torch
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
prob.kernel_timing(False)
stream = torch.cuda.ExternalStream(prob.stream)
e0.record(stream); prob.tick(200, w.dt, integrate=True, warm_start=True); e1.record(stream); prob.sync(); torch.cuda.synchronize()
print(f"   200-tick launch: {e0.elapsed_time(e1) / 200:.3f} ms per tick = {N * 200 / e0.elapsed_time(e1) / 1e3:.1f} M ticks/s")
