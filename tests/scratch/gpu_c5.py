"""C5 rollout as bench.py --config 5 sets it up: status histogram after 1,000 sticky ticks, per trajectory amplitude."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np, time
import blf_b200
from blf_b200 import capi, workloads as W, tasks as T
w = W.config(5); N = 8192
inp = W.synth_inputs(w, N, 5)
rng = np.random.Generator(np.random.Philox(key=W.SEED + 50))
phases = rng.uniform(0.0, 2.0 * np.pi, size=(N, len(w.tasks)))
model = capi.Model(w.model)
for amp, nt in ((0.05, 1000), (0.02, 1000), (0.0, 1000), (0.05, 100), (0.05, 10)):
    prob = capi.Problem(model, w.tasks, N); prob.set_inputs(inp)
    for t, task in enumerate(w.tasks):
        if task.type == T.TASK_SE3 and amp > 0:
            prob.set_setpoint_trajectory(t, [amp] * 3 + [2 * amp] * 3, 2.0, np.ascontiguousarray(phases[:, t]))
    t0 = time.time(); prob.tick(nt, w.dt, integrate=True, warm_start=True); prob.sync(); dt = time.time() - t0
    st, it = prob.get_status()
    print(f"amp {amp} ticks {nt}: status histogram {np.bincount(st, minlength=3)} matvecs/tick {it.mean() / nt:.2f} {N * nt / dt / 1e6:.1f} M ticks/s")
    prob.close()
