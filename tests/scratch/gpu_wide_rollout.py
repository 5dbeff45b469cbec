"""iCub-32 with limits: a 100-tick closed-loop launch, cold against warm-started active sets."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, tasks as T, workloads as W
N, NT = int(sys.argv[1]) if len(sys.argv) > 1 else 16384, 100
w = W.config(2); m, n = w.model, w.model.n
tasks = list(w.tasks) + [T.joint_velocity_limits_task(m, "vlim", np.full(n, 3.0), np.full(n, -3.0)),
                         T.joint_limits_task(m, "plim", np.full(n, 0.5), w.dt, use_model_limits=True)]
inp = W.synth_inputs(w, N, 2, jr=0.05, pr=0.005)
inp.setpoints = list(inp.setpoints) + [None, None]
model = capi.Model(m)
for warm in (False, True):
    prob = capi.Problem(model, tasks, N); prob.set_inputs(inp)
    stream = torch.cuda.ExternalStream(prob.stream)
    prob.tick(1, w.dt, integrate=False); prob.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); prob.tick(NT, w.dt, integrate=True, warm_start=warm); e1.record(stream); prob.sync(); torch.cuda.synchronize()
    st, it = prob.get_status()
    ms = e0.elapsed_time(e1)
    print(f"path {prob.kernel_path} warm {warm}: {NT} ticks x {N} robots in {ms:.2f} ms = {N * NT / ms / 1e3:.2f} Mticks/s, unsolved at the end {int((st != 0).sum())}, steps per tick {it.mean() / NT:.1f}")
    prob.close()
