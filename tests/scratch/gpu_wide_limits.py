"""iCub-32 (38 velocity variables) with C2's tasks plus joint limits and velocity limits: the factored-form kernel with the
pivoting active set (path 0) against the generic kernel (path 1), and a parity spot check against the oracle."""
import sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, tasks as T, workloads as W
from oracle.oracle import Oracle
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
vmax = float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
w = W.config(2)
m, n = w.model, w.model.n
tasks = list(w.tasks) + [T.joint_velocity_limits_task(m, "vlim", np.full(n, vmax), np.full(n, -vmax)),
                         T.joint_limits_task(m, "plim", np.full(n, 0.5), w.dt, use_model_limits=True)]
T.validate_tasks(m, tasks)
inp = W.synth_inputs(w, N, 2)
inp.setpoints = list(inp.setpoints) + [None, None]
model = capi.Model(m)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
out = {}
for path in (0, 1):
    prob = capi.Problem(model, tasks, N)
    prob.set_kernel_path(path)
    prob.set_inputs(inp)
    stream = torch.cuda.ExternalStream(prob.stream)
    ts = []
    for rep in range(6 if path == 0 else 3):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
        flush.fill_(rep); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    nu = prob.get_output(); st, it = prob.get_status()
    out[path] = (nu, st)
    t = np.median(ts[1:])
    nact = (np.abs(np.abs(nu[:, 6:]) - vmax) < 1e-9).sum(axis=1)
    print(f"path {path} kernel {prob.kernel_path}: {t:.3f} ms {N / t / 1e3:.2f} Mticks/s unsolved {int((st != 0).sum())} (status 1: {int((st == 1).sum())}, 2: {int((st == 2).sum())}) "
          f"steps median {np.median(it):.0f} max {it.max()}  joints at the velocity limit: mean {nact.mean():.2f} max {nact.max()}")
    prob.close()
both = (out[0][1] == 0) & (out[1][1] == 0)
print("max |nu(path 0) - nu(path 1)| where both solved:", np.abs(out[0][0] - out[1][0])[both].max(), " solved by both", both.sum())
o = Oracle(m, tasks)
sub = slice(0, min(N, 512))
nur, str_, _ = o.tick_batch(inp.base_pos[sub].copy(), inp.base_quat[sub].copy(), inp.s[sub].copy(), [None if sp is None else sp[sub] for sp in inp.setpoints], nthreads=16)
ok = (out[0][1][sub] == 0) & (str_ == 0)
print("max |nu(path 0) - oracle|:", np.abs(out[0][0][sub] - nur)[ok].max(), " compared", ok.sum(), "oracle unsolved", int((str_ != 0).sum()))
