"""A/B timing of one library build: kernel time of the C3 tick at 65,536 robots (cold start, CUDA events via the C ABI's
own stream), plus a parity spot check against the oracle.  BLFIK_LIBRARY selects the build."""
import os, sys, time; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
from oracle.oracle import Oracle
cid = int(sys.argv[1]) if len(sys.argv) > 1 else 3
N = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model); prob = capi.Problem(model, w.tasks, N)
prob.set_inputs(inp)
stream = torch.cuda.ExternalStream(prob.stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
ts = []
for rep in range(12):
    prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
    flush.fill_(rep); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
nu = prob.get_output(); st, it = prob.get_status()
o = Oracle(w.model, w.tasks)
sub = slice(0, 512)
nur, str_, _ = o.tick_batch(inp.base_pos[sub].copy(), inp.base_quat[sub].copy(), inp.s[sub].copy(), [None if sp is None else sp[sub] for sp in inp.setpoints], nthreads=16)
ok = (st[sub] == 0) & (str_ == 0)
ts = np.array(ts[2:])
print(f"{os.environ.get('BLFIK_LIBRARY', 'default')}: C{cid} N={N} min {ts.min():.3f} ms median {np.median(ts):.3f} ms  {N / np.median(ts) / 1e3:.2f} Mticks/s unsolved {int(np.sum(st != 0))} err {np.abs(nu[sub] - nur)[ok].max():.2e}")
