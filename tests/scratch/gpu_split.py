"""Timing of the two-launch path against the fused kernel: per-kernel CUDA-event times of one cold C-config tick.
usage: gpu_split.py [cid] [N]"""
import os, sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
from oracle.oracle import Oracle
cid = int(sys.argv[1]) if len(sys.argv) > 1 else 3
N = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
w = W.config(cid); inp = W.synth_inputs(w, N, cid)
model = capi.Model(w.model)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
res = {}
for path in (0, 2):
    prob = capi.Problem(model, w.tasks, N); prob.set_kernel_path(path)
    prob.set_inputs(inp)
    stream = torch.cuda.ExternalStream(prob.stream)
    ts = []
    for rep in range(10):
        prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
        flush.fill_(rep); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    nu = prob.get_output(); st, it = prob.get_status()
    res[path] = (nu, st)
    ts = np.array(ts[2:])
    print(f"path {path} kernel {prob.kernel_path}: C{cid} N={N} min {ts.min():.3f} ms median {np.median(ts):.3f} ms {N / np.median(ts) / 1e3:.2f} Mticks/s unsolved {int(np.sum(st != 0))} matvecs {it.mean():.2f}")
    prob.close()
both = (res[0][1] == 0) & (res[2][1] == 0)
print("status differs on", int((res[0][1] != res[2][1]).sum()), "robots; max |nu_split - nu_fused| where both solved:", np.abs(res[0][0] - res[2][0])[both].max())
o = Oracle(w.model, w.tasks)
sub = np.unique(np.linspace(0, N - 1, 512).astype(int))
nur, str_, _ = o.tick_batch(inp.base_pos[sub].copy(), inp.base_quat[sub].copy(), inp.s[sub].copy(), [None if sp is None else sp[sub] for sp in inp.setpoints], nthreads=16)
ok = (res[0][1][sub] == 0) & (str_ == 0)
print("split vs oracle:", np.abs(res[0][0][sub] - nur)[ok].max(), "on", int(ok.sum()))
