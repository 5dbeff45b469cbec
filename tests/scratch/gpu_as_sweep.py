"""C3, 65,536 robots: step time, iteration histogram and unsolved count for the active-set / budget settings given in the
environment (BLFIK_MAX_AS, BLFIK_MAX_ITERS)."""
import os, sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
w = W.config(3); N = 65536
inp = W.synth_inputs(w, N, 3)
prob = capi.Problem(capi.Model(w.model), w.tasks, N); prob.set_inputs(inp)
stream = torch.cuda.ExternalStream(prob.stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
prob.kernel_timing(True)
kw = {}
for a in sys.argv[1:]:
    k, v = a.split('=')
    kw[k] = int(v) if k in ('admm_iters', 'max_rounds', 'polish_iters', 'eq_iters') else float(v)
if kw: prob.set_solver(**kw)
ts, ks = [], []
for rep in range(10):
    prob.set_state(inp.base_pos, inp.base_quat, inp.s); prob.sync()
    flush.fill_(rep); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream); prob.tick(1, w.dt); e1.record(stream); prob.sync(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1)); ks.append(prob.kernel_times())
st, it = prob.get_status()
nu = prob.get_output()
ref = os.environ.get("REF_NU")
tag = f"{kw} MAX_AS={os.environ.get('BLFIK_MAX_AS','-')} MAX_ITERS={os.environ.get('BLFIK_MAX_ITERS','-')}"
extra = ""
if ref and os.path.exists(ref):
    r = np.load(ref); ok = (st == 0) & (r['st'] == 0)
    extra = f" |dnu| vs default {np.abs(nu - r['nu'])[ok].max():.2e} (solved by both {ok.sum()})"
elif ref:
    np.savez(ref, nu=nu, st=st)
k = np.median(np.array(ks[2:]), axis=0)
print(f"{tag}: step {np.median(ts[2:]):.3f} ms kernels {k.round(3).tolist()} iters mean {it.mean():.2f} >16: {(it > 16).sum()} >=48: {(it >= 48).sum()} unsolved {np.bincount(st).tolist()}{extra}")
