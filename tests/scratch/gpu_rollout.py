import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
w=W.config(5); N=int(sys.argv[1]) if len(sys.argv)>1 else 65536; T=int(sys.argv[2]) if len(sys.argv)>2 else 200
inp=W.synth_inputs(w,N,5)
model=capi.Model(w.model); prob=capi.Problem(model,w.tasks,N); prob.set_inputs(inp)
st=torch.cuda.ExternalStream(prob.stream)
for warm in (False, True):
    prob.set_inputs(inp); prob.sync()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record(st); prob.tick(T, w.dt, integrate=True, warm_start=warm); e1.record(st); e1.synchronize()
    ms=e0.elapsed_time(e1); s_,it=prob.get_status()
    print(f"rollout N={N} ticks={T} warm={warm}: {ms:.2f} ms -> {N*T/ms/1e3:.1f} Mticks/s; unsolved(last)={np.sum(s_!=0)} matvecs/tick {it.mean()/T:.2f}")
