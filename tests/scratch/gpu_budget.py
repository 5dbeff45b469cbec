import sys, time, os; sys.path.insert(0,'/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
w=W.config(3); N=65536; inp=W.synth_inputs(w,N,3)
model=capi.Model(w.model)
for rounds in (8, 3):
    prob=capi.Problem(model,w.tasks,N); prob.set_inputs(inp); prob.set_solver(max_rounds=rounds)
    st=torch.cuda.ExternalStream(prob.stream)
    ts=[]
    for rep in range(6):
        prob.set_state(inp.base_pos,inp.base_quat,inp.s); prob.sync()
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        e0.record(st); prob.tick(1,w.dt); e1.record(st); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    s_,it=prob.get_status()
    print(f"BLFIK_MAX_ITERS={os.environ.get('BLFIK_MAX_ITERS')} rounds={rounds}: {min(ts):.3f} ms unsolved={np.sum(s_!=0)} iters mean {it.mean():.2f} p99.9 {np.percentile(it,99.9):.0f} max {it.max()}")
    prob.close()
