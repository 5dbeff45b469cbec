import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
w=W.config(3); inpall=W.synth_inputs(w,65536,3)
model=capi.Model(w.model)
for N in (1776, 3552, 4096, 8192, 16384, 32768, 65536):
    inp=inpall.slice(0,N)
    prob=capi.Problem(model,w.tasks,N); prob.set_inputs(inp)
    st=torch.cuda.ExternalStream(prob.stream)
    ts=[]
    for rep in range(6):
        prob.set_state(inp.base_pos,inp.base_quat,inp.s); prob.sync()
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        e0.record(st); prob.tick(1,w.dt); e1.record(st); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    print(f"N={N:6d}: {min(ts):.3f} ms  {N/min(ts)/1e3:.1f} Mticks/s")
    prob.close()
