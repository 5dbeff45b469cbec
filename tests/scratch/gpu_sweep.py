import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W, models as M, tasks as T
def run(name, w, N, cid):
    inp=W.synth_inputs(w,N,cid); model=capi.Model(w.model)
    for path in (0,1):
        prob=capi.Problem(model,w.tasks,N); prob.set_kernel_path(path); prob.set_inputs(inp)
        st=torch.cuda.ExternalStream(prob.stream); ts=[]
        for rep in range(4):
            prob.set_state(inp.base_pos,inp.base_quat,inp.s); prob.sync()
            e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
            e0.record(st); prob.tick(1,w.dt); e1.record(st); e1.synchronize(); ts.append(e0.elapsed_time(e1))
        s_,it=prob.get_status(); nu=prob.get_output()
        print(f"{name:10s} N={N:6d} kernel={prob.kernel_path} : {min(ts):8.3f} ms {N/min(ts)/1e3:7.2f} Mticks/s unsolved={np.sum(s_!=0)} matvecs {it.mean():.1f} max|nu| {np.abs(nu[s_==0]).max():.2f}")
        prob.close()
run('C3', W.config(3), 65536, 3)
run('C4', W.config(4), 65536, 4)
run('C2(icub32)', W.config(2), 4096, 2)
m=M.icub26()
tasks=[T.com_task(m,"c",10.0),T.se3_task(m,"r","r_sole",10.0,10.0),T.se3_task(m,"l","l_sole",10.0,10.0),T.joint_tracking_task(m,"j",[5.0]*m.n,priority=1,weight=[1.0]*m.n)]
run('icub26', W.Workload('icub26',m,tasks,0.01,''), 65536, 2)
