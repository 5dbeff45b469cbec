import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np, torch
import blf_b200
from blf_b200 import capi, workloads as W
from oracle.oracle import Oracle
cid=int(sys.argv[1]) if len(sys.argv)>1 else 3
N=int(sys.argv[2]) if len(sys.argv)>2 else 65536
w=W.config(cid); inp=W.synth_inputs(w,N,cid)
model=capi.Model(w.model); prob=capi.Problem(model,w.tasks,N)
o=Oracle(w.model,w.tasks)
sub=slice(0,1024)
nur,str_,_=o.tick_batch(inp.base_pos[sub].copy(),inp.base_quat[sub].copy(),inp.s[sub].copy(),[None if sp is None else sp[sub] for sp in inp.setpoints],nthreads=16)
frames=[t.frame for t in w.tasks if t.type==0 and t.priority==0]
prob.set_inputs(inp)
k=prob.kinematics(frames[:1],jac=True,com=False)
sp=inp.setpoints[1]; tk=w.tasks[1]
settings=[dict(), dict(sigma=1e-2), dict(sigma=1e-2,rho=0.5), dict(sigma=1e-2,admm_iters=6), dict(sigma=1e-2,admm_iters=4), dict(sigma=1e-2, rho=2.0), dict(sigma=1e-1)]
for path in (0,):
  for s in settings:
    prob.set_kernel_path(path); prob.set_solver(**s)
    prob.set_inputs(inp)
    ts=[]
    for rep in range(4):
        prob.set_state(inp.base_pos,inp.base_quat,inp.s); prob.sync()
        t0=time.perf_counter(); prob.tick(1,w.dt); prob.sync(); ts.append(time.perf_counter()-t0)
    nu=prob.get_output(); st,it=prob.get_status()
    ok=(st[sub]==0)&(str_==0)
    err=np.abs(nu[sub]-nur)[ok].max()
    v=np.einsum('nrc,nc->nr',k['frame_jac'][:,0],nu)
    b_lin=sp[:,7:10]+tk.kp_linear*(sp[:,:3]-k['frame_pos'][:,0])
    eq=np.abs(v[:,:3]-b_lin)[st==0].max()
    print(f"path={path} {s}: {min(ts)*1e3:.3f} ms  {N/min(ts)/1e6:.2f} Mticks/s unsolved={np.sum(st!=0)} iters mean {it.mean():.1f} p99 {np.percentile(it,99):.0f} max {it.max()}  err_vs_oracle {err:.2e} eq_res {eq:.2e}")
