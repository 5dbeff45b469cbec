import sys, time; sys.path.insert(0,'/root/repo')
import numpy as np
import blf_b200
from blf_b200 import capi, workloads as W
from oracle.oracle import Oracle
N=int(sys.argv[1]) if len(sys.argv)>1 else 256
for cid in (3,1,4):
    w=W.config(cid); inp=W.synth_inputs(w,N,cid)
    model=capi.Model(w.model); prob=capi.Problem(model,w.tasks,N)
    print(f"== C{cid}: n={w.model.n} me={prob.me}")
    prob.set_inputs(inp)
    o=Oracle(w.model,w.tasks)
    # kinematics
    frames=[t.frame for t in w.tasks if t.frame>=0]
    k=prob.kinematics(frames)
    eJ=0; eP=0; eC=0
    for i in range(min(N,32)):
        R,P=o.fk(inp.base_pos[i],inp.base_quat[i],inp.s[i])
        for fi,f in enumerate(frames):
            Rf,pf=o.frame_pose(R,P,f); J=o.frame_jacobian(R,P,f)
            eP=max(eP,np.abs(k['frame_pos'][i,fi]-pf).max(),np.abs(k['frame_rot'][i,fi]-Rf).max())
            eJ=max(eJ,np.abs(k['frame_jac'][i,fi]-J).max())
        c,Jc=o.com(R,P)
        eC=max(eC,np.abs(k['com'][i]-c).max(),np.abs(k['com_jac'][i]-Jc).max())
    print(f" kin err pose {eP:.2e} jac {eJ:.2e} com {eC:.2e}")
    prob.debug_enable(True)
    t0=time.time(); prob.tick(1,w.dt); prob.sync(); print(' tick wall', time.time()-t0)
    nu=prob.get_output(); st,it=prob.get_status()
    H=prob.debug_get(capi.DEBUG_H); g=prob.debug_get(capi.DEBUG_G); Ce=prob.debug_get(capi.DEBUG_CEQ); be=prob.debug_get(capi.DEBUG_BEQ)
    eH=0
    for i in range(min(N,16)):
        Ho,go,Co,lo,uo=o.assemble(inp.base_pos[i],inp.base_quat[i],inp.s[i],[None if sp is None else sp[i] for sp in inp.setpoints])
        eH=max(eH,np.abs(H[i]-Ho).max(),np.abs(g[i]-go).max(),np.abs(Ce[i]-Co[:prob.me]).max(),np.abs(be[i]-uo[:prob.me]).max())
    print(f" assemble err {eH:.2e}")
    nur,str_,itr=o.tick_batch(inp.base_pos.copy(),inp.base_quat.copy(),inp.s.copy(),inp.setpoints,nthreads=8)
    ok=(st==0)&(str_==0)
    print(' status gpu',np.bincount(st,minlength=3),'oracle',np.bincount(str_,minlength=3),'iters',it.min(),np.median(it),it.max())
    d=np.abs(nu-nur).max(1)
    print(f" max|dnu| over ok: {d[ok].max():.3e}  median {np.median(d[ok]):.3e} unsolved={prob.count_unsolved()}")
    # timing
    for rep in range(3):
        t0=time.time(); prob.tick(1,w.dt); prob.sync(); dt=time.time()-t0
    print(f" tick time {dt*1e3:.3f} ms -> {N/dt:.3e} ticks/s")
    # integrate parity
    prob.tick(1,w.dt,integrate=True); p,q,s=prob.get_state()
    e=0
    for i in range(min(N,32)):
        pr,qr,sr=o.integrate(inp.base_pos[i],inp.base_quat[i],inp.s[i],nu[i],w.dt)
        e=max(e,np.abs(p[i]-pr).max(),np.abs(s[i]-sr).max(),min(np.abs(q[i]-qr).max(),np.abs(q[i]+qr).max()))
    print(f" integrate err {e:.2e}")
    prob.close()
