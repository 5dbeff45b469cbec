import sys, os; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, blf_b200
src=open('/root/repo/tests/test_gpu_parity.py').read()
i0=src.index("def test_random_task_mixes_vs_oracle"); body=src[i0:]
from blf_b200 import models as M, tasks as T, workloads as W, capi
from oracle.oracle import Oracle
pre=body[:body.index("    model, prob = make(m, tasks, N, int(os.environ")]
ns=dict(np=np,M=M,T=T,W=W,Oracle=Oracle,os=os)
exec(pre.replace("def test_random_task_mixes_vs_oracle(seed):","def build(seed):")+"    return m,tasks,o,sps,p0,q0,s0,kinds\n",ns)
for seed in [int(a) for a in sys.argv[1:]]:
    m,tasks,o,sps,p0,q0,s0,kinds=ns['build'](seed)
    N=len(p0)
    prob=capi.Problem(capi.Model(m),tasks,N); prob.set_kernel_path(int(os.environ.get("KPATH","0")))
    prob.set_state(p0,q0,s0)
    for t,sp in enumerate(sps):
        if sp is not None: prob.set_setpoint(t,sp)
    prob.debug_enable(True); prob.tick(1,0.01)
    st,it=prob.get_status(); nu=prob.get_output()
    H,Ce=prob.debug_get(capi.DEBUG_H),prob.debug_get(capi.DEBUG_CEQ); me=prob.me
    nur,str_,_=o.tick_batch(p0.copy(),q0.copy(),s0.copy(),sps)
    print("seed",seed,"n",m.n,[(t.name,t.priority,t.mask) for t in tasks],"path",prob.kernel_path)
    print(" gpu status",np.bincount(st,minlength=3),"iters",it[:8],"oracle status",np.bincount(str_,minlength=3))
    for i in range(3):
        sv=np.linalg.svd(Ce[i],compute_uv=False); Z=np.linalg.svd(Ce[i])[2][me:].T
        print("  me",me,"svmin",sv.min(),"eigmin",np.linalg.eigvalsh(Z.T@H[i]@Z).min(),"|nu|max",np.abs(nu[i]).max(),"diff",np.abs(nu[i]-nur[i]).max())
    if os.environ.get("INST"):
        from helpers import solve_qp_active_set
        i=int(os.environ["INST"])
        Ho,go,Co,l,u=o.assemble(p0[i],q0[i],s0[i],[None if sp is None else sp[i] for sp in sps])
        x=solve_qp_active_set(Ho,go,Co,l,u)
        print(" inst",i,"status",st[i],"iters",it[i],"gpu-exact",np.abs(nu[i]-x).max(),"oracle-exact",np.abs(nur[i]-x).max())
        print(" active rows (exact):", np.flatnonzero((np.abs(Co@x-u)<1e-9)|(np.abs(Co@x-l)<1e-9))[me:])
        print(" cond(H)", np.linalg.cond(Ho), "x", np.round(x,3))
        lo,hi=prob.debug_get(capi.DEBUG_LO)[i],prob.debug_get(capi.DEBUG_HI)[i]
        print(" box viol gpu", np.maximum(nu[i,6:]-hi,0).max(), np.maximum(lo-nu[i,6:],0).max(), "eq res gpu", np.abs(Ce[i]@nu[i]-prob.debug_get(capi.DEBUG_BEQ)[i]).max())
        f=lambda z: 0.5*z@Ho@z+go@z
        print(" cost gpu-exact", f(nu[i])-f(x))
    if os.environ.get("VIOL"):
        for i in np.flatnonzero(st==0)[:3]:
            Ho,go,Co,l,u=o.assemble(p0[i],q0[i],s0[i],[None if sp is None else sp[i] for sp in sps])
            cx=Co@nu[i]; v=np.maximum(cx-u,l-cx)
            print(" inst",i,"iters",it[i],"worst rows",np.argsort(-v)[:4],np.sort(-v)[:4]*-1,"me",me,"svs",np.round(np.linalg.svd(Ce[i],compute_uv=False),6))
    if os.environ.get("EQ"):
        be=prob.debug_get(capi.DEBUG_BEQ); lo=prob.debug_get(capi.DEBUG_LO); hi=prob.debug_get(capi.DEBUG_HI)
        for i in [int(a) for a in os.environ["EQ"].split(",")]:
            print(" inst",i,"st",st[i],"it",it[i],"eq res gpu rows",np.round(Ce[i]@nu[i]-be[i],5),"box viol",np.maximum(nu[i,6:]-hi[i],0).max(),np.maximum(lo[i]-nu[i,6:],0).max(), "pinned", int(np.sum(lo[i]==hi[i])), "lo>hi", int(np.sum(lo[i]>hi[i])), "oracle st", str_[i])
    if os.environ.get("ACT"):
        from helpers import solve_qp_active_set
        lo=prob.debug_get(capi.DEBUG_LO); hi=prob.debug_get(capi.DEBUG_HI); be=prob.debug_get(capi.DEBUG_BEQ)
        for i in [int(a) for a in os.environ["ACT"].split(",")]:
            x=nu[i]; j=x[6:]
            print(" inst",i,"gpu at hi",np.flatnonzero(np.abs(j-hi[i])<1e-12),"at lo",np.flatnonzero(np.abs(j-lo[i])<1e-12))
            print("   lo",np.round(lo[i],3)); print("   hi",np.round(hi[i],3)); print("   nu joints",np.round(j,3))
            Ho,go,Co,l,u=o.assemble(p0[i],q0[i],s0[i],[None if sp is None else sp[i] for sp in sps])
            try:
                xe=solve_qp_active_set(Ho,go,Co,l,u); cx=Co@xe
                print("   exact feasible? viol",max((cx-u).max(),(l-cx).max()), "exact joints", np.round(xe[6:],3))
            except Exception as e: print("   exact solver failed:",e)
            # least-violation check: is there any point satisfying equalities and box? 
            import scipy.optimize as so
            A=Ce[i]; res=so.linprog(np.zeros(A.shape[1]),A_eq=A,b_eq=be[i],bounds=[(None,None)]*6+[(lo[i][k],hi[i][k]) for k in range(len(j))],method="highs")
            print("   LP feasibility:",res.status,res.message[:60])
