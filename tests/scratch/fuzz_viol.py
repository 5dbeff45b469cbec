"""Which constraint does the certified point of robot INST violate?  python fuzz_viol.py <seed>"""
import sys, os; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, blf_b200
src=open('/root/repo/tests/test_gpu_parity.py').read()
i0=src.index("def test_random_task_mixes_vs_oracle"); body=src[i0:]
from blf_b200 import models as M, tasks as T, workloads as W, capi
from oracle.oracle import Oracle
pre=body[:body.index("    model, prob = make(m, tasks, N, int(os.environ")]
ns=dict(np=np,M=M,T=T,W=W,Oracle=Oracle,os=os)
exec(pre.replace("def test_random_task_mixes_vs_oracle(seed):","def build(seed):")+"    return m,tasks,o,sps,p0,q0,s0,kinds\n",ns)
seed=int(sys.argv[1])
m,tasks,o,sps,p0,q0,s0,kinds=ns['build'](seed)
N=len(p0)
for path in (0,1):
    prob=capi.Problem(capi.Model(m),tasks,N); prob.set_kernel_path(path)
    prob.set_state(p0,q0,s0)
    for t,sp in enumerate(sps):
        if sp is not None: prob.set_setpoint(t,sp)
    prob.debug_enable(True)
    prob.tick(1,0.01)
    st,it=prob.get_status(); nu=prob.get_output(); me=prob.me
    Ce,be=prob.debug_get(capi.DEBUG_CEQ),prob.debug_get(capi.DEBUG_BEQ)
    for i in (9,15):
        Ho,go,Co,l,u=o.assemble(p0[i],q0[i],s0[i],[None if sp is None else sp[i] for sp in sps])
        print("  robot",i,"|Ce-Co|",np.abs(Ce[i]-Co[:me]).max(),"|be-u|",np.abs(be[i]-u[:me]).max(),"GPU's own residual Ce nu - be",np.round(Ce[i]@nu[i]-be[i],6).tolist())
    print("seed",seed,"n",m.n,"path",prob.kernel_path,"me",me,[(t.name,t.priority,tuple(t.mask)) for t in tasks])
    print(" status",st.tolist())
    for i in np.flatnonzero(st==0):
        Ho,go,Co,l,u=o.assemble(p0[i],q0[i],s0[i],[None if sp is None else sp[i] for sp in sps])
        cx=Co@nu[i]; v=np.maximum(cx-u,l-cx); k=int(np.argmax(v))
        if v[k]>1e-6*(1+np.abs(cx).max()):
            sv=np.linalg.svd(Co[:me],compute_uv=False)
            lo_,hi_=l[me:],u[me:]; xj=(Co[me:]@nu[i])
            print("   joints at upper",int((np.abs(xj-hi_)<1e-9).sum()),"at lower",int((np.abs(xj-lo_)<1e-9).sum()),"of",len(xj),"max|nu|",np.abs(nu[i]).max(),"base",np.round(nu[i][:6],4).tolist(),"iters",int(it[i]))
            print("  robot",i,"violates row",k,"by",v[k],"rows 0..me-1 residual",np.round(Co[:me]@nu[i]-u[:me],6).tolist(),"sv",np.round(sv,8).tolist())
    prob.close()
