import sys; sys.path.insert(0,'/root/repo')
import numpy as np
import blf_b200
from blf_b200 import workloads as W
from oracle.oracle import Oracle
from scratch.proto_qp import setup

def run(H,g,C,b,lo,hi,sigma,rho,alpha,K0,kpol,rounds=6):
    nv=H.shape[0]; n=nv-6
    rho_vec=np.zeros(nv); rho_vec[6:]=rho
    N,Tm,_=setup(H,C,sigma,rho_vec); xb=Tm@b
    x=np.zeros(nv); z=np.zeros(n); y=np.zeros(n)
    Q=np.full(nv,sigma); Q[6:]+=rho
    tot=0
    for rd in range(rounds):
        for k in range(K0):
            r = sigma*x - g; r[6:] += rho*z - y
            xt = N@r + xb
            x = alpha*xt+(1-alpha)*x
            zr = alpha*xt[6:]+(1-alpha)*z
            zn = np.clip(zr + y/rho, lo, hi)
            y = y + rho*(zr-zn); z=zn
        tot+=K0
        up = (hi - z) < y; low = (z - lo) < -y
        A = np.nonzero(up|low)[0]+6
        bd = np.where(up,hi,lo)[A-6]
        xp = x.copy(); xp[A]=bd
        if len(A): MA=np.linalg.inv(N[np.ix_(A,A)])
        for k in range(kpol):
            r = Q*xp - g
            x1 = N@r + xb
            if len(A):
                mu = MA@(x1[A]-bd); xn = x1 - N[:,A]@mu
            else: xn=x1; mu=np.zeros(0)
            dx=np.abs(xn-xp).max(); xp=xn
            tot+=1
        yA=np.zeros(nv); 
        if len(A): yA[A]=mu
        ok = np.all(xp[6:]<=hi+1e-9) and np.all(xp[6:]>=lo-1e-9) and np.all(yA[6:][up]>=-1e-9) and np.all(yA[6:][low]<=1e-9) and dx*Q.max()<1e-8
        if ok: return xp, rd+1, tot, len(A)
    return x, -1, tot, len(A)

cid=3; NI=int(sys.argv[1]) if len(sys.argv)>1 else 256
w=W.config(cid); inp=W.synth_inputs(w,NI,cid); o=Oracle(w.model,w.tasks)
nu,st,it=o.tick_batch(inp.base_pos,inp.base_quat,inp.s,inp.setpoints,nthreads=8)
n=w.model.n
probs=[]
for i in range(NI):
    H,g,C,l,u=o.assemble(inp.base_pos[i],inp.base_quat[i],inp.s[i],[None if sp is None else sp[i] for sp in inp.setpoints])
    me=15
    hi=np.minimum(u[me:me+n],u[me+2*n:me+3*n]); lo=np.maximum(-u[me+n:me+2*n],-u[me+3*n:me+4*n])
    probs.append((H,g,C[:me],u[:me],lo,hi))
for rho in (0.03,0.1,0.3):
  for K0 in (4,8):
    for kpol in (6,10):
        errs=[];rds=[];tots=[]
        for i,(H,g,Ce,b,lo,hi) in enumerate(probs):
            x,rd,tot,na=run(H,g,Ce,b,lo,hi,1e-3,rho,1.6,K0,kpol)
            errs.append(np.abs(x-nu[i]).max()); rds.append(rd); tots.append(tot)
        rds=np.array(rds)
        print(f"rho={rho} K0={K0} kpol={kpol}: fail={np.sum(rds<0)} rounds>1: {np.sum(rds>1)} mean matvecs {np.mean(tots):.1f} max err {max(errs):.2e}")
