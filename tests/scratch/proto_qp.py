import sys; sys.path.insert(0,'/root/repo')
import numpy as np
import blf_b200
from blf_b200 import workloads as W
from oracle.oracle import Oracle

def setup(H, C, sigma, rho_vec):
    nv=H.shape[0]
    P = H + sigma*np.eye(nv) + np.diag(rho_vec)
    L = np.linalg.cholesky(P)
    Linv = np.linalg.inv(L)
    Pinv = Linv.T@Linv
    if C.shape[0]==0:
        return Pinv, np.zeros((nv,0)), Pinv
    Y = Linv@C.T
    S = Y.T@Y
    Ls = np.linalg.cholesky(S)
    V = np.linalg.solve(Ls, Y.T).T
    Wm = Linv.T@V
    N = Pinv - Wm@Wm.T
    Tm = Wm@np.linalg.inv(Ls)   # = Pinv C^T S^-1 ; x_b = Tm b
    return N, Tm, Pinv

def solve(H,g,C,b,lo,hi,sigma=1e-3,rho=1.0,alpha=1.6,iters=200,eps=1e-7,x0=None,z0=None,y0=None,check=1):
    nv=H.shape[0]; n=nv-6
    rho_vec=np.zeros(nv); rho_vec[6:]=rho
    N,Tm,_=setup(H,C,sigma,rho_vec)
    xb=Tm@b
    x=np.zeros(nv) if x0 is None else x0.copy(); z=np.zeros(n) if z0 is None else z0.copy(); y=np.zeros(n) if y0 is None else y0.copy()
    for k in range(1,iters+1):
        r = sigma*x - g
        r[6:] += rho*z - y
        xt = N@r + xb
        xn = alpha*xt+(1-alpha)*x
        zr = alpha*xt[6:]+(1-alpha)*z
        zn = np.clip(zr + y/rho, lo, hi)
        y = y + rho*(zr-zn)
        rd = max(rho*np.abs(zn-z).max(), sigma*np.abs(xn-x).max())
        x=xn; z=zn
        rp = np.abs(x[6:]-z).max()
        if k%check==0 and rp<eps and rd<eps: break
    return x,z,y,k

def polish(H,g,C,b,lo,hi,x,z,y,sigma=1e-3,nit=4):
    nv=H.shape[0]
    up = (hi - z) < y; low = (z - lo) < -y
    act = np.zeros(nv,bool); act[6:]= up|low
    bd = np.zeros(nv); bd[6:] = np.where(up,hi,np.where(low,lo,0))
    Hm=H.copy(); Hm[act,:]=0; Hm[:,act]=0; Hm[act,act]=1
    gm = g + H[:,act]@bd[act]; gm[act] = -bd[act]
    Cm = C.copy(); Cm[:,act]=0
    bm = b - C[:,act]@bd[act]
    N,Tm,_=setup(Hm,Cm,sigma,np.zeros(nv))
    xb=Tm@bm
    xp = x.copy(); xp[act]=bd[act]
    for k in range(nit):
        xp = N@(sigma*xp-gm)+xb
    # multipliers: lam from free stationarity (least squares here for the prototype)
    res = H@xp+g
    free=~act
    lam = np.linalg.lstsq(C[:,free].T, -res[free], rcond=None)[0]
    yfull = -(res + C.T@lam)
    ok = np.all(xp[6:]<=hi+1e-9) and np.all(xp[6:]>=lo-1e-9) and np.all(yfull[6:][up]>=-1e-9) and np.all(yfull[6:][low]<=1e-9)
    return xp, ok, np.abs(yfull[free]).max()

if __name__=="__main__":
    cid=int(sys.argv[1]) if len(sys.argv)>1 else 3
    NI=int(sys.argv[2]) if len(sys.argv)>2 else 128
    w=W.config(cid); inp=W.synth_inputs(w,NI,cid); o=Oracle(w.model,w.tasks)
    nu,st,it=o.tick_batch(inp.base_pos,inp.base_quat,inp.s,inp.setpoints,nthreads=8)
    n=w.model.n
    for rho in (0.3,1.0,3.0):
      for K0 in (5,10,20):
        its=[];errs=[];oks=0;errp=[]
        for i in range(NI):
            H,g,C,l,u=o.assemble(inp.base_pos[i],inp.base_quat[i],inp.s[i],[None if sp is None else sp[i] for sp in inp.setpoints])
            me=15; Ce=C[:me]; b=u[:me]
            if C.shape[0]>me:
                hi=np.minimum(u[me:me+n],u[me+2*n:me+3*n]); lo=np.maximum(-u[me+n:me+2*n],-u[me+3*n:me+4*n])
            else:
                hi=np.full(n,1e30); lo=-hi
            x,z,y,k=solve(H,g,Ce,b,lo,hi,rho=rho,iters=K0,eps=1e-9)
            xp,ok,rf=polish(H,g,Ce,b,lo,hi,x,z,y)
            oks+=ok; errs.append(np.abs(x-nu[i]).max()); errp.append(np.abs(xp-nu[i]).max() if ok else np.nan)
        print(f"rho={rho} K0={K0}: admm err max {max(errs):.2e} med {np.median(errs):.2e}; polish ok {oks}/{NI} err max {np.nanmax(errp):.2e}")
