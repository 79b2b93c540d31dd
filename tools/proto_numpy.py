import sys, importlib, time, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/oracle')
import pfml_oracle as O
syn = importlib.import_module("jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200.synthetic")

def rinv(M, base=64):
    n=M.shape[0]
    if n<=base: return np.linalg.inv(M)
    h=n//2
    A,B,C,D=M[:h,:h],M[:h,h:],M[h:,:h],M[h:,h:]
    Ai=rinv(A,base); W=Ai@B; V=C@Ai; S=D-C@W; Si=rinv(S,base)
    X12=-W@Si; X21=-Si@V; X11=Ai-X12@V
    return np.block([[X11,X12],[X21,Si]])

def pad_inv(M, base=64):
    n=M.shape[0]; npad=1
    while npad<n: npad*=2
    if npad<base: npad=n
    Mp=np.eye(npad); Mp[:n,:n]=M
    return rinv(Mp,base)[:n,:n]

def sqrt_db(A, iters=12, log=False):
    # scaled product-form DB with Byers-Xu style scalar scaling from norm bounds
    n=A.shape[0]; I=np.eye(n)
    M=A.copy(); Y=A.copy()
    Mi=pad_inv(M)
    # bounds for eig of A^(1/2)
    b=np.sqrt(np.linalg.norm(A,'fro')); a=1/np.sqrt(np.linalg.norm(Mi,'fro'))
    for k in range(iters):
        if k==0: mu=1/np.sqrt(a*b)
        else: mu=np.sqrt(2*np.sqrt(a*b)/(a+b)) if k==1 else np.sqrt(2/(zeta+1/zeta))
        # track interval: after scaling X->mu X, eigen interval [mu a, mu b]; after newton map f(x)=(x+1/x)/2
        if k==0:
            lo,hi=mu*a,mu*b; zeta=None
            a,b=1.0,0.5*(hi+1/hi)  # min is 1 (at x=1), max at ends (symmetric since lo*hi=1)
        elif k==1:
            lo,hi=mu*a,mu*b
            zeta=mu
            a,b=1.0,0.5*(hi+1/hi)
        else:
            zeta=mu
        mu2=mu*mu
        Ynew=0.5*mu*(Y+ (Y@Mi)/mu2)
        Mnew=0.5*(I+0.5*(mu2*M+Mi/mu2))
        Y,M=Ynew,Mnew
        res=np.linalg.norm(M-I,'fro')
        if log: print(k,mu,res)
        if res<1e-14*np.sqrt(n): break
        Mi=pad_inv(M)
    return Y

def sqrt_db_det(A, iters=20, log=False):
    n=A.shape[0]; I=np.eye(n); M=A.copy(); Y=A.copy()
    for k in range(iters):
        Mi=pad_inv(M)
        s,ld=np.linalg.slogdet(M); mu=np.exp(-ld/(2*n))
        res0=np.linalg.norm(M-I,'fro')
        if res0<1e-2: mu=1.0
        mu2=mu*mu
        Y=0.5*mu*(Y+(Y@Mi)/mu2); M=0.5*(I+0.5*(mu2*M+Mi/mu2))
        res=np.linalg.norm(M-I,'fro')
        if log: print(k,mu,res)
        if res<1e-13: break
    return Y,k+1

def gpu_algo(panel,t,rff_x,vs, log=False):
    lb=panel.lb; sel=np.flatnonzero(panel.valid[t]); n=len(sel); p=rff_x.shape[2]
    L=panel.fct_load[t][sel]; F=panel.fct_cov[t]; iv=panel.ivol[t,sel]; lam=panel.lam[t,sel]
    w=panel.wealth[t]; rf=panel.rf[t]; gam=panel.gamma_rel; mu=panel.mu
    Sig=L@F@L.T+np.diag(iv)
    isl=1/np.sqrt(lam)
    x=(gam/w)*Sig*np.outer(isl,isl)
    sgr=1+Sig/(1+rf+mu)**2
    y=1+np.diag(sgr)
    I=np.eye(n)
    A=x@x+4*x
    sq,nit=sqrt_db_det(A,log=log)
    mt=0.5*(x+2*I-sq)
    for it in range(panel.iterations):
        M=x+np.diag(y)-mt*sgr
        mt=pad_inv(M)
    mt=0.5*(mt+mt.T)*0+mt
    # signals tilde
    St=[]; G=[]; S0=None
    for tau in range(lb+2):
        tt=t-tau; X=rff_x[tt,sel,:]; X=X-X.mean(0)
        s=np.concatenate([np.ones((n,1)),X],axis=1)
        s=s*np.sqrt(1/(s**2).sum(0)); s=(1/vs[tt,sel])[:,None]*s
        if tau==0: S0=s
        St.append(np.sqrt(lam)[:,None]*s)
        g=(1+panel.tr_ld0[tt,sel])/(1+panel.mu_ld0[tt,sel]); G.append(np.nan_to_num(g,nan=1.0))
    g0raw=(1+panel.tr_ld0[t,sel])/(1+panel.mu_ld0[t,sel])
    # chain
    Bt=I.copy(); CB=I.copy()
    Bt=mt*G[1][None,:]; CB=CB+Bt   # B1
    for th in range(1,10):   # B2..B10
        Bt=(Bt@mt)*G[th+1][None,:]; CB=CB+Bt
    B11=(Bt@mt)*G[11][None,:]
    Cl1=CB+B11
    C=I+mt@(G[0][:,None]*CB)
    # Horner  Z'_j = g_j*(S_{j+1} + mt Z'_{j+1}),  Z'_10 = g_10*S_11 ; unscaled Z_0
    Zp=G[10][:,None]*St[11]
    for j in range(9,0,-1):
        Zp=G[j][:,None]*(St[j+1]+mt@Zp)
    Z=St[1]+mt@Zp
    Nn=St[0]+mt@(G[0][:,None]*Z)
    Nl1=Z+B11@St[12]
    om=pad_inv(C)@Nn
    oml1=pad_inv(Cl1)@Nl1
    omchg=om-g0raw[:,None]*oml1
    # denom
    Lt=isl[:,None]*L
    Yk=Lt.T@om
    risk=gam*(om.T@((iv/lam)[:,None]*om)+Yk.T@(F@Yk))
    tc=w*(omchg.T@omchg)
    rt=om.T@(isl*panel.ret_ld1[t,sel])
    return dict(r_tilde=rt,risk=risk,tc=tc,denom=risk+tc,signal_t=S0, m=isl[:,None]*mt*np.sqrt(lam)[None,:], omega=isl[:,None]*om, nit=nit)

def rel(a,b): return np.max(np.abs(a-b))/np.max(np.abs(b))

def sqrt_db_norm(A, iters=20, log=False, kind='fro'):
    n=A.shape[0]; I=np.eye(n); M=A.copy(); Y=A.copy()
    for k in range(iters):
        Mi=pad_inv(M)
        res0=np.linalg.norm(M-I,'fro')
        if kind=='fro': mu=(np.linalg.norm(Mi,'fro')/np.linalg.norm(M,'fro'))**0.25
        else: mu=(np.linalg.norm(Mi,1)*np.linalg.norm(Mi,np.inf)/(np.linalg.norm(M,1)*np.linalg.norm(M,np.inf)))**0.125
        if res0<1e-2: mu=1.0
        mu2=mu*mu
        Y=0.5*mu*(Y+(Y@Mi)/mu2); M=0.5*(I+0.5*(mu2*M+Mi/mu2))
        res=np.linalg.norm(M-I,'fro')
        if log: print(k,mu,res)
        if res<1e-13: break
    return Y,k+1
