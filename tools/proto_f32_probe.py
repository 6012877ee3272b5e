"""How accurate can a complex64 / float32 "fast mode" be?  (north_star: 1e-3 absolute.)  NumPy / LAPACK probe, CPU only.

The eigen stage (E^-1, B = Kx E^-1 Kx - I, generalised symmetric eigenproblem B w = lambda A w) is done in float32 with LAPACK
(ssygvd: the best a float32 kernel could hope for), the layer solve stays in float64, and the three transmitted / reflected
orders are compared with the all-float64 prototype on random, 8-pixel-block and 16-pixel-block patterns (order 40, 1100 nm,
60 deg).  The absolute eigenvalue error of a float32 tridiagonalisation is eps32 * |C| ~ 6e-8 * 1e3, i.e. a relative error
of 1e-5 ... 1e-4 in the propagation constants of the few propagating / slowly decaying modes that carry the efficiency.
Result (profiles/r02_f32_probe.txt): the error already touches 1e-3 before any float32 rounding in the layer solve, so a
float32 path cannot promise the 1e-3 contract; pirl_cfg.precision = 1 is therefore rejected instead of shipping a mode that
fails its stated tolerance on a few structures per thousand.
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tools'))
import numpy as np, scipy.linalg as sla
import proto_tail2 as T2, proto_kernels as PK
rng=np.random.default_rng(3)
def f32_modes(s, angle=60., order=40, n_ridge=3.551726470588235):
    N=order; ff=2*N+1
    u=(np.asarray(s,float)+1)/2
    tw,D=PK.host_tables(u.size,N); c=PK.coeffs(u,N,tw,D); Tm=PK.basis_toeplitz(c,N)
    er,eg=n_ridge**2,1.0
    E=(eg*np.eye(ff)+(er-eg)*Tm).astype(np.float32); A=((1/eg)*np.eye(ff)+(1/er-1/eg)*Tm).astype(np.float32)
    sa=np.sin(angle/180*np.pi)
    kap=np.zeros(ff); sk=np.zeros(ff); perm=np.arange(ff)
    for p in range(1,N+1):
        ic,is_=2*(p-1),2*(p-1)+1
        kap[ic]=kap[is_]=p*sa; sk[ic]=p*sa; sk[is_]=-p*sa; perm[ic],perm[is_]=is_,ic
    Einv=np.linalg.inv(E)            # float32 LAPACK
    sk32=sk.astype(np.float32)
    Bm=(sk32[:,None]*Einv[np.ix_(perm,perm)]*sk32[None,:]-np.eye(ff,dtype=np.float32)).astype(np.float32)
    lam,W=sla.eigh(Bm,A)             # float32 sygvd
    assert W.dtype==np.float32
    return lam.astype(float),W.astype(float),kap
errs=[]
for t in range(240):
    if t%3==0: s=rng.choice([-1.,1.],256)
    elif t%3==1: s=np.repeat(rng.choice([-1.,1.],32),8)
    else: s=np.repeat(rng.choice([-1.,1.],16),16)
    k0d=2*np.pi/1100*325
    lam,W,kap=T2.modes(s); der,det,_=T2.tail2(lam,W,kap,k0d,1.45,1.0,40)
    lam2,W2,kap=f32_modes(s); der2,det2,_=T2.tail2(lam2,W2,kap,k0d,1.45,1.0,40)
    errs.append(max(np.abs(det-det2).max(),np.abs(der-der2).max()))
errs=np.array(errs); print("eigen stage in f32 (tail in f64): max err %.2e median %.2e  >1e-3: %d  >1e-4: %d"%(errs.max(),np.median(errs),(errs>1e-3).sum(),(errs>1e-4).sum()))
