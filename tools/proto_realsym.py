"""Design prototype (NumPy) of the real-symmetric formulation the CUDA kernels implement.
Not product code and not the oracle: it exists to validate the math (basis change to cos/sin
harmonics, generalised symmetric eigenproblem, admittance-free tail) against the oracle on CPU
before the same steps are written as kernels.  See DESIGN.md section 3."""
import numpy as np, scipy.linalg as sla

def fourier_coeffs(u, N):
    """c_k, k=0..2N, of the R-fold up-sampled 0/1 pattern (closed form of convolution_matrix.py:101-112)."""
    n = u.size; R = 4*N+2; M = n*R
    k = np.arange(0, 2*N+1)
    p = np.arange(n)
    dft = (u[None, :] * np.exp(-2j*np.pi*np.outer(k, p)/n)).sum(1)/n
    with np.errstate(invalid='ignore', divide='ignore'):
        D = (1-np.exp(-2j*np.pi*k/n))/(R*(1-np.exp(-2j*np.pi*k/M)))
    D[k % n == 0] = 1.0
    return dft*D

def real_basis_toeplitz(c, N):
    """Q^H T(c) Q in the basis {1, sqrt2 cos(p.), sqrt2 sin(p.)}: index 0, 2p-1, 2p."""
    ff = 2*N+1
    S = np.zeros((ff, ff))
    cr, ci = c.real, c.imag
    S[0, 0] = cr[0]
    r2 = np.sqrt(2.0)
    for p in range(1, N+1):
        S[0, 2*p-1] = S[2*p-1, 0] = r2*cr[p]
        S[0, 2*p] = S[2*p, 0] = -r2*ci[p]
        for q in range(1, N+1):
            d = p-q
            crd = cr[abs(d)]; cid = ci[abs(d)]*(1 if d >= 0 else -1)
            S[2*p-1, 2*q-1] = crd + cr[p+q]
            S[2*p, 2*q] = crd - cr[p+q]
            S[2*p-1, 2*q] = -ci[p+q] + cid      # cos_p , sin_q
            S[2*p, 2*q-1] = -ci[p+q] - cid      # sin_p , cos_q
    return S

def unitary(N):
    ff = 2*N+1
    Q = np.zeros((ff, ff), complex)
    Q[N, 0] = 1
    for p in range(1, N+1):
        Q[N+p, 2*p-1] = 1/np.sqrt(2); Q[N-p, 2*p-1] = 1/np.sqrt(2)
        Q[N+p, 2*p] = 1/(1j*np.sqrt(2)); Q[N-p, 2*p] = -1/(1j*np.sqrt(2))
    return Q

def evaluate(s, wavelength=1100., angle=60., order=40, thickness=325., n_ridge=3.551726470588235,
             n_I=1.45, n_II=1.0, n_g=1.0):
    N = order; ff = 2*N+1
    u = (np.asarray(s, float)+1)/2
    c = fourier_coeffs(u, N)
    T = real_basis_toeplitz(c, N)
    er, eg = n_ridge**2, n_g**2
    E = eg*np.eye(ff) + (er-eg)*T
    A = (1/eg)*np.eye(ff) + (1/er-1/eg)*T
    sa = np.sin(angle/180*np.pi)
    # kappa per basis index (order p for cos_p / sin_p)
    porder = np.zeros(ff, int)
    for p in range(1, N+1): porder[2*p-1] = porder[2*p] = p
    kap = porder*sa
    Einv = np.linalg.inv(E)
    # B = S_K^T Einv S_K - I ;  S_K couples cos_p <-> sin_p :  (S_K)[sin_p,cos_p] = +-kappa_p
    perm = np.arange(ff)
    sign = np.ones(ff)
    for p in range(1, N+1):
        perm[2*p-1] = 2*p; perm[2*p] = 2*p-1
        sign[2*p-1] = 1.0; sign[2*p] = -1.0
    Bt = (kap*sign)[:, None]*Einv[np.ix_(perm, perm)]*(kap*sign)[None, :] - np.eye(ff)
    lam, W = sla.eigh(Bt, A)          # W^T A W = I
    q = np.sqrt(lam.astype(complex))  # lam<0 -> +i|q|
    k0 = 2*np.pi/wavelength
    X = np.exp(-k0*q*thickness)
    kI = np.conj(np.sqrt((n_I**2-kap**2).astype(complex)))
    kII = np.conj(np.sqrt((n_II**2-kap**2).astype(complex)))
    DI = 1j*kI/n_I**2      # i Z_I
    DII = 1j*kII/n_II**2   # g0
    HI = (W.T*DI[None, :]) @ W
    HII = (W.T*DII[None, :]) @ W
    P = np.linalg.inv(np.diag(q)+HII)
    Hq = HI-np.diag(q)
    S = Hq*(1-X**2)[None, :] + 2*np.diag(q) + 2*((Hq*X[None, :]) @ P)*(X*q)[None, :]
    r0 = DI[0] + 1j/n_I
    b = r0*W[0, :]
    cp = np.linalg.solve(S, b)
    d = 2*P @ (q*X*cp)
    uu = (1-X**2)*cp + X*d
    Tt = W @ d
    Rt = W @ uu; Rt[0] -= 1
    # back to orders -1,0,+1  (m = +-1 <-> basis idx 1,2)
    def to_orders(v):
        vp = (v[1] + v[2]/1j)/np.sqrt(2)   # m=+1 component  (Q[N+1,1]=1/sqrt2, Q[N+1,2]=1/(i sqrt2))
        vm = (v[1] - v[2]/1j)/np.sqrt(2)   # m=-1
        return np.array([vm, v[0], vp])
    T3 = to_orders(Tt); R3 = to_orders(Rt)
    k3I = np.array([kI[1], kI[0], kI[1]]); k3II = np.array([kII[1], kII[0], kII[1]])
    de_r = np.real(R3*np.conj(R3)*k3I/n_I)
    de_t = np.real(T3*np.conj(T3)*k3II/n_II**2)*n_I
    return de_t[0], de_r, de_t, dict(c=c, lam=lam, E=E, A=A, B=Bt, W=W)

if __name__ == '__main__':
    import sys; sys.path.insert(0, '.')
    from oracle import rcwa_oracle as O
    N = 40
    # check basis
    s = np.random.default_rng(0).choice([-1., 1.], 256)
    u = (s+1)/2
    c = fourier_coeffs(u, N)
    _, cref = O.toeplitz_from_cells(u, N, True)
    print('coeff err', np.abs(cref[2*N:]-c).max())
    idx = np.arange(2*N+1)
    cfull = np.concatenate([np.conj(c[:0:-1]), c])
    Tc = cfull[(idx[:, None]-idx[None, :])+2*N]
    Q = unitary(N)
    Tr = Q.conj().T @ Tc @ Q
    print('imag of Q^H T Q', np.abs(Tr.imag).max(), 'closed-form err', np.abs(Tr.real-real_basis_toeplitz(c, N)).max())
    for name, pat, kw in [('G1', s, {}), ('G7', np.repeat(np.random.default_rng(3).choice([-1., 1.], 16), 16), {})]:
        r = O.evaluate(pat, **kw)
        e, der, det, _ = evaluate(pat)
        print(name, r.eff, e, abs(e-r.eff), np.abs(der-r.de_ri[N-1:N+2]).max(), np.abs(det-r.de_ti[N-1:N+2]).max())
    b = np.load('tests/golden/rcwa_batch.npz')
    errs = [abs(evaluate(p.astype(float))[0]-e) for p, e in zip(b['patterns'], b['ref_eff'])]
    print('batch max err', max(errs))
