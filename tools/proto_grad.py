"""Design prototype (NumPy) of the adjoint (reverse-mode) gradient d eff / d s_p that csrc/rcwa_grad.cuh implements.
Not product code and not the oracle.  Same real-symmetric formulation and basis ordering as tools/proto_kernels.py;
the reverse sweep is written step by step the way the kernels execute it (rank-1/rank-2 adjoints of the layer solve,
one dense pass through the generalised symmetric eigenproblem, Toeplitz-diagonal sums, inverse DFT to pixels).

Convention for a real objective f and a complex intermediate z:  zbar = df/dRe z + i df/dIm z,  df = Re(conj(zbar) dz).
"""
import numpy as np
import scipy.linalg as sla

from proto_kernels import host_tables, basis_toeplitz


def coeffs_cont(v, N, tw, D):
    n = v.size
    k = np.arange(2 * N + 1)
    p = np.arange(n)
    return (v[None, :] * tw[(np.outer(k, p)) % n]).sum(1) * D


def toeplitz_adjoint(Tb, N):
    """adjoint of basis_toeplitz: Tbar (ff,ff) real -> cbar (2N+1,) complex (cbar = d/dcr + i d/dci)."""
    ff = 2 * N + 1
    z = 2 * N
    gr = np.zeros(2 * N + 1)
    gi = np.zeros(2 * N + 1)
    r2 = np.sqrt(2.0)
    gr[0] += Tb[z, z]
    for p in range(1, N + 1):
        ic, is_ = 2 * (p - 1), 2 * (p - 1) + 1
        gr[p] += r2 * (Tb[z, ic] + Tb[ic, z])
        gi[p] += -r2 * (Tb[z, is_] + Tb[is_, z])
        for q in range(1, N + 1):
            jc, js = 2 * (q - 1), 2 * (q - 1) + 1
            d = p - q
            sg = 1 if d >= 0 else -1
            gr[abs(d)] += Tb[ic, jc] + Tb[is_, js]
            gr[p + q] += Tb[ic, jc] - Tb[is_, js]
            gi[p + q] += -Tb[ic, js] - Tb[is_, jc]
            gi[abs(d)] += sg * (Tb[ic, js] - Tb[is_, jc])
    return gr + 1j * gi


def hb_vec(Hq, bb):
    return (np.conj(Hq) * bb[:, None]).sum(0)


def setup(N, angle):
    ff = 2 * N + 1
    sa = np.sin(angle / 180 * np.pi)
    kap = np.zeros(ff)
    sk = np.zeros(ff)
    perm = np.arange(ff)
    for p in range(1, N + 1):
        ic, is_ = 2 * (p - 1), 2 * (p - 1) + 1
        kap[ic] = kap[is_] = p * sa
        sk[ic] = p * sa
        sk[is_] = -p * sa
        perm[ic], perm[is_] = is_, ic
    return kap, sk, perm


def forward_backward(s, wavelength=1100., angle=60., order=40, thickness=325., n_ridge=3.551726470588235,
                     n_I=1.45, n_II=1.0, n_g=1.0, check=False):
    N = order
    ff = 2 * N + 1
    z = 2 * N
    s = np.asarray(s, float)
    n = s.size
    u = (s + 1) / 2
    er, eg = n_ridge ** 2, n_g ** 2
    eps = u * (er - eg) + eg
    tw, D = host_tables(n, N)
    cE = coeffs_cont(eps, N, tw, D)
    cA = coeffs_cont(1 / eps, N, tw, D)
    E = basis_toeplitz(cE, N)
    A = basis_toeplitz(cA, N)
    kap, sk, perm = setup(N, angle)
    Einv = np.linalg.inv(E)
    Bm = sk[:, None] * Einv[np.ix_(perm, perm)] * sk[None, :] - np.eye(ff)
    lam, W = sla.eigh(Bm, A)
    # ---------------- layer solve (forward, kept intermediates) ----------------
    k0d = 2 * np.pi / wavelength * thickness
    q = np.sqrt(lam.astype(complex))
    X = np.exp(-k0d * q)
    kI = np.conj(np.sqrt((n_I ** 2 - kap ** 2).astype(complex)))
    kII = np.conj(np.sqrt((n_II ** 2 - kap ** 2).astype(complex)))
    DI = 1j * kI / n_I ** 2
    DII = 1j * kII / n_II ** 2
    HI = (W.T * DI[None, :]) @ W
    HII = (W.T * DII[None, :]) @ W
    P = np.linalg.inv(np.diag(q) + HII)
    Hq = HI - np.diag(q)
    M1 = Hq * X[None, :]
    M2 = M1 @ P
    S = Hq * (1 - X ** 2)[None, :] + 2 * np.diag(q) + 2 * M2 * (X * q)[None, :]
    r0 = DI[z] + 1j / n_I
    b = r0 * W[z, :]
    cp = np.linalg.solve(S, b)
    y = q * X * cp
    dvec = 2 * P @ y
    el = np.zeros(ff, complex)
    el[0] = 1 / np.sqrt(2)
    el[1] = 1j / np.sqrt(2)          # order m=-1 :  (v_cos1 - v_sin1/1j)/sqrt2
    ell = W.T @ el
    t = ell @ dvec
    alpha = np.real(kII[0]) / n_II ** 2 * n_I
    eff = alpha * abs(t) ** 2
    # ---------------- reverse sweep: layer solve (all rank-1 / rank-2) ----------------
    tb = 2 * alpha * t
    db = np.conj(ell) * tb                      # dvec bar
    ellb = np.conj(dvec) * tb
    Wb_rows = {0: np.real(el[0] * np.conj(ellb)), 1: np.real(el[1] * np.conj(ellb))}   # Wbar rows 0,1
    yb = 2 * P.conj().T @ db
    qb = np.conj(X * cp) * yb
    Xb = np.conj(q * cp) * yb
    cpb = np.conj(q * X) * yb
    bb = np.linalg.solve(S.conj().T, cpb)
    Wb_rows[z] = np.real(np.conj(bb) * r0)
    # Sbar = -bb cp^H  (rank one): never formed
    # Hq D1 term
    D1b = -(np.conj(Hq) * bb[:, None]).sum(0) * np.conj(cp)      # sum_i conj(Hq_ij) Sbar_ij
    Xb += np.conj(-2 * X) * D1b
    qb += 2 * (-bb * np.conj(cp))                                  # 2 Sbar_jj
    # 2 M2 D_{Xq} term
    xqb = 2 * (-(np.conj(M2) * bb[:, None]).sum(0) * np.conj(cp))
    Xb += np.conj(q) * xqb
    qb += np.conj(X) * xqb
    # M2bar = 2 Sbar D_{Xq}^H = -bb w2^H  with  w2 = 2 cp * (X q)
    w2 = 2 * cp * (X * q)
    # M1bar = M2bar P^H = -bb (P w2)^H ; Pbar = 2 db y^H + M1^H M2bar = 2 db y^H - (M1^H bb) w2^H
    pw2 = P @ w2
    m1hb = M1.conj().T @ bb
    # Hqbar = Sbar D1^H + M1bar DX^H = -bb [cp*(1-X^2)]^H - bb [pw2 * X]^H  = -bb h^H,  h = cp(1-X^2) + X pw2
    h = cp * (1 - X ** 2) + X * pw2
    Xb += -(np.conj(Hq) * bb[:, None]).sum(0) * np.conj(pw2)       # sum_i conj(Hq_ij) M1bar_ij
    qb += bb * np.conj(h)                                          # - Hqbar_jj
    # HIbar = Hqbar = -bb h^H
    # Gbar = -P^H Pbar P^H = -2 (P^H db)(P y)^H + (P^H m1hb)(P w2)^H
    g1a, g1b = -2 * (P.conj().T @ db), P @ y
    g2a, g2b = P.conj().T @ m1hb, pw2
    qb += g1a * np.conj(g1b) + g2a * np.conj(g2b)                  # Gbar_jj
    # HIIbar = Gbar = g1a g1b^H + g2a g2b^H
    qb += np.conj(-k0d * X) * Xb
    lamb = np.real(np.conj(qb) / (2 * q))
    # Wbar = Re(DI W conj(HIbar + HIbar^T)) + Re(DII W conj(HIIbar + HIIbar^T)) + row terms
    Wb = np.zeros((ff, ff))
    for Dg, pairs in ((DI, [(-bb, h)]), (DII, [(g1a, g1b), (g2a, g2b)])):
        for a, c in pairs:
            # Hbar = a c^H  -> conj(Hbar) = conj(a) c^T ; conj(Hbar^T) = c conj(a)^T... careful: Hbar^T = conj(c) a^T, conj -> c a^H
            Wb += np.real(np.outer(Dg * (W @ np.conj(a)), c) + np.outer(Dg * (W @ c), np.conj(a)))
    for r, v in Wb_rows.items():
        Wb[r, :] += v
    # ---------------- reverse sweep: generalised symmetric eigenproblem ----------------
    # dW = W Cm, Cm_ij = (dB~_ij - lam_j dA~_ij)/(lam_j - lam_i): only K_ij = (Pm_ij - Pm_ji)/(lam_j - lam_i) is singular.
    # Invariance of the layer solve under W -> W R, diag(f(lam)) -> R^T diag(f) R gives the Daleckii-Krein form
    #   K_ij = sum_f Re[ conj(Phibar_f,ij + Phibar_f,ji) (f_j - f_i)/(lam_j - lam_i) ],
    # Phibar_f the (rank-one) adjoints of the diagonal matrices q, X, 1 - X^2, X q taken as full matrices.
    Pm = W.T @ Wb
    pairs = {
        'q': [(g1a, g1b), (g2a, g2b), (bb, h - 2 * cp)],
        'X': [(-hb_vec(Hq, bb), pw2)],
        'D1': [(-hb_vec(Hq, bb), cp)],
        'XQ': [(yb - 2 * g2a, cp)],
    }
    Ssum = {}
    for key, prs in pairs.items():
        M = np.zeros((ff, ff), complex)
        for a, c in prs:
            M += np.outer(a, np.conj(c))
        Ssum[key] = M + M.T
    qi, qj = q[:, None], q[None, :]
    Xi, Xj = X[:, None], X[None, :]
    dq = 1.0 / (qi + qj + np.eye(ff))                      # (q_j - q_i)/(lam_j - lam_i)
    zarg = -k0d * (qj - qi)
    small = np.abs(zarg) < 1e-3
    zs = np.where(small, 1.0, zarg)
    phi = np.where(small, 1 + zarg / 2 + zarg ** 2 / 6 + zarg ** 3 / 24, (np.exp(zs) - 1) / zs)
    dX = -k0d * Xi * phi * dq                                # (X_j - X_i)/(lam_j - lam_i)
    dD1 = -(Xi + Xj) * dX
    dXQ = Xj * dq + qi * dX
    K = np.real(np.conj(Ssum['q']) * dq + np.conj(Ssum['X']) * dX + np.conj(Ssum['D1']) * dD1 + np.conj(Ssum['XQ']) * dXQ)
    np.fill_diagonal(K, 0.0)
    if check:
        dl = lam[None, :] - lam[:, None] + np.eye(ff)
        Kref = (Pm - Pm.T) / dl
        np.fill_diagonal(Kref, 0)
        rel = np.abs(lam[None, :] - lam[:, None]) / (np.abs(lam[None, :]) + np.abs(lam[:, None]))
        mask = rel > 1e-4
        print('K check: max |K - Kref| on well separated pairs', np.abs(K - Kref)[mask].max(), 'max|K|', np.abs(K).max(),
              'min rel gap', rel[~np.eye(ff, dtype=bool)].min())
    Ps = 0.5 * (Pm + Pm.T)
    GB = 0.5 * K + np.diag(lamb)
    GA = -0.5 * (Ps + 0.5 * (lam[:, None] + lam[None, :]) * K)
    np.fill_diagonal(GA, -(lam * lamb + 0.5 * np.diag(Pm)))
    Bb = W @ GB @ W.T
    Ab = W @ GA @ W.T
    Eib = np.zeros((ff, ff))
    Eib[np.ix_(perm, perm)] = sk[:, None] * Bb * sk[None, :]
    Eb = -Einv @ Eib @ Einv
    # ---------------- Toeplitz adjoint and inverse DFT to pixels ----------------
    gE = np.conj(toeplitz_adjoint(Eb, N)) * D
    gA = np.conj(toeplitz_adjoint(Ab, N)) * D
    k = np.arange(2 * N + 1)
    p = np.arange(n)
    ph = tw[(np.outer(k, p)) % n]               # (2N+1, n)
    vE = np.real(gE @ ph)
    vA = np.real(gA @ ph)
    grad = 0.5 * ((er - eg) * vE - (er - eg) / eps ** 2 * vA)
    return eff, grad


if __name__ == '__main__':
    import sys
    sys.path.insert(0, '.')
    from oracle import rcwa_oracle as O
    pat = np.repeat(np.random.default_rng(3).choice([-1., 1.], 16), 16)
    eff, g = forward_backward(pat)
    print('eff', eff, 'oracle', O.efficiency(pat))
    idx = [5, 100, 200]
    fd = O.fd_gradient(pat, idx)
    print('adjoint', g[idx]); print('fd     ', fd); print('rel err', np.abs(g[idx] - fd) / np.abs(fd))
    s = np.random.default_rng(0).choice([-1., 1.], 256)
    eff, g = forward_backward(s)
    idx = [0, 17, 128, 255]
    fd = O.fd_gradient(s, idx)
    print('G1 adjoint', g[idx]); print('G1 fd     ', fd); print('rel err', np.abs(g[idx] - fd) / np.abs(fd))
    # near-degenerate pairs: two identical ridges far apart (tunnel splitting), translation equivariance
    pat = -np.ones(256); pat[20:52] = 1; pat[150:182] = 1; pat[200:210] = 1
    eff, g = forward_backward(pat, check=True)
    eff2, g2 = forward_backward(np.roll(pat, 37), check=True)
    print('two ridges: eff', eff, 'max|g|', np.abs(g).max(), 'equivariance err', np.abs(np.roll(g, 37) - g2).max())
    idx = [10, 30, 100, 160, 205]
    fd = O.fd_gradient(pat, idx)
    print('adjoint', g[idx]); print('fd     ', fd)
