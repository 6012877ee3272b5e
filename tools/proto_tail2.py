"""NumPy prototype of the REAL-arithmetic bordered layer solve (design aid for rk_tail2; NOT the oracle, NOT product code).

Modes sorted by DESCENDING eigenvalue: [I: insignificant evanescent | S: significant evanescent | P: propagating].
Unknowns: top field u and bottom field hb in the mode basis (impedance form) for evanescent modes,
          (c, hb) (amplitude form, no resonance singularity) for the few propagating modes.
   top   : (R_I  + i sum_r bI_r  w_r w_r^T) h_t + e_t = r0 w_z      h_t = u (evan.) | (1-X^2) c + X hb (prop.)
   bottom: (R_II + i sum_r bII_r w_r w_r^T) hb  + e_b = 0
   e_t = C u - Sg hb, e_b = C hb - Sg u (evan., C = q coth, Sg = q csch) ; e_t = q((1+X^2) c - X hb), e_b = q hb - 2 q X c (prop.)
The rank-3 imaginary parts become auxiliary unknowns tau_r = i bI_r (w_r . h_t), sigma_r = i bII_r (w_r . hb), so every
evanescent pivot is eliminated in REAL arithmetic on a symmetric positive definite matrix without pivoting; what is left
is a complex system of size 2 n_p + 6 whose solution carries the outputs: R_r = tau_r / (i bI_r) - delta, T_r = sigma_r / (i bII_r).
"""
import sys
import numpy as np
import scipy.linalg as sla
sys.path.insert(0, '.'); sys.path.insert(0, 'tools')
import proto_kernels as PK


def modes(s, wavelength=1100., angle=60., order=40, n_ridge=3.551726470588235, n_g=1.0):
    """Fast front: same matrices as proto_kernels.evaluate, LAPACK for the generalised eigenproblem."""
    N = order; ff = 2 * N + 1
    u = (np.asarray(s, float) + 1) / 2
    tw, D = PK.host_tables(u.size, N)
    c = PK.coeffs(u, N, tw, D)
    T = PK.basis_toeplitz(c, N)
    er, eg = n_ridge ** 2, n_g ** 2
    E = eg * np.eye(ff) + (er - eg) * T
    A = (1 / eg) * np.eye(ff) + (1 / er - 1 / eg) * T
    sa = np.sin(angle / 180 * np.pi)
    kap = np.zeros(ff); sk = np.zeros(ff); perm = np.arange(ff)
    for p in range(1, N + 1):
        ic, is_ = 2 * (p - 1), 2 * (p - 1) + 1
        kap[ic] = kap[is_] = p * sa
        sk[ic] = p * sa; sk[is_] = -p * sa
        perm[ic], perm[is_] = is_, ic
    Einv = np.linalg.inv(E)
    Bm = sk[:, None] * Einv[np.ix_(perm, perm)] * sk[None, :] - np.eye(ff)
    lam, W = sla.eigh(Bm, A)
    return lam, W, kap


def side_eliminate(R, Cdiag, Xcols, ni, ne, gj_from=None):
    """Symmetric no-pivot elimination of pivots [0, ne) of M = R + diag(Cdiag) (real), carrying the real columns Xcols.
    Pivots >= gj_from are Gauss-Jordan steps (rows above the pivot inside the window [ni, n) are updated too), so that
    afterwards M[S,S] holds the inverse of the Schur complement on S and Xcols[S] holds that inverse times the columns.
    Returns M, Xcols, Ysave (row of Xcols at each pivot), dsave (pivots)."""
    n = R.shape[0]
    M = R + np.diag(Cdiag)
    X = Xcols.copy()
    Ys = np.zeros((ne, X.shape[1])); ds = np.zeros(ne)
    for k in range(ne):
        d = M[k, k]; assert d > 0
        ds[k] = d; Ys[k] = X[k]
        rowk = M[k].copy(); xk = X[k].copy()
        if gj_from is None or k < gj_from:
            l = M[:, k] / d; l[:k + 1] = 0.0
            M -= np.outer(l, rowk); X -= np.outer(l, xk)
        else:           # Gauss-Jordan (sweep) step on the window [ni, n)
            l = M[:, k] / d; l[:ni] = 0.0; l[k] = 0.0
            M -= np.outer(l, rowk); X -= np.outer(l, xk)
            M[:, k] = -l; M[k, :] = rowk / d; M[k, k] = 1.0 / d
            X[k] = xk / d
            M[:ni, k] = 0.0
    return M, X, Ys, ds


def tail2(lam, W, kap, k0d, n_I, n_II, N, xtol=1e-12):
    ff = 2 * N + 1; z = 2 * N
    order = np.argsort(-lam, kind='stable')
    lam = lam[order]; W = W[:, order]
    ne = int(np.sum(lam > 0)); npr = ff - ne
    q = np.sqrt(lam.astype(complex))
    x = k0d * np.sqrt(np.maximum(lam, 0))
    Xr = np.exp(-x)
    ni = int(np.sum((lam > 0) & (Xr < xtol)))
    ns = ne - ni
    E = slice(0, ne); S = slice(ni, ne); P = slice(ne, ff)
    om = -np.expm1(-2 * x[:ne])                       # 1 - X^2
    Cd = np.zeros(ff); Sg = np.zeros(ff)
    Cd[:ne] = np.sqrt(lam[:ne]) * (1 + Xr[:ne] ** 2) / om
    Sg[:ne] = np.sqrt(lam[:ne]) * 2 * Xr[:ne] / om
    Cd[:ni] = np.sqrt(lam[:ni]); Sg[:ni] = 0.0
    aI = n_I ** 2 - kap ** 2; aII = n_II ** 2 - kap ** 2
    dI_re = np.where(aI < 0, np.sqrt(np.abs(aI)), 0.0) / n_I ** 2; dI_im = np.where(aI >= 0, np.sqrt(np.abs(aI)), 0.0) / n_I ** 2
    dII_re = np.where(aII < 0, np.sqrt(np.abs(aII)), 0.0) / n_II ** 2; dII_im = np.where(aII >= 0, np.sqrt(np.abs(aII)), 0.0) / n_II ** 2
    pI = np.nonzero(aI >= 0)[0]; pII = np.nonzero(aII >= 0)[0]     # propagating harmonics of the two media
    RI = (W.T * dI_re[None, :]) @ W; RII = (W.T * dII_re[None, :]) @ W
    WrI = W[pI, :].T.copy(); WrII = W[pII, :].T.copy()              # (ff, 3) real border columns
    nI_, nII_ = len(pI), len(pII)
    # ---- top: eliminate I (LDL^T), Gauss-Jordan on S
    Mt, Xt, Yt, dt = side_eliminate(RI, Cd, WrI, ni, ne, gj_from=ni)
    TinvSS = Mt[S, S]                                   # (Schur on S)^-1
    GP = Mt[S, P]                                       # Tinv * M_SP
    Gw = Xt[S, :]                                       # Tinv * Wr_S
    # ---- bottom: eliminate I, fold the top S block in, eliminate S
    #   extra columns of the bottom system: [WrII (sigma) | Sg G_P (h_t,P) | Sg G_w (tau~)]
    Xb0 = np.concatenate([WrII, np.zeros((ff, npr + nI_))], axis=1)
    Mb, Xb, Yb1, db1 = side_eliminate(RII, Cd, Xb0, ni, ni)
    Mb[S, S] -= Sg[S, None] * TinvSS * Sg[None, S]
    Xb[S, nII_:nII_ + npr] += Sg[S, None] * GP
    Xb[S, nII_ + npr:] += Sg[S, None] * Gw
    # continue the elimination over S on the window (pivots ni..ne-1); rows < ni are dead
    Ms = Mb[ni:, ni:].copy(); Xs = Xb[ni:].copy()
    Yb2 = np.zeros((ns, Xs.shape[1])); db2 = np.zeros(ns)
    for k in range(ns):
        d = Ms[k, k]; assert d > 0
        db2[k] = d; Yb2[k] = Xs[k]
        l = Ms[:, k] / d; l[:k + 1] = 0.0
        rowk = Ms[k].copy(); xk = Xs[k].copy()
        Ms -= np.outer(l, rowk); Xs -= np.outer(l, xk)
    # ---- border system.  unknown order: [h_t,P (npr) | tau~ (nI_) | hb_P (npr) | sigma (nII_) | c_P (npr)]
    nb = 3 * npr + nI_ + nII_
    o_ht, o_tau, o_hb, o_sig, o_c = 0, npr, npr + nI_, 2 * npr + nI_, 2 * npr + nI_ + nII_
    Z = np.zeros((nb, nb), complex); rhs = np.zeros(nb, complex)
    qP = q[P]; XP = np.exp(-k0d * qP)
    # top P rows (equations of the top boundary projected on the propagating modes), after elimination of u_E:
    #   Schur(M_t)[P,P] h_t,P + Schur(Xt)[P,:] tau~ + q(1+X^2) c - q X hb_P + (coupling to hb_S eliminated later) = 0
    # all couplings through eliminated unknowns are -Y^T D^-1 Y corner terms; assemble the REAL symmetric corner first:
    #   real border unknowns (columns carried): top: [h_t,P (tile cols P), tau~ (Xt cols)], bottom: [hb_P (tile cols P), sigma, h_t,P, tau~]
    corner = np.zeros((nb, nb))
    # top side corner: after the LDL^T / GJ pivots the tile block [P,P] and Xt[P] already hold the Schur complements;
    # the tau~-tau~ corner is -sum_k y_k y_k^T / d_k
    corner[o_ht:o_ht + npr, o_ht:o_ht + npr] += Mt[P, P]
    corner[o_ht:o_ht + npr, o_tau:o_tau + nI_] += Xt[P, :]
    corner[o_tau:o_tau + nI_, o_ht:o_ht + npr] += Xt[P, :].T
    corner[o_tau:o_tau + nI_, o_tau:o_tau + nI_] += -(Yt.T / dt[None, :]) @ Yt
    # bottom side: tile block [P,P] -> hb_P,hb_P ; Xs rows P -> hb_P x [sigma | h_t,P | tau~] ; Y-corner on [sigma | h_t,P | tau~]
    PP = slice(ns, ns + npr)
    corner[o_hb:o_hb + npr, o_hb:o_hb + npr] += Ms[PP, PP]
    cols_b = np.concatenate([np.arange(o_sig, o_sig + nII_), np.arange(o_ht, o_ht + npr), np.arange(o_tau, o_tau + nI_)])
    corner[np.ix_(np.arange(o_hb, o_hb + npr), cols_b)] += Xs[PP, :]
    corner[np.ix_(cols_b, np.arange(o_hb, o_hb + npr))] += Xs[PP, :].T
    Yb = np.concatenate([Yb1, Yb2]); db = np.concatenate([db1, db2])
    corner[np.ix_(cols_b, cols_b)] += -(Yb.T / db[None, :]) @ Yb
    Z[:, :] = corner
    # propagating-mode slab terms and the imaginary (radiation) parts
    for j in range(npr):
        Z[o_ht + j, o_c + j] += qP[j] * (1 + XP[j] ** 2)
        Z[o_ht + j, o_hb + j] += -qP[j] * XP[j]
        Z[o_hb + j, o_hb + j] += qP[j]
        Z[o_hb + j, o_c + j] += -2 * qP[j] * XP[j]
        # link row: h_t,P - (1 - X^2) c - X hb = 0
        Z[o_c + j, o_ht + j] = 1.0; Z[o_c + j, o_c + j] = -(1 - XP[j] ** 2); Z[o_c + j, o_hb + j] = -XP[j]
    r0 = 1j * dI_im[z] + 1j / n_I
    for a, r in enumerate(pI):
        Z[o_tau + a, o_tau + a] += 1j / dI_im[r]
        if r == z:
            rhs[o_tau + a] = -(1j / dI_im[r]) * r0
    for a, r in enumerate(pII):
        Z[o_sig + a, o_sig + a] += 1j / dII_im[r]
    sol = np.linalg.solve(Z, rhs)
    tau = sol[o_tau:o_tau + nI_].copy(); sig = sol[o_sig:o_sig + nII_]
    tau[list(pI).index(z)] += r0
    Rv = np.zeros(ff, complex); Tv = np.zeros(ff, complex)
    Rv[pI] = tau / (1j * dI_im[pI]); Tv[pII] = sig / (1j * dII_im[pII])
    Rt = Rv[[0, 1, z]]; Tt = Tv[[0, 1, z]]
    Rt[2] -= 1

    def to_orders(v):
        vp = (v[0] + v[1] / 1j) / np.sqrt(2); vm = (v[0] - v[1] / 1j) / np.sqrt(2)
        return np.array([vm, v[2], vp])
    T3 = to_orders(Tt); R3 = to_orders(Rt)
    kI = np.conj(np.sqrt((n_I ** 2 - kap ** 2).astype(complex))); kII = np.conj(np.sqrt((n_II ** 2 - kap ** 2).astype(complex)))
    k3I = np.array([kI[0], kI[z], kI[0]]); k3II = np.array([kII[0], kII[z], kII[0]])
    de_r = np.real(R3 * np.conj(R3) * k3I / n_I); de_t = np.real(T3 * np.conj(T3) * k3II / n_II ** 2) * n_I
    return de_r, de_t, dict(ni=ni, ns=ns, npr=npr, nb=nb, condZ=np.linalg.cond(Z), dmin=min(dt.min(), db.min()))


def evaluate(s, wavelength=1100., angle=60., order=40, thickness=325., n_ridge=3.551726470588235, xtol=1e-12):
    lam, W, kap = modes(s, wavelength, angle, order, n_ridge)
    k0d = 2 * np.pi / wavelength * thickness
    return tail2(lam, W, kap, k0d, 1.45, 1.0, order, xtol)


if __name__ == '__main__':
    from oracle import rcwa_oracle as O
    g = np.load('tests/golden/rcwa_golden.npz')
    names = sorted({k.split('/')[0] for k in g.files})
    for nme in names:
        p = g[nme + '/pattern']; wl = float(g[nme + '/wavelength']); od = int(g[nme + '/order'])
        der, det, info = evaluate(p, wl, float(g[nme + '/angle']), od, float(g[nme + '/thickness']), n_ridge=O.n_silicon_real(wl))
        print(f"{nme:30s} ni={info['ni']:3d} ns={info['ns']:3d} np={info['npr']:2d} nb={info['nb']:2d} condZ={info['condZ']:.1e} dmin={info['dmin']:.2e} "
              f"dT={np.abs(det[::-1]-g[nme+'/ref_tran_cut']).max():.2e} dR={np.abs(der[::-1]-g[nme+'/ref_refl_cut']).max():.2e}")
    b = np.load('tests/golden/rcwa_batch.npz')
    worst = 0; conds = []
    for p, tc, rc in zip(b['patterns'], b['ref_tran_cut'], b['ref_refl_cut']):
        der, det, info = evaluate(p.astype(float))
        worst = max(worst, np.abs(det[::-1] - tc).max(), np.abs(der[::-1] - rc).max()); conds.append(info['condZ'])
    print('batch', len(b['patterns']), 'worst', worst, 'condZ max', max(conds))
