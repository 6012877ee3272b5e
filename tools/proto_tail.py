"""NumPy prototype of the 'significant-mode' layer solve (design aid for rk_tail; NOT the oracle, NOT product code).

Modes sorted by DESCENDING eigenvalue: [insignificant evanescent (X < xtol) | significant evanescent | propagating].
Only the insignificant block (size n_i) is eliminated without pivoting -- there S_ii = (H_I + Q)_ii and
G_ii = (H_II + Q)_ii are complex symmetric with positive-definite real part.  Everything that couples top and bottom of
the layer lives on the trailing block s' (size m = n - n_i), solved densely with partial pivoting.
"""
import sys
import numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tools')
import proto_kernels as PK


def partial_elim(M, ni, cols=None, rows=None, corner=None):
    """No-pivot elimination of the first ni pivots of M (n x n) with optional extra columns (n x c), extra rows (r x n)
    and corner (r x c).  Returns trailing blocks."""
    M = M.copy(); n = M.shape[0]
    cols = None if cols is None else cols.copy()
    rows = None if rows is None else rows.copy()
    corner = None if corner is None else corner.copy()
    growth = 0.0
    for k in range(ni):
        p = 1.0 / M[k, k]
        l = M[k + 1:, k] * p
        M[k + 1:, k + 1:] -= np.outer(l, M[k, k + 1:])
        if cols is not None:
            cols[k + 1:, :] -= np.outer(l, cols[k, :])
        if rows is not None:
            lr = rows[:, k] * p
            rows[:, k + 1:] -= np.outer(lr, M[k, k + 1:])
            if corner is not None:
                corner -= np.outer(lr, cols[k, :])
        growth = max(growth, np.abs(M[k + 1:, k + 1:]).max())
    return (M[ni:, ni:], None if cols is None else cols[ni:], None if rows is None else rows[:, ni:], corner, growth)


def tail_m(lam, W, kap, k0d, n_I, n_II, N, xtol=1e-9):
    ff = 2 * N + 1; z = 2 * N
    order = np.argsort(-lam, kind='stable')
    lam = lam[order]; W = W[:, order]
    q = np.sqrt(lam.astype(complex)); X = np.exp(-k0d * q)
    ni = int(np.sum((lam > 0) & (np.abs(X) < xtol)))
    kI = np.conj(np.sqrt((n_I ** 2 - kap ** 2).astype(complex))); kII = np.conj(np.sqrt((n_II ** 2 - kap ** 2).astype(complex)))
    DI = 1j * kI / n_I ** 2; DII = 1j * kII / n_II ** 2
    HI = (W.T * DI[None, :]) @ W; HII = (W.T * DII[None, :]) @ W
    rows3 = [0, 1, z]
    Wr = W[rows3, :]                                   # (3, n) the three output harmonics
    # ---- G phase: P_s's' and g_r = (G^-1 w_r)_s'
    G = np.diag(q) + HII
    Gt, rhs_t, _, _, gg = partial_elim(G, ni, cols=Wr.T.astype(complex))
    Pss = np.linalg.inv(Gt)
    g = Pss @ rhs_t                                    # (m, 3)
    # ---- S phase
    Xs = X.copy(); Xs[:ni] = 0.0
    Hq = HI - np.diag(q)
    S = Hq * (1 - Xs ** 2)[None, :] + 2 * np.diag(q)
    S[:, ni:] += 2 * ((Hq[:, ni:] * Xs[None, ni:]) @ Pss) * (Xs * q)[None, ni:]
    r0 = DI[z] + 1j / n_I
    b = (r0 * W[z, :]).astype(complex)[:, None]
    f = (Wr * (1 - Xs ** 2)[None, :]).astype(complex)  # (3, n) functionals of c
    St, bt, ft, corner, gs = partial_elim(S, ni, cols=b, rows=f, corner=np.zeros((3, 1), complex))
    cs = np.linalg.solve(St, bt)[:, 0]                 # c on s'
    fc = -corner[:, 0] + ft @ cs                       # f_r . c  (eliminated part sits in the corner with a minus sign)
    y = (q * Xs)[ni:] * cs
    ds = 2 * Pss @ y
    Tt = 2 * g.T @ y
    Rt = fc + (Wr[:, ni:] * Xs[None, ni:]) @ ds
    Rt[2] -= 1

    def to_orders(v):
        vp = (v[0] + v[1] / 1j) / np.sqrt(2); vm = (v[0] - v[1] / 1j) / np.sqrt(2)
        return np.array([vm, v[2], vp])
    T3 = to_orders(Tt); R3 = to_orders(Rt)
    k3I = np.array([kI[0], kI[z], kI[0]]); k3II = np.array([kII[0], kII[z], kII[0]])
    de_r = np.real(R3 * np.conj(R3) * k3I / n_I); de_t = np.real(T3 * np.conj(T3) * k3II / n_II ** 2) * n_I
    return de_r, de_t, dict(ni=ni, growth=max(gg, gs), npos=int(np.sum(lam > 0)))


def evaluate(s, wavelength=1100., angle=60., order=40, thickness=325., n_ridge=3.551726470588235, xtol=1e-9):
    e, der, det, st = PK.evaluate(s, wavelength, angle, order, thickness, n_ridge=n_ridge, stages=True)
    N = order; ff = 2 * N + 1
    sa = np.sin(angle / 180 * np.pi)
    kap = np.zeros(ff)
    for p in range(1, N + 1):
        kap[2 * (p - 1)] = kap[2 * (p - 1) + 1] = p * sa
    k0d = 2 * np.pi / wavelength * thickness
    der2, det2, info = tail_m(st['lam'], st['W'], kap, k0d, 1.45, 1.0, N, xtol)
    return der, det, der2, det2, info


if __name__ == '__main__':
    from oracle import rcwa_oracle as O
    g = np.load('tests/golden/rcwa_golden.npz')
    names = sorted({k.split('/')[0] for k in g.files})
    for nme in names:
        p = g[nme + '/pattern']; wl = float(g[nme + '/wavelength']); od = int(g[nme + '/order'])
        if od > 60: continue
        der, det, der2, det2, info = evaluate(p, wl, float(g[nme + '/angle']), od, float(g[nme + '/thickness']), n_ridge=O.n_silicon_real(wl))
        print(f"{nme:30s} ni={info['ni']:3d}/{2*od+1} growth={info['growth']:.2e} dT={np.abs(det2[::-1]-g[nme+'/ref_tran_cut']).max():.2e} dR={np.abs(der2[::-1]-g[nme+'/ref_refl_cut']).max():.2e}")
    b = np.load('tests/golden/rcwa_batch.npz')
    worst = 0; nis = []
    for p, tc, rc in zip(b['patterns'][:96], b['ref_tran_cut'][:96], b['ref_refl_cut'][:96]):
        der, det, der2, det2, info = evaluate(p.astype(float))
        worst = max(worst, np.abs(det2[::-1] - tc).max(), np.abs(der2[::-1] - rc).max()); nis.append(info['ni'])
    print('batch96 worst', worst, 'ni min/mean/max', min(nis), np.mean(nis), max(nis))
