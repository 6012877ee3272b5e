"""Kernel-level NumPy prototype: the same steps, in the same order and with the same pivoting-free
factorizations, that the CUDA kernels in csrc/ execute.  Design aid + source of stage intermediates
for tests (it is NOT the oracle and NOT product code).

Basis ordering used by the kernels: idx 2(p-1)=sqrt2 cos(p G x), 2(p-1)+1 = sqrt2 sin(p G x), p=1..N; idx 2N = 1.
"""
import numpy as np


def host_tables(n_cells, N):
    """twiddle[p] = exp(-2 pi i p/n_cells);  D[k] up-sampling factor (k=0..2N)."""
    R = 4 * N + 2
    M = n_cells * R
    k = np.arange(2 * N + 1)
    tw = np.exp(-2j * np.pi * np.arange(n_cells) / n_cells)
    with np.errstate(invalid='ignore', divide='ignore'):
        D = (1 - np.exp(-2j * np.pi * k / n_cells)) / (R * (1 - np.exp(-2j * np.pi * k / M)))
    D[k % n_cells == 0] = 1.0
    return tw, D / n_cells


def coeffs(u, N, tw, D):
    n = u.size
    k = np.arange(2 * N + 1)
    p = np.arange(n)
    return (u[None, :] * tw[(np.outer(k, p)) % n]).sum(1) * D


def basis_toeplitz(c, N):
    ff = 2 * N + 1
    cr, ci = c.real, c.imag
    S = np.zeros((ff, ff))
    r2 = np.sqrt(2.0)
    z = 2 * N
    S[z, z] = cr[0]
    for p in range(1, N + 1):
        ic, is_ = 2 * (p - 1), 2 * (p - 1) + 1
        S[z, ic] = S[ic, z] = r2 * cr[p]
        S[z, is_] = S[is_, z] = -r2 * ci[p]
        for q in range(1, N + 1):
            jc, js = 2 * (q - 1), 2 * (q - 1) + 1
            d = p - q
            crd = cr[abs(d)]
            cid = ci[abs(d)] * (1 if d >= 0 else -1)
            S[ic, jc] = crd + cr[p + q]
            S[is_, js] = crd - cr[p + q]
            S[ic, js] = -ci[p + q] + cid
            S[is_, jc] = -ci[p + q] - cid
    return S


def gj_inverse(a):
    """in-place Gauss-Jordan inverse without pivoting (SPD / complex symmetric with PD rotated real part)."""
    a = a.copy()
    n = a.shape[0]
    for k in range(n):
        p = 1.0 / a[k, k]
        row = a[k, :].copy()
        col = a[:, k].copy()
        a -= np.outer(col * p, row)
        a[k, :] = row * p
        a[:, k] = -col * p
        a[k, k] = p
    return a


def ldlt(a):
    """right-looking LDL^T without pivoting on the full symmetric matrix; returns unit-lower L and D."""
    a = a.copy()
    n = a.shape[0]
    for k in range(n):
        piv = a[k, k]
        col = a[k + 1:, k].copy()
        a[k + 1:, k + 1:] -= np.outer(col, col) / piv
    D = np.diag(a).copy()
    L = np.tril(a, -1) / D[None, :] + np.eye(n)
    return L, D


def reduce_to_standard(Bm, L, D):
    """C = D^-1/2 L^-1 B L^-T D^-1/2 by row then column forward substitution."""
    X = Bm.copy()
    n = X.shape[0]
    for k in range(n):
        X[k + 1:, :] -= np.outer(L[k + 1:, k], X[k, :])
    for k in range(n):
        X[:, k + 1:] -= np.outer(X[:, k], L[k + 1:, k])
    s = 1.0 / np.sqrt(D)
    return X * s[:, None] * s[None, :]


def tridiagonalize(C):
    """Householder, lower, forward.  Returns d, e(sub-diagonal, len n-1), V (reflectors in columns), beta."""
    A = C.copy()
    n = A.shape[0]
    d = np.zeros(n)
    e = np.zeros(n - 1)
    V = np.zeros((n, n))
    beta = np.zeros(n)
    for k in range(n - 2):
        x = A[k + 1:, k].copy()
        sigma = x @ x
        x0 = x[0]
        tail2 = sigma - x0 * x0
        d[k] = A[k, k]
        if tail2 <= 0.0 or sigma == 0.0:      # nothing to annihilate
            e[k] = x0
            beta[k] = 0.0
            continue
        alpha = -np.copysign(np.sqrt(sigma), x0)
        v = x.copy()
        v[0] = x0 - alpha
        vn2 = 2.0 * (sigma - alpha * x0)
        b = 2.0 / vn2
        A22 = A[k + 1:, k + 1:]
        p = b * (A22 @ v)
        K = 0.5 * b * (p @ v)
        w = p - K * v
        A22 -= np.outer(v, w) + np.outer(w, v)
        e[k] = alpha
        V[k + 1:, k] = v
        beta[k] = b
    d[n - 2] = A[n - 2, n - 2]
    d[n - 1] = A[n - 1, n - 1]
    e[n - 2] = A[n - 1, n - 2]
    return d, e, V, beta


def form_q(V, beta):
    n = V.shape[0]
    Q = np.eye(n)
    for k in range(n - 3, -1, -1):
        if beta[k] == 0.0:
            continue
        v = V[k + 1:, k]
        w = beta[k] * (v @ Q[k + 1:, k + 1:])
        Q[k + 1:, k + 1:] -= np.outer(v, w)
    return Q


def back_transform(Q, L, D):
    """M0 = L^-T D^-1/2 Q   (so that M0^T A M0 = I-transform of Q)."""
    M = Q / np.sqrt(D)[:, None]
    n = M.shape[0]
    for k in range(n - 1, 0, -1):
        M[:k, :] -= np.outer(L[k, :k], M[k, :])
    return M


def ql_rotations(d, e, max_iter=60):
    """Implicit QL with Wilkinson shift on a symmetric tridiagonal matrix; returns eigenvalues (in place order)
    and the list of plane rotations (col i, c, s) to be applied to columns (i, i+1) of the accumulated matrix:
        z_i1' = s*z_i + c*z_i1 ; z_i' = c*z_i - s*z_i1
    """
    d = d.copy()
    n = d.size
    e = np.concatenate([e.copy(), [0.0]])
    rots = []
    eps = np.finfo(float).eps
    ok = True
    for l in range(n):
        it = 0
        while True:
            m = l
            while m < n - 1:
                dd = abs(d[m]) + abs(d[m + 1])
                if abs(e[m]) <= eps * dd:
                    break
                m += 1
            if m == l:
                break
            it += 1
            if it > max_iter:
                ok = False
                break
            g = (d[l + 1] - d[l]) / (2.0 * e[l])
            r = np.hypot(g, 1.0)
            g = d[m] - d[l] + e[l] / (g + np.copysign(r, g))
            s = 1.0
            c = 1.0
            p = 0.0
            i = m - 1
            underflow = False
            while i >= l:
                f = s * e[i]
                b = c * e[i]
                r = np.hypot(f, g)
                e[i + 1] = r
                if r == 0.0:
                    d[i + 1] -= p
                    e[m] = 0.0
                    underflow = True
                    break
                s = f / r
                c = g / r
                g = d[i + 1] - p
                r = (d[i] - g) * s + 2.0 * c * b
                p = s * r
                d[i + 1] = g + p
                g = c * r - b
                rots.append((i, c, s))
                i -= 1
            if underflow:
                continue
            d[l] -= p
            e[l] = g
            e[m] = 0.0
    return d, rots, ok


def apply_rotations(M, rots):
    Z = M.copy()
    for (i, c, s) in rots:
        zi = Z[:, i].copy()
        zi1 = Z[:, i + 1].copy()
        Z[:, i + 1] = s * zi + c * zi1
        Z[:, i] = c * zi - s * zi1
    return Z


def lu_solve_pp(S, b):
    """Gaussian elimination with partial pivoting (row swaps), then back substitution."""
    S = S.copy()
    b = b.copy()
    n = S.shape[0]
    for k in range(n):
        piv = k + np.argmax(np.abs(S[k:, k].real) + np.abs(S[k:, k].imag))
        if piv != k:
            S[[k, piv], :] = S[[piv, k], :]
            b[[k, piv]] = b[[piv, k]]
        f = S[k + 1:, k] / S[k, k]
        S[k + 1:, k + 1:] -= np.outer(f, S[k, k + 1:])
        b[k + 1:] -= f * b[k]
    x = np.zeros(n, complex)
    for k in range(n - 1, -1, -1):
        x[k] = (b[k] - S[k, k + 1:] @ x[k + 1:]) / S[k, k]
    return x


def tail(lam, W, kap, k0d, n_I, n_II, N):
    ff = 2 * N + 1
    z = 2 * N
    q = np.sqrt(lam.astype(complex))
    X = np.exp(-k0d * q)
    kI = np.conj(np.sqrt((n_I ** 2 - kap ** 2).astype(complex)))
    kII = np.conj(np.sqrt((n_II ** 2 - kap ** 2).astype(complex)))
    DI = 1j * kI / n_I ** 2
    DII = 1j * kII / n_II ** 2
    HI = (W.T * DI[None, :]) @ W
    HII = (W.T * DII[None, :]) @ W
    P = gj_inverse(np.diag(q) + HII)
    Hq = HI - np.diag(q)
    S = Hq * (1 - X ** 2)[None, :] + 2 * np.diag(q) + 2 * ((Hq * X[None, :]) @ P) * (X * q)[None, :]
    r0 = DI[z] + 1j / n_I
    b = r0 * W[z, :]
    cp = lu_solve_pp(S, b)
    dvec = 2 * P @ (q * X * cp)
    uu = (1 - X ** 2) * cp + X * dvec
    Tt = W[[0, 1, z], :] @ dvec
    Rt = W[[0, 1, z], :] @ uu
    Rt[2] -= 1

    def to_orders(v):  # v = (cos1, sin1, const) components -> orders (-1, 0, +1)
        vp = (v[0] + v[1] / 1j) / np.sqrt(2)
        vm = (v[0] - v[1] / 1j) / np.sqrt(2)
        return np.array([vm, v[2], vp])
    T3 = to_orders(Tt)
    R3 = to_orders(Rt)
    k3I = np.array([kI[0], kI[z], kI[0]])
    k3II = np.array([kII[0], kII[z], kII[0]])
    de_r = np.real(R3 * np.conj(R3) * k3I / n_I)
    de_t = np.real(T3 * np.conj(T3) * k3II / n_II ** 2) * n_I
    return de_r, de_t


def evaluate(s, wavelength=1100., angle=60., order=40, thickness=325., n_ridge=3.551726470588235,
             n_I=1.45, n_II=1.0, n_g=1.0, stages=False):
    N = order
    ff = 2 * N + 1
    u = (np.asarray(s, float) + 1) / 2
    tw, D = host_tables(u.size, N)
    c = coeffs(u, N, tw, D)
    T = basis_toeplitz(c, N)
    er, eg = n_ridge ** 2, n_g ** 2
    E = eg * np.eye(ff) + (er - eg) * T
    A = (1 / eg) * np.eye(ff) + (1 / er - 1 / eg) * T
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
    Einv = gj_inverse(E)
    Bm = sk[:, None] * Einv[np.ix_(perm, perm)] * sk[None, :] - np.eye(ff)
    L, Dg = ldlt(A)
    C = reduce_to_standard(Bm, L, Dg)
    d, e, V, beta = tridiagonalize(C)
    Q = form_q(V, beta)
    M0 = back_transform(Q, L, Dg)
    lam, rots, ok = ql_rotations(d, e)
    W = apply_rotations(M0, rots)
    k0d = 2 * np.pi / wavelength * thickness
    de_r, de_t = tail(lam, W, kap, k0d, n_I, n_II, N)
    if stages:
        return de_t[0], de_r, de_t, dict(c=c, E=E, A=A, B=Bm, C=C, d=d, e=e, M0=M0, lam=lam, W=W, nrot=len(rots), ok=ok)
    return de_t[0], de_r, de_t


if __name__ == '__main__':
    import sys
    sys.path.insert(0, '.')
    from oracle import rcwa_oracle as O
    g = np.load('tests/golden/rcwa_golden.npz')
    names = sorted({k.split('/')[0] for k in g.files})
    for nme in names:
        p = g[nme + '/pattern']
        wl = float(g[nme + '/wavelength'])
        e, der, det, st = evaluate(p, wl, float(g[nme + '/angle']), int(g[nme + '/order']), float(g[nme + '/thickness']),
                                   n_ridge=O.n_silicon_real(wl), stages=True)
        W, A = st['W'], st['A']
        print(f"{nme:30s} err={abs(e - float(g[nme + '/ref_eff'])):.2e} t3={np.abs(det[::-1] - g[nme + '/ref_tran_cut']).max():.2e} "
              f"r3={np.abs(der[::-1] - g[nme + '/ref_refl_cut']).max():.2e} nrot={st['nrot']} ({st['nrot'] / (2 * int(g[nme + '/order']) + 1) ** 2:.2f} n^2) ok={st['ok']} "
              f"orth={np.abs(W.T @ A @ W - np.eye(len(A))).max():.1e}")
    b = np.load('tests/golden/rcwa_batch.npz')
    errs = []
    nr = []
    for p, eref in zip(b['patterns'][:64], b['ref_eff'][:64]):
        e, _, _, st = evaluate(p.astype(float), stages=True)
        errs.append(abs(e - eref))
        nr.append(st['nrot'])
    print('batch64 max err', max(errs), 'rot count mean/max', np.mean(nr), max(nr))
