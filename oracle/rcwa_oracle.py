"""CPU oracle for the DeflectorGym RCWA hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT.

This file restates, in NumPy, the algorithm the reference runs for one
``MeentBase.get_efficiency`` call (1D grating, TM, one layer, ETM / "TMM"
recursion).  It deliberately keeps the reference's *formulation* (up-sampled
FFT -> Toeplitz convolution matrices -> two dense inverses -> general ``eig``
-> enhanced-transmittance update -> one dense solve), with the same LAPACK /
pocketfft library calls, because it doubles as the timed CPU baseline
(``bench.py`` ``cpu_baseline`` / ``--impl reference``, kind = "port").

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs
may import this module.  The product path (``physics_informed_reinforcement_
learning_b200``) never does; it fails loudly if its CUDA library is missing.

Parity status: PINNED.  ``oracle/make_golden.py`` runs the unmodified reference
solver (``/root/reference/deflector-gym/third_party/meent``; JLABCode called the
way ``deflector_gym/envs/meent_env.py:32-44`` calls it) and this restatement on
the same inputs; they agree to <= 1e-12 on every fixture in
``tests/golden/rcwa_golden.npz`` (see ``tests/test_oracle_golden.py``), which also
holds the Reticolo known answer 0.9134285 from
``deflector-gym/third_party/reticolo_allege/outputs_from_70_800_0.9134.mat``.

Reference lines followed (relative to deflector-gym/third_party/meent/):
  * material index            meent/on_numpy/convolution_matrix.py:40-52
  * pattern -> permittivity   JLAB/solver.py:21-22
  * convolution matrices      meent/on_numpy/convolution_matrix.py:89-114
  * incidence/exit media      meent/on_numpy/transfer_method.py:4-39
  * layer eigenmodes (TM)     meent/on_numpy/_base.py:150-159
  * ETM layer update          meent/on_numpy/transfer_method.py:42-57
  * final solve, efficiencies meent/on_numpy/transfer_method.py:60-75
  * order cut / return value  JLAB/solver.py:29-38
  * env constants             deflector_gym/envs/meent_env.py:26-46
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_TABLE = os.path.join(_HERE, "..", "physics_informed_reinforcement_learning_b200", "data", "p_si_nk.npz")

# constants the env hard-codes (meent_env.py:36-44)
N_INCIDENT = 1.45   # n_I  (glass substrate side, light comes from here)
N_EXIT = 1.0        # n_II (air)
N_GROOVE = 1.0


@dataclass
class OracleResult:
    eff: float            # de_ti[N-1]  == tran_cut[-1]   (solver.py:38)
    refl_cut: np.ndarray  # de_ri[N-1:N+2][::-1]
    tran_cut: np.ndarray  # de_ti[N-1:N+2][::-1]
    de_ri: np.ndarray     # (ff,)
    de_ti: np.ndarray     # (ff,)
    coeff_eps: np.ndarray = None      # 4N+1 central Fourier coefficients of eps(x)
    coeff_inv_eps: np.ndarray = None  # ... of 1/eps(x)
    eigvals_sorted: np.ndarray = None  # eigenvalues q^2 sorted by real part


_table_cache = None


def silicon_table():
    """(wavelength_nm, n, k) rows of the p_Si table (nk_data/matlab/p_Si.mat, 697 rows)."""
    global _table_cache
    if _table_cache is None:
        z = np.load(_TABLE)
        _table_cache = np.stack([z["wl"], z["n"], z["k"]], axis=1)
    return _table_cache


def n_silicon_real(wavelength_nm: float) -> float:
    """'p_si__real': linear interpolation of n(lambda), k dropped (convolution_matrix.py:40-52)."""
    t = silicon_table()
    return float(np.interp(wavelength_nm, t[:, 0], t[:, 1]))


def toeplitz_from_cells(values: np.ndarray, order: int, return_coeffs: bool = False):
    """Convolution (Toeplitz) matrix of a piecewise-constant function given per cell.

    Follows convolution_matrix.py:89-114: every cell is repeated R = 4N+2 times
    (n = (4N+1)*n_cells // n_cells, repeat n+1), the M-point DFT divided by M is
    centred, the 4N+1 central bins are kept, and E[i, j] = c[i-j].
    """
    values = np.asarray(values, dtype=np.float64).ravel()
    ncell = values.size
    ff = 2 * order + 1
    need = (4 * order + 1) * ncell
    if ncell < need:
        reps = need // ncell + 1
        values = np.repeat(values, reps)
    spec = np.fft.fftshift(np.fft.fft(values / values.size))
    mid = spec.size // 2
    c = spec[mid - 2 * order: mid + 2 * order + 1]          # c[2N + k] = c_k, k=-2N..2N
    idx = np.arange(ff)
    mat = c[(idx[:, None] - idx[None, :]) + 2 * order]
    if return_coeffs:
        return mat, c
    return mat


def solve_tm_single_layer(e_conv, oe_conv, wavelength, period, thickness, order,
                          n_inc=N_INCIDENT, n_exit=N_EXIT, theta=0.0, return_eigs=False):
    """ETM solve of one grating layer, TM polarisation (_base.py:119-180 with pol=1, algo='TMM')."""
    ff = 2 * order + 1
    m = np.arange(-order, order + 1)
    delta = np.zeros(ff)
    delta[order] = 1.0
    k0 = 2.0 * np.pi / wavelength

    # transfer_method.py:4-39
    kx = k0 * (n_inc * np.sin(theta) - m * (wavelength / period)).astype(complex)
    kz_inc = np.conj((k0 ** 2 * n_inc ** 2 - kx ** 2) ** 0.5)
    kz_exit = np.conj((k0 ** 2 * n_exit ** 2 - kx ** 2) ** 0.5)
    Kx = np.diag(kx / k0)
    Z_inc = np.diag(kz_inc / (k0 * n_inc ** 2))
    Z_exit = np.diag(kz_exit / (k0 * n_exit ** 2))
    f = np.eye(ff)
    g = 1j * Z_exit
    inc = 1j * delta * np.cos(theta) / n_inc
    T = np.eye(ff)

    # _base.py:150-159
    E_inv = np.linalg.inv(e_conv)
    B = Kx @ E_inv @ Kx - np.eye(ff)
    A_inv = np.linalg.inv(oe_conv)
    lam, W = np.linalg.eig(A_inv @ B)
    q = lam ** 0.5
    V = oe_conv @ W @ np.diag(q)

    # transfer_method.py:42-57
    X = np.diag(np.exp(-k0 * q * thickness))
    W_i = np.linalg.inv(W)
    V_i = np.linalg.inv(V)
    a = 0.5 * (W_i @ f + V_i @ g)
    b = 0.5 * (W_i @ f - V_i @ g)
    a_i = np.linalg.inv(a)
    f = W @ (np.eye(ff) + X @ b @ a_i @ X)
    g = V @ (np.eye(ff) - X @ b @ a_i @ X)
    T = T @ a_i @ X

    # transfer_method.py:60-75
    Tl = np.linalg.inv(g + 1j * Z_inc @ f) @ (1j * Z_inc @ delta + inc)
    R = f @ Tl - delta
    T = T @ Tl
    de_ri = np.real(R * np.conj(R) * kz_inc / (k0 * n_inc * np.cos(theta)))
    de_ti = np.real(T * np.conj(T) * kz_exit / n_exit ** 2) / (k0 * np.cos(theta) / n_inc)
    if return_eigs:
        return de_ri.real, de_ti.real, lam
    return de_ri.real, de_ti.real


def evaluate(struct, wavelength=1100.0, angle=60.0, order=40, thickness=325.0,
             n_ridge=None, intermediates=False) -> OracleResult:
    """One ``get_efficiency`` call (meent_env.py:26-46 with mfs=1 -> penalty 0).

    ``struct``: (n_cells,) of +-1 (any real values are accepted: the map to
    permittivity is affine, solver.py:21-22, which is what the finite-difference
    gradient oracle relies on).
    """
    s = np.asarray(struct, dtype=np.float64).ravel()
    period = abs(wavelength / np.sin(angle / 180.0 * np.pi))
    if n_ridge is None:
        n_ridge = n_silicon_real(wavelength)
    u = (s + 1.0) / 2.0
    eps = u * (n_ridge ** 2 - N_GROOVE ** 2) + N_GROOVE ** 2
    e_conv, c_e = toeplitz_from_cells(eps, order, True)
    oe_conv, c_oe = toeplitz_from_cells(1.0 / eps, order, True)
    out = solve_tm_single_layer(e_conv, oe_conv, wavelength, period, thickness, order,
                                return_eigs=intermediates)
    de_ri, de_ti = out[0], out[1]
    c = order
    res = OracleResult(
        eff=float(de_ti[c - 1]),
        refl_cut=de_ri[c - 1:c + 2][::-1].copy(),
        tran_cut=de_ti[c - 1:c + 2][::-1].copy(),
        de_ri=de_ri, de_ti=de_ti)
    if intermediates:
        lam = out[2]
        res.coeff_eps = c_e
        res.coeff_inv_eps = c_oe
        res.eigvals_sorted = lam[np.argsort(lam.real, kind="stable")]
    return res


def efficiency(struct, wavelength=1100.0, angle=60.0, order=40, thickness=325.0, n_ridge=None) -> float:
    return evaluate(struct, wavelength, angle, order, thickness, n_ridge).eff


def efficiency_batch(structs, wavelength=1100.0, angle=60.0, order=40, thickness=325.0) -> np.ndarray:
    structs = np.asarray(structs)
    return np.array([efficiency(s, wavelength, angle, order, thickness) for s in structs])


def fd_gradient(struct, idx, wavelength=1100.0, angle=60.0, order=40, thickness=325.0, h=1e-4):
    """Central finite difference d eff / d s_i   -- the gradient oracle."""
    s = np.asarray(struct, dtype=np.float64).copy()
    out = np.zeros(len(idx))
    for j, i in enumerate(idx):
        sp = s.copy(); sp[i] += h
        sm = s.copy(); sm[i] -= h
        out[j] = (efficiency(sp, wavelength, angle, order, thickness)
                  - efficiency(sm, wavelength, angle, order, thickness)) / (2 * h)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# General 1D stack: both polarisations, both algorithms, any number of layers (normal or oblique incidence in the plane).
# Restates _BaseRCWA.solve_1d (_base.py:119-180), transfer_1d_1/2/3 (transfer_method.py:4-75) and scattering_1d_1/2/3 with
# their helpers (scattering_method.py:12-73; smm_util.py:15-85, 136-182, 289-318).  Pinned by the multilayer / TE fixtures in
# tests/golden/rcwa_stack_golden.npz, written by oracle/make_stack_golden.py from the unmodified reference.
# NB the reference walks the layers in REVERSED list order for BOTH algorithms (_base.py:141).  For 'TMM' that builds the
# stack from the exit side up (list order = incidence side first); for 'SMM' the star products then append the layers to the
# reflection-side matrix in reversed list order, i.e. the SMM result of a list equals the TMM result of the reversed list.
# The restatement keeps that behaviour.
# ---------------------------------------------------------------------------------------------------------------------
def _homogeneous_1d(kx_norm, n_index, pol):
    """smm_util.py:136-182 on the diagonal: returns (V, Kz) as vectors (W = I)."""
    e_r = n_index ** 2
    pert = 1e-20 * (1 + 1j)
    Q = (e_r - kx_norm ** 2).astype(complex)
    Q[Q == 0] = np.conj(pert)
    Kz = np.conj(np.sqrt((e_r - kx_norm ** 2).astype(complex)))
    Kz[Kz == 0] = pert
    if pol:
        Kz = Kz * n_index ** 2
    V = Q / (-1j * Kz)
    return V, Kz


def _s_rt(A, B, ref_mode):
    """smm_util.py:70-85."""
    A_i = np.linalg.inv(A)
    S11, S12, S21, S22 = -A_i @ B, 2 * A_i, 0.5 * (A - B @ A_i @ B), B @ A_i
    if ref_mode:
        return dict(S11=S11, S22=S22, S12=S12, S21=S21)
    return dict(S11=S22, S22=S11, S12=S21, S21=S12)


def _redheffer(SA, SB):
    """smm_util.py:289-318."""
    n = len(SA['S11'])
    I = np.eye(n)
    D_i = np.linalg.inv(I - SB['S11'] @ SA['S22'])
    F_i = np.linalg.inv(I - SA['S22'] @ SB['S11'])
    return dict(S11=SA['S11'] + SA['S12'] @ D_i @ SB['S11'] @ SA['S21'], S12=SA['S12'] @ D_i @ SB['S12'],
                S21=SB['S21'] @ F_i @ SA['S21'], S22=SB['S22'] + SB['S21'] @ F_i @ SA['S22'] @ SB['S12'])


def solve_stack(e_convs, oe_convs, thicknesses, wavelength, period, order, pol=1, algo='TMM',
                n_inc=N_INCIDENT, n_exit=N_EXIT, theta=0.0):
    """de_ri, de_ti of a stack of grating layers (lists in the reference's list order)."""
    ff = 2 * order + 1
    m = np.arange(-order, order + 1)
    delta = np.zeros(ff)
    delta[order] = 1.0
    k0 = 2.0 * np.pi / wavelength
    I = np.eye(ff)
    if algo == 'TMM':                                       # transfer_1d_1
        kx = k0 * (n_inc * np.sin(theta) - m * (wavelength / period)).astype(complex)
        kz_inc = np.conj((k0 ** 2 * n_inc ** 2 - kx ** 2) ** 0.5)
        kz_exit = np.conj((k0 ** 2 * n_exit ** 2 - kx ** 2) ** 0.5)
        Kx = np.diag(kx / k0)
        f = np.eye(ff)
        if pol == 0:
            YZ_I = np.diag(kz_inc / k0); g = 1j * np.diag(kz_exit / k0); inc = 1j * n_inc * np.cos(theta) * delta
        else:
            YZ_I = np.diag(kz_inc / (k0 * n_inc ** 2)); g = 1j * np.diag(kz_exit / (k0 * n_exit ** 2))
            inc = 1j * delta * np.cos(theta) / n_inc
        T = np.eye(ff)
    elif algo == 'SMM':                                     # scattering_1d_1
        kxn = (n_inc * np.sin(theta) - m * (2 * np.pi / k0 / period)).astype(complex)
        Kx = np.diag(kxn)
        Vg, _ = _homogeneous_1d(kxn, 1.0, None)
        Vr, Kzr = _homogeneous_1d(kxn, n_inc, pol)
        Vt, Kzt = _homogeneous_1d(kxn, n_exit, pol)
        Sg = _s_rt(I + np.diag(Vr / Vg), I - np.diag(Vr / Vg), True)
    else:
        raise ValueError(algo)
    for e_conv, oe_conv, d in zip(e_convs[::-1], oe_convs[::-1], list(thicknesses)[::-1]):
        if pol == 0:                                        # _base.py:142-148
            lam, W = np.linalg.eig(Kx ** 2 - e_conv)
            q = lam ** 0.5
            V = W @ np.diag(q)
        elif pol == 1:                                      # _base.py:150-159
            E_i = np.linalg.inv(e_conv)
            B = Kx @ E_i @ Kx - I
            lam, W = np.linalg.eig(np.linalg.inv(oe_conv) @ B)
            q = lam ** 0.5
            V = oe_conv @ W @ np.diag(q)
        else:
            raise ValueError(pol)
        if algo == 'TMM':                                   # transfer_1d_2
            X = np.diag(np.exp(-k0 * q * d))
            W_i, V_i = np.linalg.inv(W), np.linalg.inv(V)
            a, b = 0.5 * (W_i @ f + V_i @ g), 0.5 * (W_i @ f - V_i @ g)
            a_i = np.linalg.inv(a)
            f, g, T = W @ (I + X @ b @ a_i @ X), V @ (I - X @ b @ a_i @ X), T @ a_i @ X
        else:                                               # scattering_1d_2: A_B_matrices, S_layer, RedhefferStar
            W_i, V_i = np.linalg.inv(W), np.linalg.inv(V)
            A = W_i + V_i @ np.diag(Vg)
            B = W_i - V_i @ np.diag(Vg)
            X = np.diag(np.exp(-q * d * k0))
            A_i = np.linalg.inv(A)
            term_i = np.linalg.inv(A - X @ B @ A_i @ X @ B)
            S11 = term_i @ (X @ B @ A_i @ X @ A - B)
            S12 = term_i @ X @ (A - B @ A_i @ B)
            Sg = _redheffer(Sg, dict(S11=S11, S22=S11, S12=S12, S21=S12))
    if algo == 'TMM':                                       # transfer_1d_3
        Tl = np.linalg.inv(g + 1j * YZ_I @ f) @ (1j * YZ_I @ delta + inc)
        R = f @ Tl - delta
        T = T @ Tl
        de_ri = np.real(R * np.conj(R) * kz_inc / (k0 * n_inc * np.cos(theta)))
        if pol == 0:
            de_ti = np.real(T * np.conj(T) * kz_exit / (k0 * n_inc * np.cos(theta)))
        else:
            de_ti = np.real(T * np.conj(T) * kz_exit / n_exit ** 2) / (k0 * np.cos(theta) / n_inc)
        return de_ri, de_ti
    St = _s_rt(I + np.diag(Vt / Vg), I - np.diag(Vt / Vg), False)      # scattering_1d_3
    Sg = _redheffer(Sg, St)
    kinc_z = n_inc * np.cos(theta)
    r = Sg['S11'] @ delta
    t = Sg['S21'] @ delta
    rsq, tsq = np.abs(r) ** 2, np.abs(t) ** 2
    if pol == 0:
        return np.real(Kzr) * rsq / np.real(kinc_z), np.real(Kzt) * tsq / np.real(kinc_z)
    return np.real(Kzr) * rsq / np.real(kinc_z) / n_inc ** 2, np.real(Kzt) * tsq / np.real(kinc_z) * n_inc ** 2 / n_exit ** 4


def evaluate_stack(structs, wavelength=1100.0, angle=60.0, order=40, thicknesses=(325.0,), pol=1, algo='TMM',
                   n_ridge=None) -> OracleResult:
    """structs: (L, n_cells) of +-1, one row per layer in the reference's list order (JLABCode with ucell of shape (L, 1, n),
    JLAB/solver.py:16-38)."""
    s = np.atleast_2d(np.asarray(structs, dtype=np.float64))
    period = abs(wavelength / np.sin(angle / 180.0 * np.pi))
    if n_ridge is None:
        n_ridge = n_silicon_real(wavelength)
    eps = (s + 1.0) / 2.0 * (n_ridge ** 2 - N_GROOVE ** 2) + N_GROOVE ** 2
    e_convs = [toeplitz_from_cells(e, order) for e in eps]
    oe_convs = [toeplitz_from_cells(1.0 / e, order) for e in eps]
    de_ri, de_ti = solve_stack(e_convs, oe_convs, thicknesses, wavelength, period, order, pol, algo)
    c = order
    return OracleResult(eff=float(de_ti[c - 1]), refl_cut=de_ri[c - 1:c + 2][::-1].copy(), tran_cut=de_ti[c - 1:c + 2][::-1].copy(),
                        de_ri=de_ri, de_ti=de_ti)
