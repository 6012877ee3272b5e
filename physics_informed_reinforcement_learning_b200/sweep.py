"""Wavelength / deflection-angle sweeps as ONE batch axis.

The reference sweeps wavelengths with a serial Python loop (meent/meent/rcwa.py:17-34 ``sweep_wavelength``) and has no angle
sweep; here every (pattern, wavelength, angle) cell is one structure of a single batched launch with its own parameter row
{wavelength, angle, thickness, n_ridge}; n_ridge = n_Si(wavelength) is interpolated on the host
(convolution_matrix.py:40-52).
"""
from __future__ import annotations

import numpy as np

from .materials import n_silicon_real


def sweep_grid(wavelengths, angles):
    wl = np.asarray(wavelengths, dtype=np.float64).ravel()
    an = np.asarray(angles, dtype=np.float64).ravel()
    W, A = np.meshgrid(wl, an, indexing="ij")
    return wl, an, W.ravel(), A.ravel()


def sweep(backend, patterns, wavelengths, angles, thickness=325.0, grad=False):
    """patterns: int8 CUDA tensor [P, n].  Returns eff [P, L, A] (and grad [P, L, A, n]) as CUDA tensors; cell order is
    pattern-major so that each pattern's L*A cells are contiguous."""
    import torch
    wl, an, Wf, Af = sweep_grid(wavelengths, angles)
    P = patterns.shape[0]
    C = Wf.size
    rows = np.empty((P * C, 4), dtype=np.float64)
    rows[:, 0] = np.tile(Wf, P)
    rows[:, 1] = np.tile(Af, P)
    rows[:, 2] = float(thickness)
    rows[:, 3] = np.tile(n_silicon_real(Wf), P)
    params = torch.from_numpy(rows).to(patterns.device)
    cells = patterns.repeat_interleave(C, dim=0).contiguous()
    if grad:
        eff, g, st = backend.evaluate_grad(cells, params=params)
        return eff.view(P, wl.size, an.size), g.view(P, wl.size, an.size, -1), st.view(P, wl.size, an.size)
    eff = backend.evaluate(cells, params=params)
    return eff.view(P, wl.size, an.size)
