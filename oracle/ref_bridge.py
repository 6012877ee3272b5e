"""Import the UNMODIFIED reference solver from /root/reference  --  TEST INFRASTRUCTURE.

Only usable in the authoring container (``/root/reference`` does not exist on
the GPU box).  Used by ``oracle/make_golden.py`` to generate the committed
fixtures and to pin ``oracle/rcwa_oracle.py``; never imported at run time by
GPU tests, ``smoke()`` or ``bench.py``.

The env module itself (deflector_gym/envs/meent_env.py) cannot be imported:
``gym`` is absent and line 58 is a SyntaxError at the pinned commit, so the
call below repeats lines 32-44 of that file around the real ``JLABCode``.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

REF_ROOT = "/root/reference"
MEENT_DIR = os.path.join(REF_ROOT, "deflector-gym", "third_party", "meent")


def available() -> bool:
    return os.path.isdir(os.path.join(MEENT_DIR, "JLAB"))


_JLAB = None


def _load():
    global _JLAB
    if _JLAB is None:
        if not available():
            raise RuntimeError("reference tree not present (expected only in the authoring container)")
        # meent/on_numpy/_base.py:2 imports matplotlib only for an unused plot helper
        if "matplotlib" not in sys.modules:
            mpl = types.ModuleType("matplotlib")
            plt = types.ModuleType("matplotlib.pyplot")
            mpl.pyplot = plt
            sys.modules["matplotlib"] = mpl
            sys.modules["matplotlib.pyplot"] = plt
        if MEENT_DIR not in sys.path:
            sys.path.insert(0, MEENT_DIR)
        from JLAB.solver import JLABCode  # noqa
        _JLAB = JLABCode
    return _JLAB


def reference_evaluate(struct, wavelength=1100, angle=60, order=40, thickness=325, algo="TMM"):
    """(eff, refl_cut, tran_cut) exactly as MeentBase.get_efficiency obtains them."""
    JLABCode = _load()
    struct = np.asarray(struct, dtype=np.float64)[np.newaxis, np.newaxis, :]
    wls = np.array([wavelength])
    period = abs(wls / np.sin(angle / 180 * np.pi))
    calc = JLABCode(grating_type=0, n_I=1.45, n_II=1., theta=0, phi=0.,
                    fourier_order=order, period=period, wls=wls, pol=1,
                    patterns=None, ucell=struct, thickness=np.array([thickness]), algo=algo)
    eff, refl_cut, tran_cut = calc.reproduce_acs_cell('p_si__real', 1)
    return float(eff), np.asarray(refl_cut, dtype=np.float64), np.asarray(tran_cut, dtype=np.float64)


def reference_material_table():
    """WL, n, k columns of nk_data/matlab/p_Si.mat (data, 697 rows)."""
    from scipy.io import loadmat
    d = loadmat(os.path.join(MEENT_DIR, "meent", "nk_data", "matlab", "p_Si.mat"))
    return d["WL"][:, 0].copy(), d["n"][:, 0].copy(), d["k"][:, 0].copy()


def reticolo_known_answer():
    from scipy.io import loadmat
    m = loadmat(os.path.join(REF_ROOT, "deflector-gym", "third_party", "reticolo_allege",
                             "outputs_from_70_800_0.9134.mat"))
    return dict(pattern=m["pattern"][:, 0].astype(np.float64),
                wavelength=float(m["wavelength"][0, 0]), angle=float(m["angle"][0, 0]),
                thickness=float(m["thickness"][0, 0]), n_Si=float(m["n_Si"][0, 0]),
                efficiency=float(m["efficiency"][0, 0]))
