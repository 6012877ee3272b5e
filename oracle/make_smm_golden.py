"""16 outputs of the UNMODIFIED reference run with algo='SMM' (scattering_method.py, smm_util.py)  --  TEST INFRASTRUCTURE.

Blocky and random 256-cell patterns at order 60 and mixed wavelengths / angles / thicknesses, each through
``ref_bridge.reference_evaluate(..., algo="SMM")`` (JLABCode -> RCWALight(algo='SMM')).  Pins the S-matrix operator
 to the reference's own SMM arithmetic on more than the single G1 case.

Run in the build container (needs /root/reference):  python -m oracle.make_smm_golden
"""
from __future__ import annotations

import os

import numpy as np

from oracle import ref_bridge

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    rng = np.random.default_rng(60)
    order, n = 60, 256
    pats = np.empty((16, n))
    for i in range(16):
        block = (32, 16, 8, 4)[i % 4]
        pats[i] = np.repeat(rng.choice([-1., 1.], n // block), block) if i < 12 else rng.choice([-1., 1.], n)
    wl = np.array([1100, 1100, 1100, 1100, 900, 1000, 1200, 950, 1100, 1050, 1150, 1100, 1100, 900, 1200, 1100])
    ang = np.array([60, 60, 50, 75, 60, 65, 55, 70, 60, 60, 60, 52, 60, 75, 50, 60])
    th = np.array([325, 325, 325, 325, 325, 300, 350, 325, 250, 400, 325, 325, 325, 325, 325, 500])
    eff, rc, tc = [], [], []
    for p, w, a, t in zip(pats, wl, ang, th):
        e, r, tt = ref_bridge.reference_evaluate(p, int(w), int(a), order, int(t), "SMM")
        eff.append(e); rc.append(r); tc.append(tt)
    np.savez_compressed(os.path.join(GOLD, "rcwa_smm_golden.npz"), patterns=pats.astype(np.int8), order=order,
                        wavelength=wl.astype(np.float64), angle=ang.astype(np.float64), thickness=th.astype(np.float64),
                        ref_eff=np.array(eff), ref_refl_cut=np.array(rc), ref_tran_cut=np.array(tc))
    print("rcwa_smm_golden.npz: eff range", min(eff), max(eff))


if __name__ == "__main__":
    main()
